"""Host side of the join: docker text -> packed fleet tables.

The reference re-reads docker state on every poll: ``docker ps`` / ``docker stats`` CSV
(gpu.py:92-103), then per container ``docker top`` (gpu.py:177), ``docker inspect``
(gpu.py:86-88,184-199) and ``docker exec .. nvidia-smi -L`` (gpu.py:221-233).  None of
that is per-sample work, so it stays on the host: this module parses those texts with the
same stdlib calls and rules as the reference and packs the outcome -- container order, pid
lists, the gate, the three pod strings, and *where the reference would have raised* --
into the arrays ``kgm_load_fleet`` (include/kgm.h) keeps resident in HBM.
"""
from __future__ import annotations

import csv
import io
import json
from dataclasses import dataclass, field

import numpy as np

from . import _ffi


@dataclass
class ContainerEntry:
    cid: str
    gated: bool = False
    err_inspect: int = 0
    err_emit: int = 0
    namespace: str = "-"
    podname: str = "-"
    user: str = "-"
    pids: list = field(default_factory=list)
    visible: list = field(default_factory=list)


@dataclass
class NodeEntry:
    """One snapshot of one node's docker state, as the join needs it."""
    name: str
    err_parse: int = 0
    err_join: int = 0
    containers: list = field(default_factory=list)


def _keyed_rows(text, columns, key):
    """docker ps / stats text -> {id: row}; ids lose the leading quote that the
    un-shell-parsed ``--format`` argument leaves in place (gpu.py:92-103)."""
    out = {}
    for cells in csv.reader(io.StringIO(text)):
        row = dict(zip(columns, cells))
        if key in row:
            row[key] = row[key][1:]
        out[row.pop(key)] = row
    return out


def _user_name(podname):
    # gpu.py:196-199
    if "turing-gpu-notebook-" in podname:
        return podname.split("turing-gpu-notebook-")[1].rsplit("-", 2)[0]
    return podname.split("-")[1]


def _node_frame(node_name, ps_text, stats_text):
    """The node-level part of node_entry: (entry, container ids in `docker stats` order); the list is
    None if the reference dies before it looks at any container."""
    node = NodeEntry(name="%s" % (node_name,))          # '%s' % None == "None"  (gpu.py:131,137)
    try:
        ps = _keyed_rows(ps_text, ["ID", "Image", "Ports"], "ID")
        stats = _keyed_rows(stats_text, ["Container", "MemUsage", "CPUPerc"], "Container")
    except Exception as exc:                            # raised before any line (gpu.py:92-103)
        node.err_parse = _ffi.doc_code_of_exception(exc)
        return node, None
    for cid in stats:                                   # gpu.py:172
        if cid not in ps:
            node.err_join = _ffi.DOC_KEY_ERROR
            return node, None
    return node, list(stats)                            # gpu.py:174: docker stats order


def _container_entry(cid, top, inspect, smi_list) -> ContainerEntry:
    """One container, with json.loads and str.split themselves (every raise point classified)."""
    ent = ContainerEntry(cid=cid)
    try:
        ent.pids = top[cid].split("\n")[1:-1]           # gpu.py:177
        squeezed = inspect[cid].replace("\n", "").replace(" ", "")   # gpu.py:87
        labels = json.loads(squeezed)[0]["Config"]["Labels"]         # gpu.py:88,184
        gate = "False"
        if "mlgpu" in labels.keys():                    # gpu.py:186
            gate = labels["mlgpu"]
        ent.gated = gate == "mlgpu"                     # gpu.py:188
    except Exception as exc:
        ent.err_inspect = _ffi.doc_code_of_exception(exc)
        ent.pids = []
        return ent
    if not ent.gated:
        return ent
    ent.visible = str(smi_list.get(cid, "")).split()    # gpu.py:224
    try:                                                # gpu.py:193-199 == 227-233
        if "io.kubernetes.pod.name" in labels.keys():
            podname = labels["io.kubernetes.pod.name"]
            namespace = labels["io.kubernetes.pod.namespace"]
            user = _user_name(podname)
            ent.podname, ent.namespace, ent.user = "%s" % podname, "%s" % namespace, "%s" % user
    except Exception as exc:
        ent.err_emit = _ffi.doc_code_of_exception(exc)
    return ent


def node_entry(node_name, ps_text, stats_text, top, inspect, smi_list) -> NodeEntry:
    """Parses one node's docker texts.  ``top`` / ``inspect`` / ``smi_list`` map a
    container id to the text the reference's subprocess would have returned."""
    node, cids = _node_frame(node_name, ps_text, stats_text)
    if cids is not None:
        node.containers = [_container_entry(cid, top, inspect, smi_list) for cid in cids]
    return node


def node_entries_device(nodes, device=0):
    """``node_entry`` for many nodes at once, the `docker inspect` texts scanned on the GPU
    (kgm_inspect_labels: one thread per container; csrc/kgm_docker.cuh).  ``nodes`` = iterable of
    (node_name, ps_text, stats_text, top, inspect, smi_list).  The scanner takes the texts it is exact
    on; every other container goes through ``_container_entry`` as before.  Returns (entries, number
    of containers the device took, number left to the host)."""
    import ctypes as C
    lib = _ffi.load()
    frames, jobs, blobs = [], [], []
    for node_name, ps_text, stats_text, top, inspect, smi_list in nodes:
        node, cids = _node_frame(node_name, ps_text, stats_text)
        frames.append((node, cids, top, inspect, smi_list))
        for k, cid in enumerate(cids or []):
            text = inspect.get(cid) if isinstance(inspect, dict) else None
            if isinstance(text, str) and isinstance(top.get(cid) if isinstance(top, dict) else None, str):
                try:
                    raw = text.encode("ascii")           # (anything else is the host's: the scanner takes ASCII only)
                except UnicodeEncodeError:
                    continue
                jobs.append((len(frames) - 1, k))
                blobs.append(raw)
    n = len(jobs)
    fields = None
    if n:
        off = np.zeros(n + 1, np.uint64)
        np.cumsum([len(b) for b in blobs], out=off[1:])
        texts = np.frombuffer(b"".join(blobs) + b"\0", np.uint8)
        u8 = [np.zeros(n, np.uint8) for _ in range(3)]
        u32 = [np.zeros(n, np.uint32) for _ in range(6)]
        out = _ffi.KgmInspectOut(*[a.ctypes.data for a in u8 + u32])
        rc = lib.kgm_inspect_labels(int(device), texts.ctypes.data, off.ctypes.data, n, C.byref(out))
        if rc != _ffi.KGM_OK:
            raise RuntimeError("kgm_inspect_labels failed: %d" % rc)
        fields = (u8, u32)
    taken = {}
    for j, (fi, k) in enumerate(jobs):
        u8, u32 = fields
        if u8[0][j]:
            continue
        raw = blobs[j]
        cut = lambda o, ln: raw[int(o):int(o) + int(ln)].decode("ascii")
        taken[(fi, k)] = (bool(u8[1][j]), bool(u8[2][j]), cut(u32[0][j], u32[1][j]), cut(u32[2][j], u32[3][j]),
                          cut(u32[4][j], u32[5][j]))
    entries = []
    for fi, (node, cids, top, inspect, smi_list) in enumerate(frames):
        for k, cid in enumerate(cids or []):
            got = taken.get((fi, k))
            if got is None:
                node.containers.append(_container_entry(cid, top, inspect, smi_list))
                continue
            gated, has_pod, pod, ns, user = got
            ent = ContainerEntry(cid=cid, gated=gated)
            ent.pids = top[cid].split("\n")[1:-1]       # gpu.py:177
            if gated:
                ent.visible = str(smi_list.get(cid, "")).split()   # gpu.py:224
                if has_pod:
                    ent.podname, ent.namespace, ent.user = pod, ns, user
            node.containers.append(ent)
        entries.append(node)
    return entries, len(taken), sum(len(c or []) for _, c, _, _, _ in frames) - len(taken)


class _Strings:
    """Back-to-back byte strings + uint32 offsets (kgm_strings)."""

    def __init__(self, items):
        enc = [s.encode() if isinstance(s, str) else bytes(s) for s in items]
        self.off = np.zeros(len(enc) + 1, np.uint32)
        if enc:
            np.cumsum([len(b) for b in enc], out=self.off[1:])
        self.bytes = np.frombuffer(b"".join(enc) + b"\0", np.uint8)

    def struct(self):
        return _ffi.KgmStrings(self.bytes.ctypes.data, self.off.ctypes.data)


class PackedFleet:
    """The arrays behind a ``kgm_fleet`` (kept alive for the duration of the call)."""

    def __init__(self, nodes):
        self.nodes = list(nodes)
        conts = [c for n in self.nodes for c in n.containers]
        self.n_nodes, self.n_containers = len(self.nodes), len(conts)
        self.node_name = _Strings([n.name for n in self.nodes])
        self.cont_begin = np.zeros(self.n_nodes + 1, np.uint32)
        if self.nodes:
            np.cumsum([len(n.containers) for n in self.nodes], out=self.cont_begin[1:])
        self.err_parse = np.asarray([n.err_parse for n in self.nodes] + [0], np.uint8)
        self.err_join = np.asarray([n.err_join for n in self.nodes] + [0], np.uint8)
        self.gated = np.asarray([c.gated for c in conts] + [0], np.uint8)
        self.err_inspect = np.asarray([c.err_inspect for c in conts] + [0], np.uint8)
        self.err_emit = np.asarray([c.err_emit for c in conts] + [0], np.uint8)
        self.cid = _Strings([c.cid for c in conts])
        self.ns = _Strings([c.namespace for c in conts])
        self.pod = _Strings([c.podname for c in conts])
        self.user = _Strings([c.user for c in conts])
        self.pid_begin = np.zeros(self.n_containers + 1, np.uint32)
        self.vis_begin = np.zeros(self.n_containers + 1, np.uint32)
        if conts:
            np.cumsum([len(c.pids) for c in conts], out=self.pid_begin[1:])
            np.cumsum([len(c.visible) for c in conts], out=self.vis_begin[1:])
        self.pid = _Strings([p for c in conts for p in c.pids])
        self.vis = _Strings([v for c in conts for v in c.visible])

    def struct(self):
        p = lambda a: a.ctypes.data
        return _ffi.KgmFleet(
            self.n_nodes, self.node_name.struct(), p(self.cont_begin), p(self.err_parse),
            p(self.err_join), self.n_containers, p(self.gated), p(self.err_inspect),
            p(self.err_emit), self.cid.struct(), self.ns.struct(), self.pod.struct(),
            self.user.struct(), p(self.pid_begin), self.pid.struct(), p(self.vis_begin),
            self.vis.struct())
