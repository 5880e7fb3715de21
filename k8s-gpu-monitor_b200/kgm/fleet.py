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


def node_entry(node_name, ps_text, stats_text, top, inspect, smi_list) -> NodeEntry:
    """Parses one node's docker texts.  ``top`` / ``inspect`` / ``smi_list`` map a
    container id to the text the reference's subprocess would have returned."""
    node = NodeEntry(name="%s" % (node_name,))          # '%s' % None == "None"  (gpu.py:131,137)
    try:
        ps = _keyed_rows(ps_text, ["ID", "Image", "Ports"], "ID")
        stats = _keyed_rows(stats_text, ["Container", "MemUsage", "CPUPerc"], "Container")
    except Exception as exc:                            # raised before any line (gpu.py:92-103)
        node.err_parse = _ffi.doc_code_of_exception(exc)
        return node
    for cid in stats:                                   # gpu.py:172
        if cid not in ps:
            node.err_join = _ffi.DOC_KEY_ERROR
            return node
    for cid in stats:                                   # gpu.py:174: docker stats order
        ent = ContainerEntry(cid=cid)
        node.containers.append(ent)
        try:
            ent.pids = top[cid].split("\n")[1:-1]       # gpu.py:177
            squeezed = inspect[cid].replace("\n", "").replace(" ", "")   # gpu.py:87
            labels = json.loads(squeezed)[0]["Config"]["Labels"]         # gpu.py:88,184
            gate = "False"
            if "mlgpu" in labels.keys():                # gpu.py:186
                gate = labels["mlgpu"]
            ent.gated = gate == "mlgpu"                 # gpu.py:188
        except Exception as exc:
            ent.err_inspect = _ffi.doc_code_of_exception(exc)
            ent.pids = []
            continue
        if not ent.gated:
            continue
        ent.visible = str(smi_list.get(cid, "")).split()   # gpu.py:224
        try:                                            # gpu.py:193-199 == 227-233
            if "io.kubernetes.pod.name" in labels.keys():
                podname = labels["io.kubernetes.pod.name"]
                namespace = labels["io.kubernetes.pod.namespace"]
                user = _user_name(podname)
                ent.podname, ent.namespace, ent.user = "%s" % podname, "%s" % namespace, "%s" % user
        except Exception as exc:
            ent.err_emit = _ffi.doc_code_of_exception(exc)
    return node


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
