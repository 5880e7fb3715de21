"""Synthetic replay of the texts one poll of the reference consumes.

The reference reads four subprocess outputs per poll (``gpu.py:92-116``) plus three
per-container ones (``gpu.py:177,184,221``).  There is no docker / nvidia-smi on the build
or GPU box, so fleet-scale input is *replayed*: this module produces those texts for a
synthetic fleet, deterministically.  Every value is a pure function
``f(seed, node, gpu-or-proc, tick, field)`` of a counter-based mixer (splitmix64), so a
document can be regenerated anywhere from its coordinates.

Text shapes follow real tool output: they are generator
assumptions, not parity facts -- parity is always "same bytes as the oracle on the same
texts", whatever the texts are.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field

import numpy as np

GPU_HEADER = ("uuid, utilization.gpu [%], utilization.memory [%], memory.total [MiB], "
              "memory.used [MiB], memory.free [MiB]\n")
APPS_HEADER = "pid, process_name, gpu_uuid, used_gpu_memory [MiB]\n"
MEM_TOTAL_MIB = 183359          # what nvidia-smi prints for a B200

_PROC_NAMES = (
    "/opt/conda/envs/py3/bin/python",
    "/opt/conda/bin/python3.10",
    "python3 train.py",
    "/usr/bin/python3",
    "/usr/local/bin/tritonserver",
    "/opt/conda/envs/tf2/bin/python3.9",
    "python",
    "/workspace/bin/worker --rank 0",
    "/usr/bin/ffmpeg",
    "/opt/app/venv/bin/python -m serve",
)
_NAMESPACES = tuple(["default", "kube-system", "ml-platform", "research"] +
                    ["10007%03d-team" % i for i in range(46)])
_APPS = ("jupyter", "trainer", "serving", "batch", "notebook")
_M64 = (1 << 64) - 1


def mix64(x):
    """splitmix64 finaliser; works on python ints and on uint64 numpy arrays."""
    if isinstance(x, np.ndarray):
        x = x.astype(np.uint64)
        with np.errstate(over="ignore"):
            x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return x ^ (x >> np.uint64(31))
    x &= _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def h(seed, *coords):
    """Counter-based hash of a coordinate tuple (python ints)."""
    x = mix64(seed + 0x9E3779B97F4A7C15)
    for c in coords:
        x = mix64(x ^ ((c + 0x9E3779B97F4A7C15) & _M64))
    return x


def _hex(x, n):
    return ("%016x" % x)[:n]


def _uuid(seed, node, gpu):
    a, b = h(seed, 1, node, gpu), h(seed, 2, node, gpu)
    s = "%016x%016x" % (a, b)
    return "GPU-%s-%s-%s-%s-%s" % (s[0:8], s[8:12], s[12:16], s[16:20], s[20:32])


_ALNUM = "abcdefghijklmnopqrstuvwxyz0123456789"


def _alnum(x, n):
    out = []
    for _ in range(n):
        out.append(_ALNUM[x % 36])
        x //= 36
    return "".join(out)


@dataclass
class ContainerSnap:
    cid: str
    gated: bool
    labels: dict                    # the Labels object docker inspect would show (or None)
    pids: list                      # pid strings as `docker top -eo pid` prints them
    visible: list                   # uuids `nvidia-smi -L` shows inside the container
    gpu_procs: list = field(default_factory=list)   # [(pid:int, gpu:int, name_idx:int)]


@dataclass
class NodeSnap:
    """Docker-side state of one node, as the texts the reference would read."""
    name: str
    uuids: list
    containers: list                # ContainerSnap, in `docker stats` order
    mig: int = 0                    # 0 = plain GPUs; 1 = MIG parents: utilization prints "[N/A]" and idle
                                    # containers see MIG instance uuids; 2 = additionally the process
                                    # rows name MIG instance uuids (the reference dies: gpu.py:151)

    @property
    def ps_text(self):              # gpu.py:92-97; argv is not shell-parsed -> literal quotes
        return "".join("'%s,registry.local/ml/%s:latest,'\n" % (c.cid, "img%d" % (i % 7))
                       for i, c in enumerate(self.containers))

    @property
    def stats_text(self):           # gpu.py:98-103
        return "".join("'%s,%d.%dGiB / 503.5GiB,%d.%02d%%'\n" %
                       (c.cid, 1 + i % 50, i % 10, i % 400, i % 100)
                       for i, c in enumerate(self.containers))

    @property
    def top(self):                  # gpu.py:177
        return {c.cid: "PID\n" + "".join(p + "\n" for p in c.pids) for c in self.containers}

    @property
    def inspect(self):              # gpu.py:86-88
        return {c.cid: json.dumps([{"Id": c.cid * 5 + c.cid[:4], "Path": "/bin/sh",
                                    "Config": {"Hostname": c.cid, "Labels": c.labels}}],
                                   indent=4)
                for c in self.containers}

    @property
    def smi_list(self):             # gpu.py:221-224
        return {c.cid: "".join(u + "\n" for u in c.visible) for c in self.containers}


def make_node(seed, node, *, gpus=8, containers=8, procs_per_container=4,
              idle_every=0, ungated_every=0, name_fmt="k8s-node-gpu-%05d"):
    """One node snapshot.  Container ``i`` is pinned to GPU ``i % gpus``; it runs
    ``procs_per_container`` GPU processes plus a couple of non-GPU pids unless it is an
    idle one (every ``idle_every``-th) or an ungated system container."""
    uuids = [_uuid(seed, node, g) for g in range(gpus)]
    used_pids = set()
    snaps = []
    for i in range(containers):
        x = h(seed, 3, node, i)
        cid = _hex(x, 12)
        gated = not (ungated_every and i % ungated_every == ungated_every - 1)
        idle = bool(idle_every and i % idle_every == idle_every - 1)
        user = "user%05d" % (h(seed, 4, node, i) % 20000)
        if (x >> 20) % 10 == 0:
            pod = "%s-%s-%d" % (_APPS[(x >> 8) % len(_APPS)], user, (x >> 12) % 8)
        else:
            pod = "turing-gpu-notebook-%s-%s-%s" % (user, _hex(h(seed, 5, node, i), 6),
                                                    _alnum(h(seed, 6, node, i), 5))
        labels = {"annotation.io.kubernetes.container.restartCount": "0",
                  "io.kubernetes.container.name": "main",
                  "io.kubernetes.pod.name": pod,
                  "io.kubernetes.pod.namespace": _NAMESPACES[(x >> 32) % len(_NAMESPACES)],
                  "io.kubernetes.pod.uid": _hex(h(seed, 7, node, i), 16)}
        if gated:
            labels["mlgpu"] = "mlgpu"
        gpu = i % gpus
        pids, gpu_procs = [], []
        nproc = 0 if (idle or not gated) else procs_per_container
        for j in range(nproc + 2):
            k = 0
            while True:
                pid = 1000 + h(seed, 8, node, i, j, k) % 4193304
                if pid not in used_pids:
                    break
                k += 1
            used_pids.add(pid)
            pids.append(str(pid))
            if j < nproc:
                gpu_procs.append((pid, gpu, h(seed, 9, node, i, j) % len(_PROC_NAMES)))
        snaps.append(ContainerSnap(cid=cid, gated=gated, labels=labels, pids=pids,
                                   visible=[uuids[gpu]] if gated else [],
                                   gpu_procs=gpu_procs))
    return NodeSnap(name=name_fmt % node, uuids=uuids, containers=snaps)


def make_texts(seed, node_idx, snap: NodeSnap, tick, power=False):
    """The two nvidia-smi texts of one poll of ``snap`` at ``tick`` (gpu.py:104-116).
    ``power`` appends a 7th ``power.draw`` column (an extended query; the reference's
    ``zip`` ignores extra cells, gpu.py:49, so its lines do not change)."""
    ng = len(snap.uuids)
    g = np.arange(ng, dtype=np.uint64)
    base = np.uint64(h(seed, 10, node_idx, tick))
    util = (mix64(base ^ (g * np.uint64(4) + np.uint64(1))) % np.uint64(101)).astype(np.int64)
    umem = (mix64(base ^ (g * np.uint64(4) + np.uint64(2))) % np.uint64(101)).astype(np.int64)
    used = (mix64(base ^ (g * np.uint64(4) + np.uint64(3))) %
            np.uint64(MEM_TOTAL_MIB + 1)).astype(np.int64)
    rows = [GPU_HEADER[:-1] + ", power.draw [W]\n" if power else GPU_HEADER]
    watts = (mix64(base ^ (g * np.uint64(4))) % np.uint64(100001)).astype(np.int64)   # 0.00 .. 1000.00 W
    for i in range(ng):
        if snap.mig:                # a MIG parent reports no utilization
            row = "%s, [N/A], [N/A], %d MiB, %d MiB, %d MiB" % (snap.uuids[i], MEM_TOTAL_MIB, used[i],
                                                                MEM_TOTAL_MIB - used[i])
        else:
            row = "%s, %d %%, %d %%, %d MiB, %d MiB, %d MiB" % (snap.uuids[i], util[i], umem[i], MEM_TOTAL_MIB,
                                                                used[i], MEM_TOTAL_MIB - used[i])
        if power:
            row += ", %d.%02d W" % (watts[i] // 100, watts[i] % 100)
        rows.append(row + "\n")
    gpu_csv = "".join(rows)
    rows = [APPS_HEADER]
    # nvidia-smi lists compute apps GPU-major
    procs = sorted((gpu, ci, j, pid, name)
                   for ci, c in enumerate(snap.containers)
                   for j, (pid, gpu, name) in enumerate(c.gpu_procs))
    for gpu, ci, j, pid, name in procs:
        mem = 100 + h(seed, 11, node_idx, tick, ci, j) % 39901
        uuid = snap.uuids[gpu]
        if snap.mig == 2:
            uuid = "MIG-" + uuid[4:]
        rows.append("%d, %s, %s, %d MiB\n" % (pid, _PROC_NAMES[name], uuid, mem))
    return gpu_csv, "".join(rows)


@dataclass
class Replay:
    """A fleet batch: node snapshots + documents (node x tick), document-major by tick
    within node (all ticks of node 0, then node 1, ...)."""
    seed: int
    nodes: list                      # NodeSnap
    doc_node: np.ndarray             # int32 [n_docs] index into nodes
    doc_tick: np.ndarray             # int32 [n_docs]
    gpu_texts: list                  # str per doc
    apps_texts: list                 # str per doc

    @property
    def n_docs(self):
        return len(self.gpu_texts)

    def n_samples(self):
        return sum(t.count("\n") - 1 for t in self.gpu_texts)


def make_replay(seed=0, *, nodes=1, ticks=1, node_base=0, power=False, **node_kw) -> Replay:
    snaps = [make_node(seed, node_base + n, **node_kw) for n in range(nodes)]
    doc_node, doc_tick, gt, at = [], [], [], []
    for n, snap in enumerate(snaps):
        for t in range(ticks):
            a, b = make_texts(seed, node_base + n, snap, t, power=power)
            doc_node.append(n)
            doc_tick.append(t)
            gt.append(a)
            at.append(b)
    return Replay(seed, snaps, np.asarray(doc_node, np.int32), np.asarray(doc_tick, np.int32),
                  gt, at)


def make_skewed_replay(seed=0, *, hosts=40, epochs=3, ticks_per_epoch=4, gpus=8, max_containers=160,
                       zipf=1.2, scale=6, mig_every=3, node_base=0) -> Replay:
    """BASELINE configs[4] in miniature: short-lived pods and skewed series keys.  Every host
    has ``epochs`` docker snapshots (its pods are replaced from one epoch to the next), the
    number of containers per host follows a Zipf law (a few hosts carry most pods), every
    ``mig_every``-th host is a MIG host (utilization "[N/A]", idle pods see MIG uuids ->
    ``gpu_id=None``) and every other one of those also reports MIG uuids in its process rows,
    on which the reference raises after the node lines.  Documents are tick-major inside an
    epoch, i.e. consecutive documents belong to different hosts."""
    snaps, doc_node, doc_tick, gt, at = [], [], [], [], []
    for e in range(epochs):
        first = len(snaps)
        for hst in range(hosts):
            u = (h(seed, 20, hst) % 100000 + 1) / 100001.0
            nc = max(1, min(max_containers, int(scale / (u ** (1.0 / zipf)))))
            mig = 0
            if mig_every and hst % mig_every == mig_every - 1:
                mig = 2 if (hst // mig_every) % 2 else 1
            snap = make_node(seed + 1000 * e, node_base + hst, gpus=gpus, containers=nc,
                             procs_per_container=1 + h(seed, 21, hst, e) % 3, idle_every=3, ungated_every=5)
            snap.mig = mig
            if mig:
                for c in snap.containers:
                    c.visible = ["MIG-" + v[4:] for v in c.visible]
            snaps.append(snap)
        for t in range(ticks_per_epoch):
            tick = e * ticks_per_epoch + t
            for hst in range(hosts):
                a, b = make_texts(seed, node_base + hst, snaps[first + hst], tick)
                doc_node.append(first + hst)
                doc_tick.append(tick)
                gt.append(a)
                at.append(b)
    return Replay(seed, snaps, np.asarray(doc_node, np.int32), np.asarray(doc_tick, np.int32), gt, at)


# BASELINE.json `configs`, as generator parameters
CONFIGS = {
    # single node, 8 GPUs, 32 processes
    "config1": dict(nodes=1, ticks=1, gpus=8, containers=8, procs_per_container=4),
    # 64-node fleet batch: 512 GPUs, 2k pods, 60 s at 1 Hz
    "config2": dict(nodes=64, ticks=60, gpus=8, containers=32, procs_per_container=1),
    # 1k-node fleet: 8k GPUs, 40k processes, 10-min window
    "config3": dict(nodes=1000, ticks=600, gpus=8, containers=10, procs_per_container=4),
}
