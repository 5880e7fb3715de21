"""The replay of kgm/replay.py generated on the GPU (kgm_replay_offsets / kgm_replay_fill).

``DeviceReplay(nodes, node_ids, seed)`` uploads what the generator needs to know about a set of
hosts (uuids, the GPU-major process list, MIG mode: the ``NodeSnap`` objects of replay.make_node);
``window(tick0, n_ticks)`` then writes the nvidia-smi texts of those hosts' polls for the ticks
``tick0 .. tick0 + n_ticks - 1`` into device memory, host-major, byte for byte what
``replay.make_texts`` returns for the same coordinates (tests/test_replay_gpu.py), and
``DeviceWindow.upload(mon)`` makes it the resident batch of a ``Monitor`` without the texts ever
visiting the host.  Bench / test plumbing: the reference reads live tool output (gpu.py:104-116).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _ffi, replay


class DeviceWindow:
    """The texts of one window on the device + their offsets on the host."""

    def __init__(self, gpu_csv, apps_csv, gpu_off, apps_off, doc_node, doc_tick):
        self.gpu_csv, self.apps_csv = gpu_csv, apps_csv          # torch uint8, device
        self.gpu_off, self.apps_off = gpu_off, apps_off          # numpy uint64 [n_docs + 1], host
        self.doc_node, self.doc_tick = doc_node, doc_tick        # numpy uint32 [n_docs], host
        self.n_docs = len(doc_node)

    @property
    def gpu_bytes(self):
        return int(self.gpu_off[-1])

    @property
    def apps_bytes(self):
        return int(self.apps_off[-1])

    def texts(self, d):
        """(gpu text, apps text) of document d as str (copies from the device: tests)."""
        g = self.gpu_csv[int(self.gpu_off[d]):int(self.gpu_off[d + 1])].cpu().numpy().tobytes().decode()
        a = self.apps_csv[int(self.apps_off[d]):int(self.apps_off[d + 1])].cpu().numpy().tobytes().decode()
        return g, a

    def upload(self, mon):
        """Makes the window the resident batch of ``mon`` (device-to-device copy of the texts)."""
        st = _ffi.KgmBatch()
        st.n_docs = self.n_docs
        for name, ptr in (("gpu_csv", self.gpu_csv.data_ptr()), ("apps_csv", self.apps_csv.data_ptr()),
                          ("gpu_off", self.gpu_off.ctypes.data), ("apps_off", self.apps_off.ctypes.data),
                          ("doc_node", self.doc_node.ctypes.data), ("doc_tick", self.doc_tick.ctypes.data)):
            setattr(st, name, ptr)
        rc = mon._lib.kgm_upload_batch(mon._ctx, C.byref(st))
        if rc != _ffi.KGM_OK:
            raise RuntimeError("kgm_upload_batch failed: %d" % rc)
        mon._batch = self                               # keeps the device texts alive


class DeviceReplay:
    def __init__(self, nodes, node_ids, seed, device=0):
        import torch
        self.torch = torch
        self.device = torch.device("cuda", device)
        self.lib = _ffi.load()
        self.nodes, self.node_ids, self.seed = list(nodes), [int(n) for n in node_ids], int(seed)
        gpu_begin, proc_begin = [0], [0]
        uuid = bytearray()
        pid, ci_, j_, gpu_, name_ = [], [], [], [], []
        for snap in self.nodes:
            for u in snap.uuids:
                b = u.encode()
                if len(b) != 40:
                    raise ValueError("uuid of %d bytes (the generator writes 40)" % len(b))
                uuid += b
            gpu_begin.append(gpu_begin[-1] + len(snap.uuids))
            procs = sorted((gpu, ci, j, p, name) for ci, c in enumerate(snap.containers)
                           for j, (p, gpu, name) in enumerate(c.gpu_procs))           # replay.make_texts order
            for gpu, ci, j, p, name in procs:
                if ci >= 1 << 16 or j >= 256 or gpu >= 256:
                    raise ValueError("container / process ordinal beyond the spec's field widths")
                pid.append(p); ci_.append(ci); j_.append(j); gpu_.append(gpu); name_.append(name)
            proc_begin.append(proc_begin[-1] + len(procs))
        names = [s.encode() for s in replay._PROC_NAMES]
        name_off = np.cumsum([0] + [len(s) for s in names]).astype(np.uint32)

        def dev(a, dtype):
            return torch.from_numpy(np.array(a, dtype=dtype)).to(self.device)
        self._keep = {
            "node_id": dev(self.node_ids, np.uint32), "mig": dev([s.mig for s in self.nodes], np.uint8),
            "gpu_begin": dev(gpu_begin, np.uint32), "uuid": dev(np.frombuffer(bytes(uuid) or b"\0", np.uint8), np.uint8),
            "proc_begin": dev(proc_begin, np.uint32), "proc_pid": dev(pid or [0], np.uint32),
            "proc_ci": dev(ci_ or [0], np.uint16), "proc_j": dev(j_ or [0], np.uint8),
            "proc_gpu": dev(gpu_ or [0], np.uint8), "proc_name": dev(name_ or [0], np.uint8),
            "names": dev(np.frombuffer(b"".join(names), np.uint8), np.uint8), "name_off": dev(name_off, np.uint32),
        }
        sp = _ffi.KgmReplaySpec()
        sp.seed, sp.n_nodes, sp.mem_total_mib = self.seed, len(self.nodes), replay.MEM_TOTAL_MIB
        for k, t in self._keep.items():
            setattr(sp, k, t.data_ptr())
        self.spec = sp

    def window(self, tick0, n_ticks) -> DeviceWindow:
        torch = self.torch
        n = len(self.nodes) * int(n_ticks)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        scratch = torch.empty(2 * n, dtype=torch.int32, device=self.device)
        offs = torch.empty((2, n + 1), dtype=torch.int64, device=self.device)
        rc = self.lib.kgm_replay_offsets(C.byref(self.spec), int(tick0), int(n_ticks), scratch.data_ptr(),
                                         offs[0].data_ptr(), offs[1].data_ptr(), stream)
        if rc != _ffi.KGM_OK:
            raise RuntimeError("kgm_replay_offsets failed: %d" % rc)
        host = offs.cpu().numpy().astype(np.uint64)               # (the sizes: one small copy per window)
        gpu_off, apps_off = np.ascontiguousarray(host[0]), np.ascontiguousarray(host[1])
        gpu_csv = torch.empty(max(int(gpu_off[-1]), 1), dtype=torch.uint8, device=self.device)
        apps_csv = torch.empty(max(int(apps_off[-1]), 1), dtype=torch.uint8, device=self.device)
        rc = self.lib.kgm_replay_fill(C.byref(self.spec), int(tick0), int(n_ticks), offs[0].data_ptr(),
                                      offs[1].data_ptr(), gpu_csv.data_ptr(), apps_csv.data_ptr(), stream)
        if rc != _ffi.KGM_OK:
            raise RuntimeError("kgm_replay_fill failed: %d" % rc)
        torch.cuda.current_stream(self.device).synchronize()      # offs / scratch may be freed
        doc_node = np.repeat(np.arange(len(self.nodes), dtype=np.uint32), int(n_ticks))
        doc_tick = np.tile(np.arange(int(tick0), int(tick0) + int(n_ticks), dtype=np.uint32), len(self.nodes))
        return DeviceWindow(gpu_csv, apps_csv, gpu_off, apps_off, doc_node, doc_tick)
