"""Python host API over libkgm.so -- the reference's ``main()`` seam, batched.

``gpu.main()`` (gpu.py:90-256) takes no arguments: it shells out for its texts and POSTs
each line.  Here the texts are arguments and the lines are the return value:

    mon = Monitor()                                  # one per process / GPU
    mon.load_fleet([node_entry(...), ...])           # docker-side state (kgm/fleet.py)
    res = mon.encode(gpu_texts, apps_texts, doc_node)   # one entry per poll ("document")
    res.body        -> bytes: every line + "\\n", in the reference's POST order
    res.lines(d)    -> the lines of document d
    res.status[d]   -> 0, or the class of the exception the reference would have died with

``poll()`` wraps that for a single document with the reference's own inputs.  All compute
happens in the CUDA library; without it (or without a GPU) construction fails.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _ffi
from .fleet import NodeEntry, PackedFleet, node_entry


def _check(rc, what, ctx=None, lib=None):
    if rc != _ffi.KGM_OK:
        detail = ""
        if lib is not None:
            detail = lib.kgm_strerror(rc).decode()
            if ctx and rc in (_ffi.E_CUDA, _ffi.E_NOMEM):
                detail += ": " + lib.kgm_last_cuda_error(ctx).decode()
        raise _ffi.KgmError(rc, what, detail)


class PinnedArray:
    """A numpy view over page-locked host memory from kgm_host_alloc."""

    def __init__(self, lib, nbytes):
        self._lib = lib
        self.ptr = lib.kgm_host_alloc(max(int(nbytes), 1))
        if not self.ptr:
            raise MemoryError("kgm_host_alloc(%d)" % nbytes)
        self.array = np.ctypeslib.as_array((C.c_uint8 * max(int(nbytes), 1)).from_address(self.ptr))

    def close(self):
        if self.ptr:
            self.array = None
            self._lib.kgm_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Batch:
    """Concatenated texts + offsets of a list of documents (kgm_batch), host side."""

    def __init__(self, gpu_texts, apps_texts, doc_node, doc_tick=None, lib=None, pinned=False):
        enc = lambda t: t.encode() if isinstance(t, str) else bytes(t)
        g = [enc(t) for t in gpu_texts]
        a = [enc(t) for t in apps_texts]
        if not (len(g) == len(a) == len(doc_node)):
            raise ValueError("gpu_texts, apps_texts and doc_node differ in length")
        self.n_docs = len(g)
        self.gpu_off = np.zeros(self.n_docs + 1, np.uint64)
        self.apps_off = np.zeros(self.n_docs + 1, np.uint64)
        if self.n_docs:
            np.cumsum([len(t) for t in g], out=self.gpu_off[1:])
            np.cumsum([len(t) for t in a], out=self.apps_off[1:])
        gb, ab = b"".join(g), b"".join(a)
        self._pins = []
        if pinned and lib is not None:
            pg, pa = PinnedArray(lib, len(gb)), PinnedArray(lib, len(ab))
            pg.array[:len(gb)] = np.frombuffer(gb, np.uint8)
            pa.array[:len(ab)] = np.frombuffer(ab, np.uint8)
            self._pins = [pg, pa]
            self.gpu_csv, self.apps_csv = pg.array, pa.array
        else:
            self.gpu_csv = np.frombuffer(gb + b"\0", np.uint8)
            self.apps_csv = np.frombuffer(ab + b"\0", np.uint8)
        self.gpu_bytes, self.apps_bytes = len(gb), len(ab)
        self.doc_node = np.ascontiguousarray(doc_node, np.uint32)
        self.doc_tick = None if doc_tick is None else np.ascontiguousarray(doc_tick, np.uint32)

    def struct(self):
        return _ffi.KgmBatch(
            self.n_docs, self.gpu_csv.ctypes.data, self.gpu_off.ctypes.data,
            self.apps_csv.ctypes.data, self.apps_off.ctypes.data, self.doc_node.ctypes.data,
            None if self.doc_tick is None else self.doc_tick.ctypes.data)


class HostResult:
    """A result copied to host memory (views are copied out of the context's buffers)."""

    def __init__(self, r: _ffi.KgmResult):
        self.n_docs, self.n_lines, self.n_bytes = int(r.n_docs), int(r.n_lines), int(r.n_bytes)
        view = lambda ptr, ctype, n: (np.ctypeslib.as_array((ctype * n).from_address(ptr)).copy()
                                      if n else np.zeros(0, ctype))
        self.body = bytes(view(r.lines, C.c_uint8, self.n_bytes))
        self.line_off = view(r.line_off, C.c_uint64, self.n_lines + 1)
        self.doc_line_off = view(r.doc_line_off, C.c_uint64, self.n_docs + 1)
        self.status = view(r.doc_status, C.c_int32, self.n_docs)

    def lines(self, d=None):
        """Lines (without the trailing newline) of document ``d``, or of the whole batch."""
        lo, hi = (0, self.n_lines) if d is None else (int(self.doc_line_off[d]),
                                                      int(self.doc_line_off[d + 1]))
        off = self.line_off
        return [self.body[int(off[i]):int(off[i + 1]) - 1].decode() for i in range(lo, hi)]

    def error(self, d):
        """Exception class name the reference would have raised for document ``d``."""
        code = int(self.status[d])
        if code == _ffi.DOC_OK:
            return None
        return _ffi.EXC_OF_DOC_CODE.get(code, "kgm:%d" % code)


class Monitor:
    def __init__(self, device=0):
        self._lib = _ffi.load()
        if self._lib.kgm_abi_version() != _ffi.ABI_VERSION:
            raise RuntimeError("libkgm.so ABI mismatch")
        self._ctx = C.c_void_p()
        _check(self._lib.kgm_create(C.byref(self._ctx), device), "kgm_create", None, self._lib)
        self._fleet = None
        self._batch = None
        self._timing_buf = _ffi.KgmTiming()
        self._result_buf = _ffi.KgmResult()
        self._async_buf = _ffi.KgmResult()

    # -- lifecycle ------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.kgm_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- fleet ----------------------------------------------------------------------------
    def load_fleet(self, nodes):
        packed = nodes if isinstance(nodes, PackedFleet) else PackedFleet(nodes)
        st = packed.struct()
        _check(self._lib.kgm_load_fleet(self._ctx, C.byref(st)), "kgm_load_fleet", self._ctx, self._lib)
        self._fleet = packed
        return packed

    # -- batches --------------------------------------------------------------------------
    def make_batch(self, gpu_texts, apps_texts, doc_node, doc_tick=None, pinned=False):
        return Batch(gpu_texts, apps_texts, doc_node, doc_tick, lib=self._lib, pinned=pinned)

    def encode(self, gpu_texts, apps_texts=None, doc_node=None, doc_tick=None) -> HostResult:
        """Host texts in, host lines out (kgm_encode_batch with host copy)."""
        batch = gpu_texts if isinstance(gpu_texts, Batch) else \
            self.make_batch(gpu_texts, apps_texts, doc_node, doc_tick)
        st, res = batch.struct(), _ffi.KgmResult()
        _check(self._lib.kgm_encode_batch(self._ctx, C.byref(st), 1, C.byref(res)),
               "kgm_encode_batch", self._ctx, self._lib)
        return HostResult(res)

    def encode_raw(self, batch: Batch, want_host_copy=True) -> _ffi.KgmResult:
        """kgm_encode_batch without copying the result out of the context's buffers."""
        st, res = batch.struct(), _ffi.KgmResult()
        _check(self._lib.kgm_encode_batch(self._ctx, C.byref(st), 1 if want_host_copy else 0,
                                          C.byref(res)), "kgm_encode_batch", self._ctx, self._lib)
        return res

    def upload(self, batch: Batch):
        st = batch.struct()
        _check(self._lib.kgm_upload_batch(self._ctx, C.byref(st)), "kgm_upload_batch", self._ctx, self._lib)
        self._batch = batch

    def encode_resident(self) -> _ffi.KgmResult:
        """Runs the pipeline on the resident batch; the result holds DEVICE pointers."""
        res = self._result_buf                       # reused: valid until the next call
        rc = self._lib.kgm_encode_resident(self._ctx, C.byref(res))
        if rc != _ffi.KGM_OK:
            _check(rc, "kgm_encode_resident", self._ctx, self._lib)
        return res

    def encode_resident_async(self) -> _ffi.KgmResult:
        """Queues the pipeline on the resident batch and returns at once: the result holds the
        DEVICE pointers, sizes 0; the 16 bytes in front of ``lines`` receive {n_bytes,
        n_lines} in stream order.  ``encode_finish`` waits and returns the sized result."""
        res = self._async_buf
        rc = self._lib.kgm_encode_resident_async(self._ctx, C.byref(res))
        if rc != _ffi.KGM_OK:
            _check(rc, "kgm_encode_resident_async", self._ctx, self._lib)
        return res

    def encode_finish(self) -> _ffi.KgmResult:
        res = self._result_buf
        rc = self._lib.kgm_encode_finish(self._ctx, C.byref(res))
        if rc != _ffi.KGM_OK:
            _check(rc, "kgm_encode_finish", self._ctx, self._lib)
        return res

    def set_output_buffer(self, dev_ptr: int, nbytes: int):
        """kgm_set_output_buffer: resident encodes write {header, lines} at ``dev_ptr`` (0: own buffer)."""
        _check(self._lib.kgm_set_output_buffer(self._ctx, C.c_void_p(dev_ptr or None), int(nbytes)),
               "kgm_set_output_buffer", self._ctx, self._lib)

    def push_step(self, slot_ptr: int, slot_bytes: int, wait_flag: int, wait_value: int, signal_flag: int,
                  signal_value: int, copy_engine=False) -> _ffi.KgmResult:
        """kgm_push_step: one step of a multi-GPU job in one call -- wait for a counter, encode the
        resident batch into ``slot_ptr``, bump a counter -- all on the context's stream.  With
        ``copy_engine`` (kgm_push_copy_step) the kernel writes the context's own buffer and a copy
        engine carries the result to the slot behind it."""
        res = self._async_buf
        fn = self._lib.kgm_push_copy_step if copy_engine else self._lib.kgm_push_step
        rc = fn(self._ctx, C.c_void_p(slot_ptr), int(slot_bytes), C.c_void_p(wait_flag or None),
                                     int(wait_value) & 0xFFFFFFFF, C.c_void_p(signal_flag or None),
                                     int(signal_value) & 0xFFFFFFFF, C.byref(res))
        if rc != _ffi.KGM_OK:
            _check(rc, "kgm_push_step", self._ctx, self._lib)
        return res

    @property
    def stream_handle(self) -> int:
        """The cudaStream_t the context queues its work on (kgm_stream)."""
        return int(self._lib.kgm_stream(self._ctx) or 0)

    @property
    def output_capacity(self) -> int:
        return int(self._lib.kgm_output_capacity(self._ctx))

    def fetch(self) -> HostResult:
        res = _ffi.KgmResult()
        _check(self._lib.kgm_fetch_result(self._ctx, C.byref(res)), "kgm_fetch_result", self._ctx, self._lib)
        return HostResult(res)

    def window_reduce(self, window_ticks) -> dict:
        """Stage 3 (kgm_window_reduce) on the resident batch, which must carry ``doc_tick``:
        per (series, window) count / min / max / sum / mean / last of gpu_util, mem and
        power, as numpy arrays (copies), ascending (kind, node|container, gpu_id, window)."""
        w = _ffi.KgmWindows()
        _check(self._lib.kgm_window_reduce(self._ctx, int(window_ticks), 1, C.byref(w)), "kgm_window_reduce",
               self._ctx, self._lib)
        n = int(w.n)
        view = lambda ptr, ctype: (np.ctypeslib.as_array((ctype * n).from_address(ptr)).copy()
                                   if n else np.zeros(0, ctype))
        out = {"n": n, "kind": view(w.kind, C.c_uint8)}
        for name in ("node", "container", "gpu_id", "window", "count"):
            out[name] = view(getattr(w, name), C.c_uint32)
        for m, metric in enumerate(("gpu_util", "mem", "power")):
            for stat, ctype in (("count", C.c_uint32), ("min", C.c_uint32), ("max", C.c_uint32),
                                ("last", C.c_uint32), ("sum", C.c_uint64), ("mean", C.c_double)):
                out["%s_%s" % (metric, stat)] = view(getattr(w, "m_" + stat)[m], ctype)
        return out

    def kernel_times(self):
        """(encode_ms, total_ms) of the last run -- the cheap accessor for timed loops."""
        t = self._timing_buf
        self._lib.kgm_get_timing(self._ctx, C.byref(t))
        return t.tokenize_ms, t.join_ms, t.scan_ms, t.encode_ms, t.total_ms

    def timing(self) -> dict:
        t = _ffi.KgmTiming()
        _check(self._lib.kgm_get_timing(self._ctx, C.byref(t)), "kgm_get_timing", self._ctx, self._lib)
        return {k: getattr(t, k) for k, _ in _ffi.KgmTiming._fields_}

    def window_front(self) -> str:
        """The front end the last window_reduce took its samples from."""
        return {1: "k_poll2_samples", 0: "k_tokenize, k_join"}[self._lib.kgm_window_front(self._ctx)]

    def last_kernel(self) -> str:
        """The kernel(s) the last run went through."""
        k = self._lib.kgm_last_kernel(self._ctx)
        return {0: "k_tokenize, k_join, k_scan, k_encode", 1: "k_poll", 2: "k_poll2"}[k]

    def set_pipeline(self, mode):
        """PIPELINE_AUTO (fused kernel when the batch fits), PIPELINE_STAGED, PIPELINE_FUSED_GENERIC."""
        _check(self._lib.kgm_set_pipeline(self._ctx, int(mode)), "kgm_set_pipeline", self._ctx, self._lib)

    def flush_l2(self, nbytes=256 << 20):
        _check(self._lib.kgm_flush_l2(self._ctx, nbytes), "kgm_flush_l2", self._ctx, self._lib)


def poll(mon: Monitor, node_name, gpu_csv, apps_csv, ps_text, stats_text, top, inspect, smi_list):
    """One document through the library, with the inputs of one reference ``main()`` call.
    Returns ``(lines, error_class_name_or_None)`` like the test oracle does."""
    mon.load_fleet([node_entry(node_name, ps_text, stats_text, top, inspect, smi_list)])
    res = mon.encode([gpu_csv], [apps_csv], [0])
    return res.lines(0), res.error(0)


__all__ = ["Monitor", "Batch", "HostResult", "NodeEntry", "node_entry", "poll"]
