"""ctypes view of tests/native/_hostcheck.so (host build of the kernels' leaf functions)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-monitor_b200"))
from kgm import build  # noqa: E402

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build.build_hostcheck())
        _lib.hc_hash_bytes.restype = C.c_uint64
    return _lib


def float_repr(x: float) -> str:
    buf = (C.c_ubyte * 32)()
    n = lib().hc_float_repr(C.c_double(x), buf)
    return bytes(buf[:n]).decode()


def mem_util_repr(used: int, total: int) -> str:
    buf = (C.c_ubyte * 32)()
    n = lib().hc_mem_util_repr(C.c_uint64(used), C.c_uint64(total), buf)
    return bytes(buf[:n]).decode()


def float_repr_many(xs):
    x = np.ascontiguousarray(xs, np.float64)
    out = np.zeros((len(x), 24), np.uint8)
    ln = np.zeros(len(x), np.int32)
    lib().hc_float_repr_many(C.c_void_p(x.ctypes.data), C.c_int64(len(x)), C.c_void_p(out.ctypes.data),
                             C.c_void_p(ln.ctypes.data))
    return [bytes(out[i, :ln[i]]).decode() for i in range(len(x))]


def scan_gpu_row(row: bytes):
    out = (C.c_uint32 * 32)()
    u64 = (C.c_uint64 * 4)()
    lib().hc_scan_gpu_row(row + b"\0" * 16, 0, len(row), out, u64)
    return dict(ncells=out[0], flags=out[1], key_lo=out[2], cell_end=list(out[3:9]), nd=list(out[9:13]),
                util=out[13], parse_error=out[14], node_error=out[15],
                key_hash=u64[0], mem_total=u64[1], mem_used=u64[2])


def scan_proc_row(row: bytes):
    out = (C.c_uint32 * 32)()
    u64 = (C.c_uint64 * 4)()
    lib().hc_scan_proc_row(row + b"\0" * 16, 0, len(row), out, u64)
    return dict(ncells=out[0], flags=out[1], key_lo=out[2], cell_end=list(out[3:7]), name_len=out[7],
                nd_mem=out[8], pid_len=out[9], used_mem=out[10], parse_error=out[11],
                uuid_hash=u64[0], pid_lo=u64[1], pid_hi=u64[2])


def node_line(host: str, row: bytes, sid: int):
    out = (C.c_ubyte * 65536)()
    planned = C.c_uint32()
    n = lib().hc_node_line(host.encode(), row + b"\0" * 16, len(row), sid, out, C.byref(planned))
    return (bytes(out[:n]) if n >= 0 else None), planned.value, n


def busy_line(host, ns, cid, user, pod, prow: bytes, grow: bytes, sid: int):
    out = (C.c_ubyte * 65536)()
    planned = C.c_uint32()
    n = lib().hc_busy_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                           prow + b"\0" * 16, len(prow), grow + b"\0" * 16, len(grow), sid, out,
                           C.byref(planned))
    return (bytes(out[:n]) if n >= 0 else None), planned.value, n


def idle_line(host, ns, cid, user, pod, sid):
    out = (C.c_ubyte * 65536)()
    planned = C.c_uint32()
    n = lib().hc_idle_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                           0xFFFFFFFF if sid is None else sid, out, C.byref(planned))
    return bytes(out[:n]), planned.value


def hash_bytes(b: bytes) -> int:
    return lib().hc_hash_bytes(b, len(b))
