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
        _lib.hc_key_hash_bytes.restype = C.c_uint64
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
                util=out[13], parse_error=out[14], node_error=out[15], power=out[16],
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


# ---- fast path of the fused kernel -----------------------------------------------------------
def _line_or_none(n, out):
    """(line bytes, guards intact) from a buffer with 8 guard bytes either side."""
    if n < 0:
        return None, True
    raw = bytes(out[:n + 16])
    return raw[8:8 + n], raw[:8] == b"\x5a" * 8 and raw[8 + n:] == b"\x5a" * 8


def fast_node_line(host, row: bytes, sid, shift=0, align=0):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_fast_node_line(host.encode(), row + b"\0" * 16, len(row), sid, shift, align, out,
                                C.byref(planned))
    line, guards = _line_or_none(n, out)
    return line, planned.value, guards


def fast_busy_line(host, ns, cid, user, pod, prow: bytes, grow: bytes, sid, shift=0, align=0):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_fast_busy_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                                prow + b"\0" * 16, len(prow), grow + b"\0" * 16, len(grow), sid, shift,
                                align, out, C.byref(planned))
    line, guards = _line_or_none(n, out)
    return line, planned.value, guards


def fast_idle_line(host, ns, cid, user, pod, sid, align=0):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_fast_idle_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                                0xFFFFFFFF if sid is None else sid, align, out, C.byref(planned))
    line, guards = _line_or_none(n, out)
    return line, planned.value, guards


def fast_gpu_row(row: bytes, shift=0):
    out = (C.c_uint32 * 16)()
    h = C.c_uint64()
    if not lib().hc_fast_gpu_row(row + b"\0" * 16, len(row), shift, out, C.byref(h)):
        return None
    f = lambda i: row[out[2 * i]:out[2 * i] + out[2 * i + 1]]
    return dict(key=f(0), util=f(1), mt=f(2), mu=f(3), mf=f(4), hash=h.value)


def fast_proc_row(row: bytes, shift=0):
    out = (C.c_uint32 * 16)()
    key = (C.c_ubyte * 48)()
    u64 = (C.c_uint64 * 2)()
    if not lib().hc_fast_proc_row(row + b"\0" * 16, len(row), shift, out, key, u64):
        return None
    f = lambda i: row[out[2 * i]:out[2 * i] + out[2 * i + 1]]
    return dict(pid=f(0), name=b"".join(f(1 + i) for i in range(4)), mem=f(5), key=bytes(key[:out[12]]),
                hash=u64[0], pidkey=u64[1])


# ---- k_poll2: lines written by the streaming line writer ---------------------------------------
def _two_lines_or_none(n, out, align):
    """(line, ok): the buffer holds 8 guard bytes, the line, and a second copy of it right behind
    (the next lane's line); ok = nothing touched in front of the word the line starts in (a line
    that starts inside a word stores zeros in its predecessor's bytes, which the predecessor
    ORs its last bytes into) and both copies identical."""
    if n < 0:
        return None, True
    raw = bytes(out[:2 * n + 16])
    z = align & 3
    front = raw[:8 - z] == b"\x5a" * (8 - z) and raw[8 - z:8] == b"\0" * z
    return raw[8:8 + n], front and raw[8 + n:8 + 2 * n] == raw[8:8 + n]


def p2_node_line(host, row: bytes, sid, shift=0, align=0, jobs=4):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_p2_node_line(host.encode(), row + b"\0" * 16, len(row), sid, shift, align, jobs, out,
                              C.byref(planned))
    line, guards = _two_lines_or_none(n, out, align)
    return line, planned.value, guards


def p2_busy_line(host, ns, cid, user, pod, prow: bytes, grow: bytes, sid, shift=0, align=0, jobs=4):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_p2_busy_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                              prow + b"\0" * 16, len(prow), grow + b"\0" * 16, len(grow), sid, shift,
                              align, jobs, out, C.byref(planned))
    line, guards = _two_lines_or_none(n, out, align)
    return line, planned.value, guards


def p2_idle_line(host, ns, cid, user, pod, sid, align=0, jobs=4):
    out = (C.c_ubyte * 20000)()
    planned = C.c_uint32()
    n = lib().hc_p2_idle_line(host.encode(), ns.encode(), cid.encode(), user.encode(), pod.encode(),
                              0xFFFFFFFF if sid is None else sid, align, jobs, out, C.byref(planned))
    line, guards = _two_lines_or_none(n, out, align)
    return line, planned.value, guards


def p2_gpu_row(row: bytes, shift=0):
    out = (C.c_uint32 * 16)()
    u64 = (C.c_uint64 * 3)()
    if not lib().hc_p2_gpu_row(row + b"\0" * 16, len(row), shift, out, u64):
        return None
    f = lambda i: row[out[2 * i]:out[2 * i] + out[2 * i + 1]]
    return dict(key=f(0), util=f(1), mt=f(2), mu=f(3), mf=f(4), hash=u64[0], total=u64[1], used=u64[2])


def key_hash_bytes(b: bytes) -> int:
    return lib().hc_key_hash_bytes(b, len(b))


def inspect_fields(text: bytes):
    """kgm_docker.cuh inspect_fields: None if the scanner leaves the container to the host, else
    (gated, podname or None, namespace or None, user or None) -- user only for a gated container."""
    out = (C.c_uint32 * 9)()
    buf = (C.c_ubyte * max(len(text), 1)).from_buffer_copy(text or b"\0")
    lib().hc_inspect_fields(buf, C.c_uint32(len(text)), out)
    if out[0]:
        return None
    cut = lambda o, n: text[o:o + n].decode()
    gated, has_pod = bool(out[1]), bool(out[2])
    return (gated, cut(out[3], out[4]) if has_pod else None, cut(out[5], out[6]) if has_pod else None,
            cut(out[7], out[8]) if gated and has_pod else None)


def p2_gpu_power(row: bytes, shift=0):
    """gpu_row_power of kgm_poll2_leaf.cuh: the value (0xffffffff = none), or None if it declines."""
    out = C.c_uint32()
    ok = lib().hc_p2_gpu_power(row, C.c_uint32(len(row)), C.c_uint32(shift), C.byref(out))
    return out.value if ok else None
