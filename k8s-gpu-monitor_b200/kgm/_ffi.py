"""ctypes binding of include/kgm.h.

The shared library is built in-tree by ``__graft_entry__.build()`` (``kgm/build.py``) as
``kgm/libkgm.so``.  There is no fallback: if the library is missing or no CUDA device is
usable, importing callers get an exception, never a CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkgm.so")

KGM_OK = 0
E_INVALID, E_CUDA, E_NOMEM, E_LIMIT, E_STATE = -1, -2, -3, -4, -5

DOC_OK = 0
DOC_KEY_ERROR = 1
DOC_VALUE_ERROR = 2
DOC_ZERO_DIVISION = 3
DOC_ATTRIBUTE_ERROR = 4
DOC_INDEX_ERROR = 5
DOC_CSV_ERROR = 6
DOC_JSON_ERROR = 7
DOC_TYPE_ERROR = 8
DOC_OTHER_ERROR = 9
DOC_UNSUPPORTED_QUOTE = 20
DOC_UNSUPPORTED_BYTE = 21
DOC_UNSUPPORTED_RANGE = 22

# exception class name (as the reference would die with) <-> status code
DOC_CODE_OF_EXC = {
    "KeyError": DOC_KEY_ERROR, "ValueError": DOC_VALUE_ERROR,
    "ZeroDivisionError": DOC_ZERO_DIVISION, "AttributeError": DOC_ATTRIBUTE_ERROR,
    "IndexError": DOC_INDEX_ERROR, "Error": DOC_CSV_ERROR, "JSONDecodeError": DOC_JSON_ERROR,
    "TypeError": DOC_TYPE_ERROR,
}
EXC_OF_DOC_CODE = {v: k for k, v in DOC_CODE_OF_EXC.items()}


def doc_code_of_exception(exc: BaseException) -> int:
    return DOC_CODE_OF_EXC.get(type(exc).__name__, DOC_OTHER_ERROR)


class KgmStrings(C.Structure):
    _fields_ = [("bytes", C.c_void_p), ("off", C.c_void_p)]


class KgmFleet(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_uint32), ("node_name", KgmStrings), ("cont_begin", C.c_void_p),
        ("node_err_parse", C.c_void_p), ("node_err_join", C.c_void_p),
        ("n_containers", C.c_uint32), ("gated", C.c_void_p), ("err_inspect", C.c_void_p),
        ("err_emit", C.c_void_p), ("cid", KgmStrings), ("ns", KgmStrings), ("pod", KgmStrings),
        ("user", KgmStrings), ("pid_begin", C.c_void_p), ("pid", KgmStrings),
        ("vis_begin", C.c_void_p), ("vis", KgmStrings),
    ]


class KgmBatch(C.Structure):
    _fields_ = [
        ("n_docs", C.c_uint32), ("gpu_csv", C.c_void_p), ("gpu_off", C.c_void_p),
        ("apps_csv", C.c_void_p), ("apps_off", C.c_void_p), ("doc_node", C.c_void_p),
        ("doc_tick", C.c_void_p),
    ]


class KgmResult(C.Structure):
    _fields_ = [
        ("lines", C.c_void_p), ("n_bytes", C.c_uint64), ("line_off", C.c_void_p),
        ("n_lines", C.c_uint64), ("doc_line_off", C.c_void_p), ("doc_status", C.c_void_p),
        ("n_docs", C.c_uint32),
    ]


class KgmWindows(C.Structure):
    _fields_ = [("n", C.c_uint64), ("kind", C.c_void_p), ("node", C.c_void_p), ("container", C.c_void_p),
                ("gpu_id", C.c_void_p), ("window", C.c_void_p), ("count", C.c_void_p),
                ("m_count", C.c_void_p * 3), ("m_min", C.c_void_p * 3), ("m_max", C.c_void_p * 3),
                ("m_last", C.c_void_p * 3), ("m_sum", C.c_void_p * 3), ("m_mean", C.c_void_p * 3)]


class KgmReplaySpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_nodes", C.c_uint32), ("mem_total_mib", C.c_uint32),
                ("node_id", C.c_void_p), ("mig", C.c_void_p), ("gpu_begin", C.c_void_p), ("uuid", C.c_void_p),
                ("proc_begin", C.c_void_p), ("proc_pid", C.c_void_p), ("proc_ci", C.c_void_p),
                ("proc_j", C.c_void_p), ("proc_gpu", C.c_void_p), ("proc_name", C.c_void_p),
                ("names", C.c_void_p), ("name_off", C.c_void_p)]


class KgmInspectOut(C.Structure):
    _fields_ = [("status", C.c_void_p), ("gated", C.c_void_p), ("has_pod", C.c_void_p), ("pod_off", C.c_void_p),
                ("pod_len", C.c_void_p), ("ns_off", C.c_void_p), ("ns_len", C.c_void_p), ("user_off", C.c_void_p),
                ("user_len", C.c_void_p)]


class KgmTiming(C.Structure):
    _fields_ = [
        ("h2d_ms", C.c_float), ("tokenize_ms", C.c_float), ("join_ms", C.c_float),
        ("scan_ms", C.c_float), ("encode_ms", C.c_float), ("d2h_ms", C.c_float),
        ("total_ms", C.c_float), ("launches", C.c_uint32), ("fused", C.c_uint32),
        ("slow_docs", C.c_uint32), ("fused_ctas_per_sm", C.c_uint32), ("docs_in_flight_per_sm", C.c_uint32),
    ]


# every symbol include/kgm.h declares: (restype, argtypes)
SIGNATURES = {
    "kgm_abi_version": (C.c_int, []),
    "kgm_strerror": (C.c_char_p, [C.c_int]),
    "kgm_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "kgm_destroy": (None, [C.c_void_p]),
    "kgm_last_cuda_error": (C.c_char_p, [C.c_void_p]),
    "kgm_load_fleet": (C.c_int, [C.c_void_p, C.POINTER(KgmFleet)]),
    "kgm_encode_batch": (C.c_int, [C.c_void_p, C.POINTER(KgmBatch), C.c_int, C.POINTER(KgmResult)]),
    "kgm_upload_batch": (C.c_int, [C.c_void_p, C.POINTER(KgmBatch)]),
    "kgm_encode_resident": (C.c_int, [C.c_void_p, C.POINTER(KgmResult)]),
    "kgm_encode_resident_async": (C.c_int, [C.c_void_p, C.POINTER(KgmResult)]),
    "kgm_encode_finish": (C.c_int, [C.c_void_p, C.POINTER(KgmResult)]),
    "kgm_set_output_buffer": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "kgm_stream": (C.c_void_p, [C.c_void_p]),
    "kgm_stream_signal_flags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "kgm_stream_wait_flags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "kgm_push_copy_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p,
                                     C.c_uint32, C.POINTER(KgmResult)]),
    "kgm_push_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                C.POINTER(KgmResult)]),
    "kgm_output_capacity": (C.c_uint64, [C.c_void_p]),
    "kgm_fetch_result": (C.c_int, [C.c_void_p, C.POINTER(KgmResult)]),
    "kgm_get_timing": (C.c_int, [C.c_void_p, C.POINTER(KgmTiming)]),
    "kgm_last_kernel": (C.c_int, [C.c_void_p]),
    "kgm_set_pipeline": (C.c_int, [C.c_void_p, C.c_int]),
    "kgm_debug_stage_clocks": (C.c_int, [C.c_void_p]),
    "kgm_debug_chunk_times": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "kgm_concat_shards": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64,
                                    C.c_void_p, C.c_void_p]),
    "kgm_window_reduce": (C.c_int, [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]),
    "kgm_window_front": (C.c_int, [C.c_void_p]),
    "kgm_inspect_labels": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "kgm_replay_offsets": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "kgm_replay_fill": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p]),
    "kgm_flush_l2": (C.c_int, [C.c_void_p, C.c_uint64]),
    "kgm_host_alloc": (C.c_void_p, [C.c_uint64]),
    "kgm_host_free": (None, [C.c_void_p]),
    "kgm_tokenize_gpu_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p]),
}
NOALLOC_MAX_ROWS = 16384
ABI_VERSION = 3
PIPELINE_AUTO, PIPELINE_STAGED, PIPELINE_FUSED_GENERIC, PIPELINE_FUSED_V1, PIPELINE_FUSED_V2 = 0, 1, 2, 3, 4

_lib = None


class KgmError(RuntimeError):
    def __init__(self, code, what, detail=""):
        self.code = code
        super().__init__("%s failed: %d%s" % (what, code, (" (" + detail + ")") if detail else ""))


def load():
    """Loads libkgm.so and types every entry point.  Raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise FileNotFoundError(
                "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError = header / library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
