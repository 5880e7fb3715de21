"""CPU: the C-ABI library builds, loads and exports every symbol include/kgm.h declares."""
import ctypes as C
import os
import re

from kgm import _ffi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kgm.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(kgm_[a-z0-9_]+)\s*\(", text))


def test_library_builds_and_exports_the_declared_abi():
    path = build.build_library()
    assert os.path.isfile(path)
    lib = C.CDLL(path)
    declared = _declared_symbols()
    assert declared == set(_ffi.SIGNATURES), declared ^ set(_ffi.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name


def test_version_strings_and_struct_layout_without_a_gpu():
    lib = _ffi.load()
    assert lib.kgm_abi_version() == _ffi.ABI_VERSION == 3
    assert lib.kgm_strerror(0) == b"ok"
    assert lib.kgm_strerror(_ffi.DOC_KEY_ERROR) == b"KeyError"
    assert lib.kgm_strerror(_ffi.DOC_CSV_ERROR) == b"Error"
    for name, code in _ffi.DOC_CODE_OF_EXC.items():
        assert lib.kgm_strerror(code).decode() == name
    # layout the C side assumes (x86-64)
    assert C.sizeof(_ffi.KgmStrings) == 16
    assert C.sizeof(_ffi.KgmBatch) == 56
    assert C.sizeof(_ffi.KgmResult) == 56
    assert C.sizeof(_ffi.KgmTiming) == 48
    assert C.sizeof(_ffi.KgmReplaySpec) == 8 + 4 + 4 + 12 * 8
    assert C.sizeof(_ffi.KgmInspectOut) == 9 * 8
    assert C.sizeof(_ffi.KgmFleet) == 8 + 16 + 8 * 3 + 8 + 8 * 3 + 16 * 4 + 8 + 16 + 8 + 16


def test_every_kernel_in_the_library_is_the_repository_s_own():
    """No library kernels (CUB sort / scan and the like): every device function of libkgm.so is
    defined in csrc/ under the kgm namespace."""
    import shutil
    import subprocess
    import pytest
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    out = subprocess.run(["cuobjdump", "-res-usage", build.build_library()], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL, text=True).stdout
    funcs = re.findall(r"Function (\S+):", out)
    assert len(funcs) >= 10
    foreign = [f for f in funcs if not f.startswith("_ZN3kgm")]
    assert not foreign, foreign
