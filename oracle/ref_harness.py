"""TEST INFRASTRUCTURE ONLY -- drives the UNMODIFIED reference script in-process.

Runs `/root/reference/gpu.py` `main()` (gpu.py:90-256) for one poll with its three I/O
seams replaced:

  * ``gpu.command``        (gpu.py:36-37)   -> canned docker / nvidia-smi texts
  * ``gpu.os.popen``       (gpu.py:221-223) -> canned ``nvidia-smi -L | cut`` list
  * ``gpu.requests.post``  (gpu.py:145,218,252) -> capture of the POSTed line

The reference tree exists only in the build container.  ``oracle/make_ref.py`` (run by
``__graft_entry__.build()``) puts an unmodified copy of the script into the git-ignored
``oracle/_ref/``, which travels to the GPU box; there this harness is used by ``bench.py``'s
two CPU legs only (``--impl reference`` and ``cpu_baseline``), never by the product.  In the
build container it is also used by ``tests/golden/make_golden.py`` (fixture generation) and
by the CPU-only differential test that pins ``oracle/poll_oracle.py`` against the real script.
"""
from __future__ import annotations

import contextlib
import importlib.util
import io
import os
import warnings

import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_DIR = os.environ.get("KGM_REFERENCE_DIR", "/root/reference")
if not os.path.isfile(os.path.join(REFERENCE_DIR, "gpu.py")):
    REFERENCE_DIR = os.path.join(_HERE, "_ref")          # the copy oracle/make_ref.py made
_gpu = None


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "gpu.py"))


def _ensure_requests():
    """gpu.py imports `requests` only to POST (gpu.py:145); the harness replaces that call, so
    a box without the package gets an empty stand-in module."""
    try:
        import requests  # noqa: F401
    except Exception:
        stub = types.ModuleType("requests")
        stub.post = lambda *a, **k: None
        sys.modules["requests"] = stub


def _load():
    global _gpu
    if _gpu is None:
        _ensure_requests()
        spec = importlib.util.spec_from_file_location(
            "_kgm_reference_gpu", os.path.join(REFERENCE_DIR, "gpu.py"))
        mod = importlib.util.module_from_spec(spec)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")          # '\D' SyntaxWarnings on 3.12
            with contextlib.redirect_stdout(io.StringIO()):
                spec.loader.exec_module(mod)         # prints node_name / influxdb_url
        _gpu = mod
    return _gpu


class _Resp:
    status_code = 204


class _Pipe:
    def __init__(self, text):
        self._text = text

    def read(self):
        return self._text

    def close(self):
        return None


def run_reference(node_name, gpu_csv, apps_csv, ps_text, stats_text, top, inspect, smi_list):
    """One `main()` call.  Returns (lines POSTed in order, exception class name or None)."""
    gpu = _load()
    posted = []

    def fake_command(args):
        if args[0] == "docker":
            if args[1] == "ps":
                return ps_text
            if args[1] == "stats":
                return stats_text
            if args[1] == "top":
                return top[args[2]]
            if args[1] == "inspect":
                return inspect[args[2]]
        if args[0] == "nvidia-smi":
            if args[2].startswith("--query-gpu="):
                return gpu_csv
            if args[2].startswith("--query-compute-apps="):
                return apps_csv
        raise AssertionError("unexpected argv %r" % (args,))

    def fake_popen(cmd):
        # "docker exec -i <cid> nvidia-smi -L |cut ..." (gpu.py:221)
        cid = cmd.split("docker exec -i ", 1)[1].split(" ", 1)[0]
        return _Pipe(smi_list.get(cid, ""))

    def fake_post(url, data=None, **kw):
        posted.append(data)
        return _Resp()

    saved = (gpu.command, gpu.os.popen, gpu.requests.post, os.environ.get("NODE_NAME"))
    gpu.command, gpu.os.popen, gpu.requests.post = fake_command, fake_popen, fake_post
    if node_name is None:
        os.environ.pop("NODE_NAME", None)
    else:
        os.environ["NODE_NAME"] = node_name
    err = None
    try:
        with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            try:
                gpu.main()
            except Exception as exc:  # the reference has no try/except: capture the class
                err = type(exc).__name__
    finally:
        gpu.command, gpu.os.popen, gpu.requests.post = saved[:3]
        if saved[3] is None:
            os.environ.pop("NODE_NAME", None)
        else:
            os.environ["NODE_NAME"] = saved[3]
    return posted, err
