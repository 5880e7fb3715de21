"""GPU: kgm_encode_resident_async / kgm_encode_finish and the size header in front of the
lines (what the multi-GPU gather consumes on the device)."""
import ctypes as C

import numpy as np
import pytest
import torch

from fleet_util import entry_of_snap
from kgm import _ffi, monitor, replay, shard

pytestmark = pytest.mark.gpu


def _header(res):
    dev = torch.device("cuda", 0)
    t = shard.tensor_of_device_buffer(int(res.lines) - 16, 16, dev)
    return [int(x) for x in t.view(torch.int64).cpu().tolist()]


@pytest.mark.parametrize("pipeline", [_ffi.PIPELINE_AUTO, _ffi.PIPELINE_STAGED, _ffi.PIPELINE_FUSED_GENERIC])
def test_async_then_finish_equals_the_synchronous_call(pipeline):
    rp = replay.make_replay(seed=8, nodes=16, ticks=6, gpus=8, containers=10, procs_per_container=4)
    with monitor.Monitor(0) as mon:
        mon.set_pipeline(pipeline)
        mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
        mon.upload(mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
        mon.encode_resident()
        want = mon.fetch()
        for _ in range(3):
            early = mon.encode_resident_async()
            assert int(early.n_bytes) == 0 and int(early.n_lines) == 0 and int(early.lines) != 0
            done = mon.encode_finish()
            assert int(done.n_bytes) == want.n_bytes and int(done.n_lines) == want.n_lines
            assert int(done.lines) == int(early.lines)
            assert _header(done) == [want.n_bytes, want.n_lines]
            got = mon.fetch()
            assert got.body == want.body and np.array_equal(got.line_off, want.line_off)
        # the synchronous call writes the header too
        r = mon.encode_resident()
        assert _header(r) == [want.n_bytes, want.n_lines]
        assert mon.output_capacity >= want.n_bytes + 16 and mon.stream_handle != 0


def test_async_on_a_fresh_context_grows_the_workspace_in_finish():
    rp = replay.make_replay(seed=9, nodes=8, ticks=4, gpus=8, containers=40, procs_per_container=1, idle_every=1)
    with monitor.Monitor(0) as mon, monitor.Monitor(0) as ref:
        for m in (mon, ref):
            m.load_fleet([entry_of_snap(s) for s in rp.nodes])
            m.upload(m.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
        ref.encode_resident()
        want = ref.fetch()
        mon.encode_resident_async()                  # first run ever on this context: guesses only
        done = mon.encode_finish()
        assert int(done.n_bytes) == want.n_bytes
        assert _header(done) == [want.n_bytes, want.n_lines]
        assert mon.fetch().body == want.body


def test_finish_without_async_is_a_state_error():
    with monitor.Monitor(0) as mon:
        res = _ffi.KgmResult()
        assert mon._lib.kgm_encode_finish(mon._ctx, C.byref(res)) == _ffi.E_STATE
        assert mon._lib.kgm_encode_resident_async(mon._ctx, C.byref(res)) == _ffi.E_STATE   # nothing loaded


def test_output_into_a_callers_buffer():
    """kgm_set_output_buffer: header at the pointer, lines 16 bytes in; too small a buffer is
    KGM_E_LIMIT, never a write past its end; NULL goes back to the context's own buffer."""
    rp = replay.make_replay(seed=12, nodes=12, ticks=5, gpus=8, containers=10, procs_per_container=4)
    dev = torch.device("cuda", 0)
    with monitor.Monitor(0) as mon:
        mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
        mon.upload(mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
        own = mon.encode_resident()
        want = mon.fetch()
        own_ptr = int(own.lines)
        guard = 4096
        ext = torch.full((want.n_bytes + 16 + guard,), 0xEE, dtype=torch.uint8, device=dev)
        mon.set_output_buffer(ext.data_ptr(), want.n_bytes + 32)          # just enough (whole 16-byte units count)
        for _ in range(2):
            mon.encode_resident_async()
            r = mon.encode_finish()
            assert int(r.lines) == ext.data_ptr() + 16 and int(r.n_bytes) == want.n_bytes
            host = ext.cpu().numpy()
            assert host[:16].view(np.int64).tolist() == [want.n_bytes, want.n_lines]
            assert host[16:16 + want.n_bytes].tobytes() == want.body
            assert (host[32 + want.n_bytes:] == 0xEE).all()               # nothing past the end
            assert mon.fetch().body == want.body
        assert want.n_bytes <= mon.output_capacity < want.n_bytes + 32
        small = torch.full((want.n_bytes // 2 + guard,), 0xEE, dtype=torch.uint8, device=dev)
        mon.set_output_buffer(small.data_ptr(), want.n_bytes // 2)
        res = _ffi.KgmResult()
        assert mon._lib.kgm_encode_resident(mon._ctx, C.byref(res)) == _ffi.E_LIMIT
        torch.cuda.synchronize()
        assert (small.cpu().numpy()[(want.n_bytes // 2) // 16 * 16:] == 0xEE).all()
        assert mon._lib.kgm_set_output_buffer(mon._ctx, C.c_void_p(ext.data_ptr() + 4), 1 << 20) == _ffi.E_INVALID
        mon.set_output_buffer(0, 0)
        r = mon.encode_resident()
        assert int(r.lines) == own_ptr and mon.fetch().body == want.body
