"""GPU (two or more): contexts on different devices in ONE process give the same bytes.  The kernels'
shared-memory opt-in (cudaFuncSetAttribute) is per device; a process that had only set it on the
first device it used would fail every launch on the second one."""
import pytest

from fleet_util import entry_of_snap, oracle_of_replay_doc
from kgm import _ffi, monitor, replay
from oracle import poll_oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("pipeline", ["fused-v1", "fused-v2", "staged"])
def test_contexts_on_two_devices_in_one_process(pipeline):
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two GPUs")
    mode = {"fused-v1": _ffi.PIPELINE_FUSED_V1, "fused-v2": _ffi.PIPELINE_FUSED_V2, "staged": _ffi.PIPELINE_STAGED}[pipeline]
    rp = replay.make_replay(seed=17, nodes=6, ticks=5, gpus=8, containers=9, procs_per_container=3, idle_every=4)
    want = b"".join(poll_oracle.body_of(oracle_of_replay_doc(rp, d)[0]) for d in range(rp.n_docs))
    bodies = []
    mons = [monitor.Monitor(0), monitor.Monitor(1)]
    try:
        for rounds in range(2):                      # device 0, device 1, then both again
            for m in mons:
                m.set_pipeline(mode)
                m.load_fleet([entry_of_snap(s) for s in rp.nodes])
                res = m.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
                assert not res.status.any()
                bodies.append(res.body)
                if pipeline != "staged":
                    m.window_reduce(2)               # k_poll2_samples opts in on its own
    finally:
        for m in mons:
            m.close()
    assert all(b == want for b in bodies)
