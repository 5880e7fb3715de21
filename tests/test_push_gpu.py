"""GPU, two or more devices: the multi-GPU step on hardware.  Every rank encodes its own hosts
straight into its slot on the step's gathering GPU (peer memory, kgm_push_step) and the ranks
synchronise through counters in that memory; the gathered slots, packed, must be the oracle's
rank-major concatenation byte for byte."""
import os
import socket

import pytest
import torch

from fleet_util import entry_of_snap, oracle_of_replay_doc
from oracle import poll_oracle

pytestmark = pytest.mark.gpu


NCTX = 4


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_replay(rank):
    from kgm import replay
    return replay.make_replay(seed=31, nodes=3 + rank, ticks=4, gpus=4, containers=5, procs_per_container=2,
                              idle_every=4, node_base=10 * rank)


def _worker(rank, world, port, steps, q):
    import torch.distributed as dist
    from kgm import monitor, shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    try:
        rp = _rank_replay(rank)
        mons = []
        for _ in range(NCTX):                            # several contexts per rank, visited in turn as bench.py does
            m = monitor.Monitor(rank)
            m.load_fleet([entry_of_snap(s) for s in rp.nodes])
            m.upload(m.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
            m.encode_resident()                          # sizes the workspace
            mons.append(m)
        cap = int(mons[0].encode_resident().n_bytes) + (1 << 16)
        gather = shard.ShardGather(cap, device, push=True, rotate_root=True, contexts=NCTX)
        assert gather.mode == "push", "CUDA IPC between the ranks is needed"
        pending = [False] * NCTX
        for k in range(steps):
            m = mons[k % NCTX]
            if pending[k % NCTX]:
                m.encode_finish()
            gather.push_step(m)
            pending[k % NCTX] = True
        for i, m in enumerate(mons):
            if pending[i]:
                m.encode_finish()
        gather.finish()
        got = None
        if rank == gather.last_root:
            got = (gather.shard_offsets(), gather.body().cpu().numpy().tobytes())
        q.put((rank, "ok", gather.last_root, got))
        dist.barrier()
        for m in mons:
            m.set_output_buffer(0, 0)
            m.close()
    except Exception as exc:  # pragma: no cover - reported to the parent
        import traceback
        q.put((rank, "%s: %s\n%s" % (type(exc).__name__, exc, traceback.format_exc()), None, None))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("steps", [1, 7, 150])
def test_push_gather_is_the_rank_major_concatenation(steps):
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, steps, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in results), results
    root = results[0][2]
    assert root == (steps - 1) % world
    offsets, body = next(r[3] for r in results if r[0] == root)
    want = []
    for rank in range(world):
        rp = _rank_replay(rank)
        b = b""
        for d in range(rp.n_docs):
            lines, err = oracle_of_replay_doc(rp, d)
            assert err is None
            b += poll_oracle.body_of(lines)
        want.append(b)
    assert [offsets[r + 1] - offsets[r] for r in range(world)] == [len(b) for b in want]
    assert body == b"".join(want)
