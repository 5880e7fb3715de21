"""CPU: the multi-GPU plumbing (kgm/shard.py) with two gloo ranks.

Shards are whole hosts: every rank encodes its own hosts and the only
exchange is the all-gather of {bytes, lines} plus the rank-major concatenation of the
encoded buffers.  Here the per-rank "encoded buffer" comes from the oracle (no GPU in this
container); what is under test is the partition and the assembly, which are the same code
the NCCL run uses.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fleet_util import oracle_of_replay_doc
from kgm import replay, shard
from oracle import poll_oracle


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard_bodies(world):
    """Oracle bodies of a small fleet, split by assign_hosts: (per-rank bytes, per-rank lines)."""
    rp = replay.make_replay(seed=21, nodes=5, ticks=3, gpus=4, containers=6, procs_per_container=2,
                            idle_every=3)
    names = [s.name for s in rp.nodes]
    weights = [sum(len(rp.gpu_texts[d]) + len(rp.apps_texts[d]) for d in range(rp.n_docs)
                   if rp.doc_node[d] == n) for n in range(len(names))]
    owner = shard.assign_hosts(names, weights, world)
    bodies, lines = [b""] * world, [0] * world
    for d in range(rp.n_docs):
        got, err = oracle_of_replay_doc(rp, d)
        assert err is None
        r = owner[int(rp.doc_node[d])]
        bodies[r] += poll_oracle.body_of(got)
        lines[r] += len(got)
    return bodies, lines, owner, names, weights


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bodies, lines, owner, names, weights = _shard_bodies(world)
        # every rank derives the same host -> rank map without talking
        assert owner == shard.assign_hosts(list(names), list(weights), world)
        mine = np.frombuffer(bodies[rank] + b"\0" * 7, np.uint8).copy()
        local = torch.from_numpy(mine)
        counts = shard.gather_counts(len(bodies[rank]), lines[rank], torch.device("cpu"))
        assert counts[:, 0].tolist() == [len(b) for b in bodies]
        assert counts[:, 1].tolist() == lines
        off = shard.slice_offsets(counts)
        assert off[:, 0].tolist() == list(np.cumsum([0] + [len(b) for b in bodies]))
        whole, off2 = shard.concat_encoded(local, counts)
        assert torch.equal(off, off2)
        assert bytes(whole.numpy().tobytes()) == b"".join(bodies)
        # second round into a caller-provided buffer (bench.py reuses it every step)
        out = torch.zeros(len(b"".join(bodies)) + 64, dtype=torch.uint8)
        whole2, _ = shard.concat_encoded(local, counts, out=out)
        assert bytes(whole2.numpy().tobytes()) == b"".join(bodies)
        # asynchronous gather of the shards to rank 0 (what bench.py runs every step), twice
        sg = shard.ShardGather(max(len(b) for b in bodies) + 16, torch.device("cpu"))
        for _ in range(3):
            sg.start(local, len(bodies[rank]), lines[rank])
        sg.finish()
        if rank == 0:
            assert bytes(sg.body().numpy().tobytes()) == b"".join(bodies)
        q.put((rank, "ok"))
    except Exception as exc:  # pragma: no cover - reported to the parent
        q.put((rank, "%s: %s" % (type(exc).__name__, exc)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_ranks_gather_counts_and_concatenate_rank_major():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=150) for _ in range(world))
    for p in procs:
        p.join(timeout=30)
    assert results == {0: "ok", 1: "ok"}


def test_assign_hosts_is_balanced_and_deterministic():
    names = ["k8s-node-gpu-%05d" % i for i in range(37)]
    weights = [1000 + (i * 7919) % 5000 for i in range(37)]
    weights[3] = 60000                                # one dominant host (config 5 skew)
    for world in (1, 2, 4, 8):
        a = shard.assign_hosts(names, weights, world)
        assert a == shard.assign_hosts(list(names), list(weights), world)
        assert set(a) <= set(range(world))
        load = [sum(w for w, r in zip(weights, a) if r == k) for k in range(world)]
        # largest-first onto the least loaded rank: no rank exceeds the mean by more than the largest host
        assert max(load) <= sum(weights) / world + max(weights)
    perm = list(reversed(range(37)))                  # order of presentation does not matter
    b = shard.assign_hosts([names[i] for i in perm], [weights[i] for i in perm], 4)
    a = shard.assign_hosts(names, weights, 4)
    assert [b[perm.index(i)] for i in range(37)] == a


def test_shards_cover_every_document_once():
    bodies, lines, owner, names, _ = _shard_bodies(2)
    assert len(owner) == len(names) and set(owner) == {0, 1}
    assert sum(lines) > 0 and all(len(b) > 0 for b in bodies)


def test_push_mode_slot_arithmetic():
    """Push gather (CUDA only at run time): which slot a rank writes at step k, on which
    gathering rank, and that a slot set is never reused inside any rank's run-ahead window."""
    from kgm.shard import ShardGather
    for world in (2, 3, 4, 8):
        for rotate in (False, True):
            stride = 4096
            sets = 4
            ahead = [2 * world if rotate else 3 for r in range(world)]     # the widest window shard.py uses
            gs = []
            for r in range(world):
                g = ShardGather.__new__(ShardGather)
                g.world, g.rank, g.stride, g.rotate_root, g.PUSH_SETS = world, r, stride, rotate, sets
                gs.append(g)
            last_use = {}                                # (root, set offset) -> step
            for k in range(6 * world * sets):
                roots = {g.push_root(k) for g in gs}
                assert len(roots) == 1                   # every rank agrees on the gathering rank
                root = roots.pop()
                assert root == (k % world if rotate else 0)
                off = {g._push_offset(k) for g in gs}
                assert len(off) == 1                     # ... and on the slot set
                base = off.pop()
                assert base % (world * stride) == 0 and base + world * stride <= sets * world * stride
                spans = sorted((base + g.rank * stride, base + (g.rank + 1) * stride) for g in gs)
                assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))     # the ranks' slots are disjoint
                if (root, base) in last_use:
                    # a writer of step k has waited for the gathering rank's acknowledgement of step
                    # k - ahead (every shard of that step has landed, so every rank is past the
                    # set's last use)
                    assert k - max(ahead) >= last_use[(root, base)]
                last_use[(root, base)] = k


def test_push_mode_counter_arithmetic():
    """Every (rank, context) pair has its own "landed" counter on the gathering rank (so a plain
    store is as good as an atomic max), the acknowledgement counter lies behind them, a counter's
    values only grow, and every rank waits on the counters the writers of the same step raise."""
    from kgm.shard import ShardGather
    for world in (2, 3, 4, 8):
        for contexts in (1, 2, 4, 8):
            gs = []
            for r in range(world):
                g = ShardGather.__new__(ShardGather)
                g.world, g.rank, g.rotate_root, g.contexts = world, r, True, contexts
                gs.append(g)
            g0 = gs[0]
            assert 4 * (g0._ack_idx() + 1) <= g0.FLAG_BYTES and g0.FLAG_BYTES % 256 == 0
            last = {}                                    # (root, counter index) -> last value written
            for k in range(5 * world * contexts):
                root, value = g0.push_root(k), g0._uses(k) + 1
                idx = [g._landed_idx(k, g.rank) for g in gs]
                assert len(set(idx)) == world and max(idx) < g0._ack_idx()
                # the gathering rank waits on `world` consecutive counters: exactly the writers' ones
                assert idx == list(range(g0._landed_idx(k, 0), g0._landed_idx(k, 0) + world))
                for i in idx:
                    assert last.get((root, i), 0) < value     # written by one (rank, context) stream, in its order
                    last[(root, i)] = value
