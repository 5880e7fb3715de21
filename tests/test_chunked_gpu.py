"""GPU: the chunked end-to-end path of kgm_encode_batch (large host batches: text copies,
kernels and result copies of consecutive chunks overlap, every kernel continues the line and
byte count of its predecessors, the host learns the sizes from totals the kernels post into
mapped memory) gives the bytes of the one-launch path and of the oracle.  Batches mix healthy
documents with damaged ones (documents that die early, emit nothing or take the generic
path inside the kernel)."""
import random

import numpy as np
import pytest

from fleet_util import UNSUPPORTED, entry_of_snap
from kgm import monitor, replay
from oracle import poll_oracle
from test_parity_gpu import _damage

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mon():
    m = monitor.Monitor(0)
    yield m
    m.close()


def _big_batch(seed, damaged_every):
    rng = random.Random(seed)
    rp = replay.make_replay(seed=seed, nodes=64, ticks=30, gpus=8, containers=10, procs_per_container=4)
    gt, at = list(rp.gpu_texts), list(rp.apps_texts)
    if damaged_every:
        for d in range(rp.n_docs):
            if d % damaged_every == 0:
                gt[d] = _damage(rng, gt[d])
                at[d] = _damage(rng, at[d])
        gt[5] = ""                                   # no rows at all
        at[7] = ""
        gt[rp.n_docs - 1] = _damage(rng, gt[rp.n_docs - 1])      # the last document of the last chunk
    return rp, gt, at


def _same(a, b):
    assert a.n_docs == b.n_docs and a.n_lines == b.n_lines and a.n_bytes == b.n_bytes
    assert a.body == b.body
    assert np.array_equal(a.line_off, b.line_off)
    assert np.array_equal(a.doc_line_off, b.doc_line_off)
    assert np.array_equal(a.status, b.status)


@pytest.mark.parametrize("damaged_every", [0, 3])
@pytest.mark.parametrize("pinned", [True, False])
def test_chunked_and_one_launch_results_are_identical(mon, damaged_every, pinned):
    rp, gt, at = _big_batch(11 + damaged_every, damaged_every)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    batch = mon.make_batch(gt, at, rp.doc_node, rp.doc_tick, pinned=pinned)
    assert batch.gpu_bytes + batch.apps_bytes >= (4 << 20) and rp.n_docs >= 1024

    mon.upload(batch)
    mon.encode_resident()
    one = mon.fetch()
    assert mon.timing()["launches"] == 1

    for _ in range(3):                               # (posted totals and counter sets are reused call after call)
        chunked = mon.encode(batch)
        t = mon.timing()
        assert t["launches"] > 1 and t["fused"] == 1
        _same(chunked, one)
    mon.encode_resident()                            # the batch is resident afterwards
    _same(mon.fetch(), one)

    # and the bytes are the reference's
    rng = random.Random(1)
    docs = range(rp.n_docs) if damaged_every and pinned else rng.sample(range(rp.n_docs), 100)
    raised = 0
    for d in docs:
        if int(chunked.status[d]) in UNSUPPORTED:
            continue
        s = rp.nodes[int(rp.doc_node[d])]
        want_lines, want_err = poll_oracle.poll(s.name, gt[d], at[d], s.ps_text, s.stats_text, s.top,
                                                s.inspect, s.smi_list)
        assert chunked.lines(d) == want_lines, d
        assert chunked.error(d) == want_err, d
        raised += want_err is not None
    if damaged_every and pinned:
        assert raised > 100


def test_chunked_first_call_on_a_fresh_context():
    """A fresh context has no idea of the output size: the first chunked call may find its
    guesses too small, must fall back to the plain path and still return the right bytes."""
    rp, gt, at = _big_batch(31, 5)
    with monitor.Monitor(0) as fresh, monitor.Monitor(0) as ref:
        for m in (fresh, ref):
            m.load_fleet([entry_of_snap(s) for s in rp.nodes])
        got = fresh.encode(fresh.make_batch(gt, at, rp.doc_node, rp.doc_tick, pinned=True))
        ref.upload(ref.make_batch(gt, at, rp.doc_node, rp.doc_tick))
        ref.encode_resident()
        _same(got, ref.fetch())
