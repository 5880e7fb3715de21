"""GPU: BASELINE configs[2] and configs[4] shapes.

configs[2] (1k-node fleet, 10-minute window) is 600k documents / 2.7 GB of text in all; it
is streamed as batches of ticks.  One such batch is built here at full width (1000 nodes)
from 4 distinct ticks replicated 10 times, so that the host-side generation stays in
seconds; it is checked through size-independent properties (the offsets partition the body
into newline-terminated lines, every document has its 8 node + 40 pod lines, replicas of a
document are byte-identical wherever they land in the output) plus sampled byte parity
against the oracle.  configs[4] (short-lived pods, MIG, skewed keys) runs in miniature with
full byte parity: Zipf-sized hosts, several docker snapshots per host, gpu_id=None lines
and documents on which the reference dies after the node lines."""
import random
import zlib

import numpy as np
import pytest

from fleet_util import entry_of_snap, oracle_of_replay_doc
from kgm import _ffi, monitor, replay
from oracle import poll_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["fused", "staged", "fused-v1", "fused-v2"])
def mon(request):
    m = monitor.Monitor(0)
    m.set_pipeline({"fused": _ffi.PIPELINE_AUTO, "staged": _ffi.PIPELINE_STAGED,
                    "fused-v1": _ffi.PIPELINE_FUSED_V1, "fused-v2": _ffi.PIPELINE_FUSED_V2}[request.param])
    m.pipeline_name = request.param
    yield m
    m.close()


def test_config3_shape_one_streamed_batch_at_full_width(mon):
    distinct, replicas = 4, 10
    rp = replay.make_replay(seed=3, nodes=1000, ticks=distinct, gpus=8, containers=10, procs_per_container=4)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    # batch order: tick-major (all nodes at tick 0, then tick 1, ...), each distinct tick 10 times
    order = [n * distinct + (t % distinct) for t in range(distinct * replicas) for n in range(1000)]
    gt = [rp.gpu_texts[d] for d in order]
    at = [rp.apps_texts[d] for d in order]
    dn = [int(rp.doc_node[d]) for d in order]
    res = mon.encode(mon.make_batch(gt, at, dn, pinned=True))
    n_docs = len(order)
    assert res.n_docs == n_docs == 40000 and not res.status.any()
    lo = res.line_off.astype(np.int64)
    body = np.frombuffer(res.body, np.uint8)
    assert lo[0] == 0 and lo[-1] == res.n_bytes == len(res.body)
    assert np.all(np.diff(lo) > 0) and np.all(body[lo[1:] - 1] == 10)
    assert int((body == 10).sum()) == res.n_lines == n_docs * 48
    assert np.all(np.diff(res.doc_line_off.astype(np.int64)) == 48)
    # replicas of a document produce the same bytes
    dstart = lo[res.doc_line_off.astype(np.int64)]
    crc = {}
    rng = random.Random(2)
    for i in rng.sample(range(n_docs), 3000):
        c = zlib.crc32(res.body[dstart[i]:dstart[i + 1]])
        assert crc.setdefault(order[i], c) == c, i
    # sampled byte parity with the oracle
    for i in rng.sample(range(n_docs), 60):
        lines, err = oracle_of_replay_doc(rp, order[i])
        assert err is None and res.body[dstart[i]:dstart[i + 1]] == poll_oracle.body_of(lines), i
    if mon.pipeline_name == "fused":
        t = mon.timing()
        assert t["fused"] == 1 and t["slow_docs"] == 0
    # one launch over the resident batch (the end-to-end call above cut it into chunks, the small ones
    # of which AUTO gives to the team kernel): 40 000 documents -> the warp kernel
    mon.encode_resident()
    assert mon.last_kernel() == {"fused": "k_poll2", "fused-v1": "k_poll", "fused-v2": "k_poll2",
                                 "staged": "k_tokenize, k_join, k_scan, k_encode"}[mon.pipeline_name]


def test_config5_shape_skewed_short_lived_pods_and_mig(mon):
    rp = replay.make_skewed_replay(seed=5, hosts=40, epochs=3, ticks_per_epoch=4)
    assert max(len(s.containers) for s in rp.nodes) > 100          # a few hosts carry most pods
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    res = mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
    body, died, none_lines, empty_util = b"", 0, 0, 0
    for d in range(rp.n_docs):
        lines, err = oracle_of_replay_doc(rp, d)
        assert res.lines(d) == lines, d
        assert res.error(d) == err, d
        died += err is not None
        none_lines += sum("gpu_id=None" in x for x in lines)
        empty_util += sum(" gpu_util=," in x for x in lines)
        body += poll_oracle.body_of(lines)
    assert res.body == body
    assert died > 20 and none_lines > 100 and empty_util > 100
    if mon.pipeline_name == "fused":
        t = mon.timing()
        # the dying documents go generic, and the few whose texts exceed a warp's slice
        assert t["fused"] == 1 and died <= t["slow_docs"] < rp.n_docs // 3
        mon.encode_resident()
        assert mon.last_kernel() == "k_poll"         # 480 documents, less than one wave of k_poll2: the team kernel
    if mon.pipeline_name == "fused-v2":
        # k_poll2 answers the KeyError of gpu.py:151 itself (node lines, then the status): only the
        # documents whose texts exceed a warp's slice go generic
        t = mon.timing()
        assert t["fused"] == 1 and t["slow_docs"] < died
