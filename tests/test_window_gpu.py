"""GPU: stage 3, the per-series time-window reductions (kgm_window_reduce), against
oracle/window_oracle.py.  Parity is UNPINNED by the reference (it has no such stage); the
samples being reduced are the oracle's own lines, so what is checked here is that the
device groups, sorts and reduces exactly those: integers bit-exact, the mean as the one
IEEE division sum/count (equal, tolerance 0)."""
import random

import numpy as np
import pytest

from fleet_util import entry_of_snap
from kgm import _ffi, monitor, replay
from oracle import poll_oracle, window_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mon():
    m = monitor.Monitor(0)
    yield m
    m.close()


def _oracle_groups(rp, entries, gpu_texts, apps_texts, doc_node, doc_tick, window_ticks):
    cont_base = np.concatenate([[0], np.cumsum([len(e.containers) for e in entries])])
    index_of = [{c.cid: int(cont_base[n]) + i for i, c in enumerate(e.containers)} for n, e in enumerate(entries)]
    docs = []
    for d in range(len(gpu_texts)):
        n = int(doc_node[d])
        s = rp.nodes[n]
        lines, _ = poll_oracle.poll(s.name, gpu_texts[d], apps_texts[d], s.ps_text, s.stats_text, s.top,
                                    s.inspect, s.smi_list)
        docs.append((n, int(doc_tick[d]), lines, gpu_texts[d], index_of[n]))
    return window_oracle.windows(docs, window_ticks)


def _compare(got, want):
    assert got["n"] == len(want)
    for i, (key, g) in enumerate(want.items()):
        kind, entity, gid, window = key
        assert (int(got["kind"][i]), int(got["gpu_id"][i]), int(got["window"][i])) == (kind, gid, window), key
        assert int(got["node"][i]) == g["node"], key
        assert int(got["container"][i]) == (entity if kind else 0xFFFFFFFF), key
        assert int(got["count"][i]) == g["count"], key
        for m in window_oracle.METRICS:
            a = g[m]
            assert int(got[m + "_count"][i]) == a["count"], (key, m)
            if a["count"]:
                assert (int(got[m + "_min"][i]), int(got[m + "_max"][i]), int(got[m + "_sum"][i]),
                        int(got[m + "_last"][i])) == (a["min"], a["max"], a["sum"], a["last"]), (key, m)
                assert float(got[m + "_mean"][i]) == a["mean"], (key, m)


@pytest.mark.parametrize("window_ticks,power", [(5, True), (1, False), (60, True), (7, True)])
def test_windows_of_a_clean_fleet_match_the_oracle(mon, window_ticks, power):
    rp = replay.make_replay(seed=41, nodes=5, ticks=23, gpus=4, containers=7, procs_per_container=2,
                            idle_every=3, ungated_every=4, power=power)
    entries = [entry_of_snap(s) for s in rp.nodes]
    mon.load_fleet(entries)
    res = mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
    assert not res.status.any()
    got = mon.window_reduce(window_ticks)
    assert mon.window_front() == "k_poll2_samples"   # a clean fleet: the planning half of k_poll2 feeds stage 3
    want = _oracle_groups(rp, entries, rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, window_ticks)
    _compare(got, want)
    # the staged front end (tokenizer + join) gives the same groups
    import os
    os.environ["KGM_WINDOW_STAGED_FRONT"] = "1"
    try:
        staged = mon.window_reduce(window_ticks)
        assert mon.window_front() == "k_tokenize, k_join"
    finally:
        del os.environ["KGM_WINDOW_STAGED_FRONT"]
    _compare(staged, want)
    got = mon.window_reduce(window_ticks)
    # ... and in line form (node_gpu_w / pod_gpu_w with nanosecond timestamps)
    from kgm import sink
    hosts = [s.name for s in rp.nodes]
    cids = [c.cid for e in entries for c in e.containers]
    assert sink.window_lines(got, hosts, cids, window_ticks, t0_ns=1_700_000_000 * 10 ** 9) == \
        window_oracle.lines(want, hosts, cids, window_ticks, t0_ns=1_700_000_000 * 10 ** 9)
    if power:
        assert int(got["power_count"].sum()) > 0
    else:
        assert int(got["power_count"].sum()) == 0
    # the line protocol itself is untouched by stage 3
    assert mon.fetch().body == res.body


def test_windows_with_damaged_shuffled_and_repeated_documents(mon):
    """Documents out of tick order, the same (node, tick) twice, rows the reference dies on
    (only the lines POSTed before the raise are samples), digit-free and huge cells."""
    rng = random.Random(7)
    rp = replay.make_replay(seed=43, nodes=3, ticks=9, gpus=3, containers=5, procs_per_container=2,
                            idle_every=2, power=True)
    entries = [entry_of_snap(s) for s in rp.nodes]
    order = list(range(rp.n_docs)) + [rng.randrange(rp.n_docs) for _ in range(10)]
    rng.shuffle(order)
    gt, at, dn, dt = [], [], [], []
    for d in order:
        g, a = rp.gpu_texts[d], rp.apps_texts[d]
        k = rng.randrange(8)
        if k == 0:
            g = g.replace(" %,", " [N/A],", 1)                       # gpu_util without digits
        elif k == 1:
            g = g.replace(" W\n", " [N/A]\n", 1)                     # power without digits
        elif k == 2:
            rows = g.split("\n")
            rows[2] = rows[2].replace("MiB", "", 1).replace("183359", "0")   # ZeroDivisionError mid-way
            g = "\n".join(rows)
        elif k == 3:
            a = a.replace(" MiB\n", " 99999999999 MiB\n", 1)         # saturates at 2^32-2
        elif k == 4:
            a = a + "77, x, GPU-unknown, 1 MiB\n"                    # KeyError gpu.py:151: node lines only
        gt.append(g); at.append(a); dn.append(int(rp.doc_node[d])); dt.append(int(rp.doc_tick[d]))
    mon.load_fleet(entries)
    res = mon.encode(gt, at, dn, dt)
    assert res.status.any() and not res.status.all()
    for w in (4, 2):
        _compare(mon.window_reduce(w), _oracle_groups(rp, entries, gt, at, dn, dt, w))


def test_fused_front_end_with_missing_values_and_dying_documents(mon):
    """Documents the planning half of k_poll2 takes although they are not clean: utilisation or power
    without digits, and the KeyError of an unknown process uuid (node lines only)."""
    rp = replay.make_replay(seed=44, nodes=6, ticks=9, gpus=8, containers=6, procs_per_container=3, idle_every=4,
                            power=True)
    entries = [entry_of_snap(s) for s in rp.nodes]
    rng = random.Random(9)
    gt, at = [], []
    for d in range(rp.n_docs):
        g, a = rp.gpu_texts[d], rp.apps_texts[d]
        k = rng.randrange(5)
        if k == 0:
            g = g.replace(" %,", " [N/A],", 1)
        elif k == 1:
            g = g.replace(" W\n", " [N/A]\n", 1)
        elif k == 2:
            a = a + "77, x, GPU-unknown, 1 MiB\n"
        gt.append(g); at.append(a)
    mon.load_fleet(entries)
    res = mon.encode(gt, at, rp.doc_node, rp.doc_tick)
    assert res.status.any() and not res.status.all()
    for w in (3, 9):
        got = mon.window_reduce(w)
        assert mon.window_front() == "k_poll2_samples"
        _compare(got, _oracle_groups(rp, entries, gt, at, rp.doc_node, rp.doc_tick, w))


def test_window_reduce_needs_ticks_and_a_batch(mon):
    rp = replay.make_replay(seed=2, nodes=1, ticks=2, gpus=2, containers=2, procs_per_container=1)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node)             # no doc_tick
    with pytest.raises(_ffi.KgmError):
        mon.window_reduce(5)
    mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
    with pytest.raises(_ffi.KgmError):
        mon.window_reduce(0)
    assert mon.window_reduce(1)["n"] == 2 * (2 + 2)                  # 2 ticks x (2 node + 2 pod series)
