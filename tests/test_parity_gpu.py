"""GPU: the CUDA path, called through the C ABI, against the oracle and the frozen
reference outputs.  Bit-exact: every byte of every line, the line order, the per-document
line counts and the class of the exception the reference would have died with."""
import random
import zlib

import numpy as np
import pytest

import cases
import golden_io
from fleet_util import (UNSUPPORTED, entry_of_snap, oracle_of_case, oracle_of_replay_doc,
                        run_cases_as_one_batch, run_replay)
from kgm import _ffi, monitor, replay
from oracle import poll_oracle

pytestmark = pytest.mark.gpu
GOLD = golden_io.load()

# cases on which the library reports "unsupported" instead of the reference's exception class
# (documented divergence: DESIGN.md "Limits"); the emitted lines must still agree.
UNSUPPORTED_OK = {"err_quoted_first_cell": _ffi.DOC_UNSUPPORTED_QUOTE,     # quotes that never close
                  "mem_20_digits": _ffi.DOC_UNSUPPORTED_RANGE}             # memory.used of 10^19 MiB and more


PIPELINES = {"fused": _ffi.PIPELINE_AUTO, "staged": _ffi.PIPELINE_STAGED,
             "fused-generic": _ffi.PIPELINE_FUSED_GENERIC, "fused-v2": _ffi.PIPELINE_FUSED_V2}


@pytest.fixture(scope="module", params=list(PIPELINES))
def mon(request):
    """Every test runs four times: through the fused single-pass kernel the library picks (the
    team kernel k_poll for batches of this size; fast path, with its in-kernel generic path for
    irregular documents), through the four staged kernels, through the fused kernel with the fast
    path switched off, and through the warp kernel k_poll2."""
    m = monitor.Monitor(0)
    m.set_pipeline(PIPELINES[request.param])
    m.pipeline_name = request.param
    yield m
    m.close()


def _check_unsupported(got_lines, status, want_lines, gpu_csv, apps_csv):
    """A document outside the supported dialect: the status names the reason, the reason is
    really in the text, and the lines emitted up to there are the reference's."""
    assert got_lines == want_lines[:len(got_lines)]
    if status == _ffi.DOC_UNSUPPORTED_QUOTE:
        assert '"' in gpu_csv + apps_csv
    elif status == _ffi.DOC_UNSUPPORTED_BYTE:
        assert not (gpu_csv + apps_csv).isascii()
    else:
        assert status == _ffi.DOC_UNSUPPORTED_RANGE
        import re
        assert any(len(re.sub(r"\D", "", cell)) > 19 for row in gpu_csv.split("\n") for cell in row.split(",")) or \
            max(map(len, (gpu_csv + apps_csv).split("\n"))) > 8000


@pytest.mark.parametrize("g", GOLD[:64], ids=[g["name"] for g in GOLD[:64]])
def test_single_poll_matches_frozen_reference(mon, g):
    lines, err = monitor.poll(mon, *cases.inputs_of(g))
    if g["name"] in UNSUPPORTED_OK:
        assert err == "kgm:%d" % UNSUPPORTED_OK[g["name"]]
        _check_unsupported(lines, UNSUPPORTED_OK[g["name"]], g["lines"], g["gpu_csv"], g["apps_csv"])
    else:
        assert lines == g["lines"]
        assert err == g["err"]


def test_all_golden_cases_in_one_batch(mon):
    res = run_cases_as_one_batch(mon, GOLD)
    assert res.n_docs == len(GOLD)
    want_body = b""
    for d, g in enumerate(GOLD):
        if g["name"] in UNSUPPORTED_OK:
            assert int(res.status[d]) == UNSUPPORTED_OK[g["name"]]
            _check_unsupported(res.lines(d), int(res.status[d]), g["lines"], g["gpu_csv"], g["apps_csv"])
        else:
            assert res.lines(d) == g["lines"], g["name"]
            assert res.error(d) == g["err"], g["name"]
        want_body += poll_oracle.body_of(res.lines(d))
    assert res.body == want_body
    assert int(res.doc_line_off[-1]) == res.n_lines == sum(len(res.lines(d)) for d in range(len(GOLD)))
    assert sum(len(g["lines"]) for g in GOLD) - res.n_lines == sum(
        len(g["lines"]) - len(res.lines(d)) for d, g in enumerate(GOLD) if g["name"] in UNSUPPORTED_OK)
    assert int(res.line_off[-1]) == res.n_bytes == len(want_body)


def test_empty_and_ragged_batches(mon):
    rp = replay.make_replay(seed=3, nodes=2, ticks=2, gpus=2, containers=2, procs_per_container=1)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    res = mon.encode([], [], [])
    assert (res.n_docs, res.n_lines, res.n_bytes, res.body) == (0, 0, 0, b"")
    texts_g = ["", rp.gpu_texts[0], "only a header", rp.gpu_texts[3], "\n", ""]
    texts_a = ["", rp.apps_texts[0], "", "h\n", rp.apps_texts[2], "x"]
    nodes = [0, 0, 1, 1, 1, 0]
    res = mon.encode(texts_g, texts_a, nodes)
    for d in range(len(nodes)):
        s = rp.nodes[nodes[d]]
        want = poll_oracle.poll(s.name, texts_g[d], texts_a[d], s.ps_text, s.stats_text, s.top,
                                s.inspect, s.smi_list)
        assert (res.lines(d), res.error(d)) == want, d


@pytest.mark.parametrize("shape", [
    dict(nodes=1, ticks=1, gpus=8, containers=8, procs_per_container=4),                    # BASELINE configs[0]
    dict(nodes=6, ticks=5, gpus=8, containers=32, procs_per_container=1),                   # configs[1] shape
    dict(nodes=3, ticks=4, gpus=8, containers=10, procs_per_container=4),                   # configs[2] shape
    dict(nodes=4, ticks=3, gpus=8, containers=12, procs_per_container=2, idle_every=3, ungated_every=4),
    dict(nodes=2, ticks=2, gpus=16, containers=70, procs_per_container=1, idle_every=7),
    dict(nodes=2, ticks=2, gpus=40, containers=3, procs_per_container=45),                  # > 32 rows per warp chunk
    dict(nodes=3, ticks=2, gpus=1, containers=1, procs_per_container=1),
], ids=["config1", "config2-shape", "config3-shape", "idle-ungated", "wide", "tall", "tiny"])
def test_replay_batches_match_oracle(mon, shape):
    rp = replay.make_replay(seed=11, **shape)
    res = run_replay(mon, rp)
    body = b""
    for d in range(rp.n_docs):
        lines, err = oracle_of_replay_doc(rp, d)
        assert err is None and int(res.status[d]) == 0
        assert res.lines(d) == lines, d
        body += poll_oracle.body_of(lines)
    assert res.body == body


def _damage(rng, text):
    for _ in range(rng.randrange(4)):
        if not text:
            break
        i = rng.randrange(len(text))
        k = rng.randrange(7)
        if k == 0:
            text = text[:i] + text[i + 1:]
        elif k == 1:
            text = text[:i] + rng.choice(",\n %9a0[\r") + text[i:]
        elif k == 2:
            text = text[:i] + text[min(len(text), i + rng.randrange(1, 40)):]
        elif k == 3:
            text = text[:i] + "\n" + text[i:]
        elif k == 4:
            text = text[:i] + rng.choice("0123456789") + text[i + 1:]
        elif k == 5:
            text = text[:i] + rng.choice(text.split("\n")) + "\n" + text[i:]
        else:
            text = text[:i] + " " + text[i:]
    return text


def test_damaged_inputs_match_oracle(mon):
    """Malformed rows: same lines before the raise point, same exception class."""
    rng = random.Random(99)
    rp = replay.make_replay(seed=5, nodes=3, ticks=1, gpus=4, containers=6, procs_per_container=2,
                            idle_every=3, ungated_every=5)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    gt, at, nodes = [], [], []
    for _ in range(600):
        d = rng.randrange(rp.n_docs)
        gt.append(_damage(rng, rp.gpu_texts[d]))
        at.append(_damage(rng, rp.apps_texts[d]))
        nodes.append(int(rp.doc_node[d]))
    res = mon.encode(gt, at, nodes)
    raised = unsupported = 0
    for d in range(len(gt)):
        s = rp.nodes[nodes[d]]
        want_lines, want_err = poll_oracle.poll(s.name, gt[d], at[d], s.ps_text, s.stats_text, s.top,
                                                s.inspect, s.smi_list)
        if int(res.status[d]) in UNSUPPORTED:
            unsupported += 1
            _check_unsupported(res.lines(d), int(res.status[d]), want_lines, gt[d], at[d])
            continue
        assert res.lines(d) == want_lines, (d, gt[d], at[d])
        assert res.error(d) == want_err, (d, gt[d], at[d])
        raised += want_err is not None
    assert 50 < raised < 560 and unsupported < 30


def test_docker_side_errors_match_oracle(mon):
    """Raise points inside the container loop, many containers, in one batch with good ones."""
    sel = [g for g in GOLD if g["name"].startswith(("err_labels", "err_pod", "ok_pod", "err_no_", "err_bad",
                                                    "err_stats", "gates", "pid_twice", "dup_keys"))]
    res = run_cases_as_one_batch(mon, sel * 3)
    for d, g in enumerate(sel * 3):
        assert (res.lines(d), res.error(d)) == (g["lines"], g["err"]), g["name"]


def test_full_config2_properties_and_sampled_parity(mon):
    """BASELINE configs[1] at full size: 64 nodes x 60 ticks = 3840 documents."""
    rp = replay.make_replay(seed=0, **replay.CONFIGS["config2"])
    assert rp.n_docs == 3840 and rp.n_samples() == 30720
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
    res = mon.encode(batch)
    assert not res.status.any()
    # structure: offsets are a partition of the body into '\n'-terminated lines
    lo = res.line_off
    assert lo[0] == 0 and int(lo[-1]) == res.n_bytes == len(res.body)
    assert np.all(np.diff(lo.astype(np.int64)) > 0)
    body = np.frombuffer(res.body, np.uint8)
    assert np.all(body[lo[1:].astype(np.int64) - 1] == 10)
    assert int((body == 10).sum()) == res.n_lines
    # 8 node lines + 32 pod lines per document, node lines first
    dl = np.diff(res.doc_line_off.astype(np.int64))
    assert np.all(dl == 40)
    starts = lo[:-1].astype(np.int64)
    first5 = np.stack([body[starts + i] for i in range(5)], 1)
    is_node = (first5 == np.frombuffer(b"node_", np.uint8)).all(1)
    is_pod = (first5 == np.frombuffer(b"pod_g", np.uint8)).all(1)
    assert np.all(is_node | is_pod) and int(is_node.sum()) == 30720
    assert np.all(is_node.reshape(-1, 40)[:, :8]) and not is_node.reshape(-1, 40)[:, 8:].any()
    # sampled byte parity against the oracle + checksum of the sampled documents
    rng = random.Random(1)
    for d in sorted(rng.sample(range(rp.n_docs), 160)):
        lines, err = oracle_of_replay_doc(rp, d)
        a, b = int(res.doc_line_off[d]), int(res.doc_line_off[d + 1])
        got = res.body[int(lo[a]):int(lo[b])]
        assert err is None and zlib.crc32(got) == zlib.crc32(poll_oracle.body_of(lines)), d
    # idempotence, and the resident (device-pointer) path gives the same bytes
    mon.upload(batch)
    mon.encode_resident()
    again = mon.fetch()
    assert again.body == res.body and np.array_equal(again.line_off, res.line_off)
    mon.encode_resident()
    assert mon.fetch().body == res.body
    t = mon.timing()
    assert t["total_ms"] > 0
    if mon.pipeline_name == "staged":
        assert t["launches"] == 4 and t["fused"] == 0
    else:
        # the clean fleet batch goes through the fast path entirely (or not at all when it is off)
        assert t["launches"] == 1 and t["fused"] == 1
        assert t["slow_docs"] == (0 if mon.pipeline_name in ("fused", "fused-v2") else rp.n_docs)
        assert t["fused_ctas_per_sm"] in (1, 2)                     # 1 x 7 or 2 x 3/4 teams per SM


def test_mem_util_floats_parse_back_within_tolerance(mon):
    """North-star tolerance for the only floating-point field: 1e-6 relative after parse-back
    (the bytes are in fact identical; this is the looser, representation-free check)."""
    rp = replay.make_replay(seed=21, nodes=4, ticks=6, gpus=8, containers=4, procs_per_container=2)
    res = run_replay(mon, rp)
    n = 0
    for d in range(rp.n_docs):
        want, _ = oracle_of_replay_doc(rp, d)
        for a, b in zip(res.lines(d), want):
            if a.startswith("node_gpu"):
                x = float(a.split("mem_util=")[1].split(",")[0])
                y = float(b.split("mem_util=")[1].split(",")[0])
                assert abs(x - y) <= 1e-6 * max(abs(y), 1e-300)
                n += 1
    assert n == rp.n_samples()


def test_missing_library_or_bad_calls_fail_loudly(mon):
    lib = _ffi.load()
    import ctypes as C
    res = _ffi.KgmResult()
    fresh = C.c_void_p()
    assert lib.kgm_create(C.byref(fresh), 0) == 0
    assert lib.kgm_encode_resident(fresh, C.byref(res)) == _ffi.E_STATE       # nothing loaded
    lib.kgm_destroy(fresh)
    bad = C.c_void_p()
    assert lib.kgm_create(C.byref(bad), 4096) == _ffi.E_CUDA                  # no such device
    rp = replay.make_replay(seed=1, nodes=1, ticks=1)
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    with pytest.raises(_ffi.KgmError):
        mon.encode(rp.gpu_texts, rp.apps_texts, [7])                          # unknown node id
