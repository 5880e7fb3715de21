"""CPU: the oracle restatement against (a) the frozen reference outputs and (b), where the
reference tree exists (build container only), the live unmodified script."""
import random

import pytest

import cases
import golden_io
from oracle import poll_oracle, ref_harness

GOLD = golden_io.load()


def test_golden_is_current():
    """The fixture file was generated from exactly today's case list."""
    assert [g["name"] for g in GOLD] == [c["name"] for c in cases.all_cases()]
    for g, c in zip(GOLD, cases.all_cases()):
        assert cases.inputs_of(g) == cases.inputs_of(c), g["name"]


@pytest.mark.parametrize("g", GOLD, ids=[g["name"] for g in GOLD])
def test_oracle_matches_frozen_reference(g):
    lines, err = poll_oracle.poll(*cases.inputs_of(g))
    assert lines == g["lines"]
    assert err == g["err"]


def test_appendix_b_known_answers():
    g = {x["name"]: x for x in GOLD}
    b = g["appendix_b"]["lines"]
    assert b[0] == ('node_gpu,host=node-a,gpu_id=0 gpu_util=49,mem_util=96.06435643564356,'
                    'mem_used=15524,mem_total=16160,node="node-a",gid=0,mem_free=636')
    assert b[3].startswith("pod_gpu,host=node-a,namespace=10007089-name,gpu_id=0,"
                           "containerid=16ee8a182726,pid_t=71169,gpu_util_t=49,"
                           'user_name_t=alice-bob procressname="python3train.py",')
    assert b[5].count("None") == 2 and len(b) == 6
    assert g["node_name_unset"]["lines"][0].startswith("node_gpu,host=None,gpu_id=0 ")
    reprs = [l.split("mem_util=")[1].split(",")[0] for l in g["mem_util_reprs"]["lines"]]
    assert reprs[:9] == ["0.0005453781925075944", "100.0", "33.333333333333336",
                         "77.77777777777777", "74.609375", "9.99999000001e-05",
                         "76.39232673267327", "66.66666666666667", "0.015869140625"]


needs_ref = pytest.mark.skipif(not ref_harness.available(),
                               reason="reference tree only exists in the build container")


def _mutate(rng, text):
    """Random byte-level damage: the kinds of malformed rows SURVEY.md A.6 lists."""
    if not text:
        return text
    ops = rng.randrange(4)
    for _ in range(ops):
        i = rng.randrange(len(text))
        kind = rng.randrange(7)
        if kind == 0:
            text = text[:i] + text[i + 1:]
        elif kind == 1:
            text = text[:i] + rng.choice(",\n %9a0[") + text[i:]
        elif kind == 2:
            j = min(len(text), i + rng.randrange(1, 40))
            text = text[:i] + text[j:]
        elif kind == 3:
            text = text[:i] + "\n" + text[i:]
        elif kind == 4:
            text = text[:i] + rng.choice("0123456789") + text[i + 1:]
        elif kind == 5:
            line = rng.choice(text.split("\n"))
            text = text[:i] + line + "\n" + text[i:]
        else:
            text = text[:i] + " " + text[i:]
    return text


@needs_ref
def test_oracle_matches_live_reference_on_damaged_inputs():
    rng = random.Random(1234)
    base = [g for g in GOLD if g["name"].startswith(("replay_cfg1", "replay_idle", "appendix"))]
    checked = raised = 0
    for trial in range(400):
        g = dict(rng.choice(base))
        g["gpu_csv"] = _mutate(rng, g["gpu_csv"])
        g["apps_csv"] = _mutate(rng, g["apps_csv"])
        if rng.randrange(4) == 0:
            cid = rng.choice(sorted(g["top"]))
            g["top"] = dict(g["top"])
            g["top"][cid] = _mutate(rng, g["top"][cid])
        want = ref_harness.run_reference(*cases.inputs_of(g))
        got = poll_oracle.poll(*cases.inputs_of(g))
        assert got == want, (trial, g["gpu_csv"], g["apps_csv"])
        checked += 1
        raised += want[1] is not None
    assert checked == 400 and 20 < raised < 380
