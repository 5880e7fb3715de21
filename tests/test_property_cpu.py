"""CPU: property tests (hypothesis) over the generator's parameter space and over raw values.

  * the oracle port equals the unmodified reference on replayed fleets of arbitrary shape, clean
    and damaged (skipped where /root/reference is absent: the GPU box);
  * the device float formatter (host build of the same header) equals Python's repr on arbitrary
    used / total pairs, and the row scanners agree with csv / re on arbitrary printable rows.
"""
import csv
import io
import re

import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

import hostcheck as hc
from kgm import replay
from oracle import poll_oracle, ref_harness

COMMON = dict(deadline=None, suppress_health_check=list(HealthCheck))


@settings(max_examples=40, **COMMON)
@given(seed=st.integers(0, 2 ** 31), gpus=st.integers(1, 9), containers=st.integers(0, 7), procs=st.integers(0, 5),
       idle_every=st.integers(0, 4), ungated_every=st.integers(0, 4), tick=st.integers(0, 10 ** 6),
       damage=st.lists(st.tuples(st.integers(0, 10 ** 6), st.sampled_from("\n,\r\" 7x")), max_size=3))
def test_oracle_equals_reference_on_arbitrary_fleets(seed, gpus, containers, procs, idle_every, ungated_every, tick,
                                                     damage):
    if not ref_harness.available():
        pytest.skip("reference script not present")
    snap = replay.make_node(seed, 0, gpus=gpus, containers=containers, procs_per_container=procs,
                            idle_every=idle_every, ungated_every=ungated_every)
    g, a = replay.make_texts(seed, 0, snap, tick)
    for pos, ch in damage:                               # a few bytes replaced anywhere in either text
        if pos % 2:
            g = g[:pos % max(len(g), 1)] + ch + g[pos % max(len(g), 1) + 1:]
        else:
            a = a[:pos % max(len(a), 1)] + ch + a[pos % max(len(a), 1) + 1:]
    args = (snap.name, g, a, snap.ps_text, snap.stats_text, snap.top, snap.inspect, snap.smi_list)
    assert poll_oracle.poll(*args) == ref_harness.run_reference(*args)


@settings(max_examples=400, **COMMON)
@given(used=st.integers(0, 10 ** 18), total=st.integers(1, 10 ** 18))
def test_mem_util_repr_is_pythons(used, total):
    assert hc.mem_util_repr(used, total) == repr(used * 100.0 / total)


ROW_TEXT = st.text(alphabet=st.sampled_from(list("0123456789abcxyzGPU-/._ ,%[]MiB\t")), max_size=120)


@settings(max_examples=300, **COMMON)
@given(row=ROW_TEXT)
def test_row_scanners_agree_with_csv_and_re_on_arbitrary_rows(row):
    cells = next(csv.reader(io.StringIO(row)), [])
    got = hc.scan_gpu_row(row.encode())
    assert got["ncells"] == min(len(cells), 255)
    for slot, col in enumerate((1, 3, 4, 5)):
        if col < len(cells):
            assert got["nd"][slot] == min(len(re.sub(r"\D", "", cells[col])), 255)
    got = hc.scan_proc_row(row.encode())
    assert got["ncells"] == min(len(cells), 255)
    if len(cells) >= 2:
        assert got["name_len"] == len(cells[1].replace(" ", ""))
    fast = hc.fast_gpu_row(row.encode())
    if fast is not None:                                 # the fast parser may decline, never disagree
        assert len(cells) >= 6 and fast["key"].decode() == cells[0].lstrip()
        assert fast["mu"].decode() == re.sub(r"\D", "", cells[4])
