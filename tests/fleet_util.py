"""Helpers shared by the GPU parity tests: run cases / replays through the library and
through the oracle in the same shape."""
from kgm import _ffi
from kgm.fleet import node_entry
from oracle import poll_oracle

import cases

UNSUPPORTED = {_ffi.DOC_UNSUPPORTED_QUOTE, _ffi.DOC_UNSUPPORTED_BYTE, _ffi.DOC_UNSUPPORTED_RANGE}


def entry_of_case(c):
    return node_entry(c["node_name"], c["ps_text"], c["stats_text"], c["top"], c["inspect"], c["smi_list"])


def entry_of_snap(snap):
    return node_entry(snap.name, snap.ps_text, snap.stats_text, snap.top, snap.inspect, snap.smi_list)


def oracle_of_case(c):
    return poll_oracle.poll(*cases.inputs_of(c))


def oracle_of_replay_doc(rp, d):
    snap = rp.nodes[int(rp.doc_node[d])]
    return poll_oracle.poll(snap.name, rp.gpu_texts[d], rp.apps_texts[d], snap.ps_text, snap.stats_text,
                            snap.top, snap.inspect, snap.smi_list)


def run_cases_as_one_batch(mon, case_list):
    """Every case becomes its own node + one document; one library call for all."""
    mon.load_fleet([entry_of_case(c) for c in case_list])
    return mon.encode([c["gpu_csv"] for c in case_list], [c["apps_csv"] for c in case_list],
                      list(range(len(case_list))))


def run_replay(mon, rp):
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    return mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
