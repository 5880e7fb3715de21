"""GPU: the device-side replay generator writes the same bytes as kgm/replay.py make_texts, and a
window generated on the device encodes to the oracle's lines without its texts visiting the host."""
import numpy as np
import pytest

from kgm import monitor, replay, replay_device
from fleet_util import entry_of_snap, oracle_of_replay_doc
from oracle import poll_oracle

pytestmark = pytest.mark.gpu


def _nodes(seed):
    nodes, ids = [], [3, 17, 12499, 0, 256]
    for k, n in enumerate(ids):
        snap = replay.make_node(seed, n, gpus=8 if k != 3 else 3, containers=8 + 3 * k, procs_per_container=1 + k % 4,
                                idle_every=3 if k == 1 else 0, ungated_every=4 if k == 2 else 0)
        nodes.append(snap)
    nodes[1].mig = 1                                  # MIG parents: "[N/A]" utilization
    nodes[4].mig = 2                                  # ... and MIG uuids in the process rows
    return nodes, ids


@pytest.mark.parametrize("tick0,n_ticks", [(0, 1), (0, 5), (3599, 3), (1000000, 2)])
def test_device_texts_are_make_texts_byte_for_byte(tick0, n_ticks):
    seed = 7
    nodes, ids = _nodes(seed)
    gen = replay_device.DeviceReplay(nodes, ids, seed)
    win = gen.window(tick0, n_ticks)
    assert win.n_docs == len(nodes) * n_ticks
    d = 0
    for k, snap in enumerate(nodes):
        for t in range(tick0, tick0 + n_ticks):
            g, a = replay.make_texts(seed, ids[k], snap, t)
            dg, da = win.texts(d)
            assert dg == g, (k, t)
            assert da == a, (k, t)
            assert win.doc_node[d] == k and win.doc_tick[d] == t
            d += 1
    assert win.gpu_bytes == sum(len(replay.make_texts(seed, ids[k], s, t)[0])
                                for k, s in enumerate(nodes) for t in range(tick0, tick0 + n_ticks))


def test_a_window_generated_on_the_device_encodes_to_the_oracle_s_lines():
    seed = 11
    ids = list(range(40, 52))
    nodes = [replay.make_node(seed, n, gpus=8, containers=8, procs_per_container=4) for n in ids]
    gen = replay_device.DeviceReplay(nodes, ids, seed)
    win = gen.window(8, 4)
    with monitor.Monitor(0) as mon:
        mon.load_fleet([entry_of_snap(s) for s in nodes])
        win.upload(mon)
        mon.encode_resident()
        res = mon.fetch()
        body = b""
        d = 0
        for k, snap in enumerate(nodes):
            for t in range(8, 12):
                g, a = replay.make_texts(seed, ids[k], snap, t)
                lines, err = poll_oracle.poll(snap.name, g, a, snap.ps_text, snap.stats_text, snap.top, snap.inspect,
                                              snap.smi_list)
                assert err is None and res.lines(d) == lines, (k, t)
                body += poll_oracle.body_of(lines)
                d += 1
        assert res.body == body
