"""GPU: kgm_inspect_labels (the `docker inspect` texts of a fleet scanned on the device) builds the
same fleet entries as kgm/fleet.py node_entry with json.loads; what the scanner does not take goes
through the host code path unchanged."""
import json
import random

import pytest

from kgm import fleet, replay

pytestmark = pytest.mark.gpu


def _inputs(seed, n_nodes, damage):
    rng = random.Random(seed)
    out = []
    for n in range(n_nodes):
        s = replay.make_node(seed, n, gpus=8, containers=6 + n % 7, procs_per_container=1 + n % 3,
                             idle_every=3 if n % 2 else 0, ungated_every=4 if n % 3 == 0 else 0)
        inspect = dict(s.inspect)
        if damage:
            for cid in list(inspect):
                k = rng.randrange(12)
                text = inspect[cid]
                if k == 0:
                    inspect[cid] = text[:-3]                                        # JSONDecodeError
                elif k == 1:
                    inspect[cid] = text.replace('"Labels": {', '"Labels": null, "x": {', 1)   # AttributeError
                elif k == 2:
                    inspect[cid] = text.replace("io.kubernetes.pod.namespace", "io.kubernetes.pod.ns")   # KeyError at emit
                elif k == 3:
                    inspect[cid] = text.replace('"mlgpu": "mlgpu"', '"mlgpu": "ml\\u0067pu"')   # an escape: the host's
                elif k == 4:
                    inspect[cid] = json.dumps(json.loads(text), indent=None)        # another layout
                elif k == 5:
                    inspect[cid] = text.replace("turing-gpu-notebook-", "nodashatall")
        out.append((s.name, s.ps_text, s.stats_text, s.top, inspect, s.smi_list))
    return out


@pytest.mark.parametrize("damage", [False, True])
def test_device_scanned_fleet_equals_the_host_s(damage):
    inputs = _inputs(23, 40, damage)
    want = [fleet.node_entry(*i) for i in inputs]
    got, taken, left = fleet.node_entries_device(inputs)
    assert got == want
    total = sum(len(n.containers) for n in want)
    assert taken + left == total
    if damage:
        assert 0 < left < total // 2 and any(c.err_inspect for n in want for c in n.containers)
    else:
        assert left == 0
