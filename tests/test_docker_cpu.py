"""CPU: the device-side `docker inspect` scanner (csrc/kgm_docker.cuh, host build) against the
reference's own rules as kgm/fleet.py states them with json.loads and str.split (gpu.py:86-88,184-199).

The scanner may leave any container to the host (None); whatever it takes must be exactly what the
reference computes, and the reference must raise nothing on it."""
import json
import random

from hypothesis import given, settings, strategies as st

import hostcheck as hc
from kgm import fleet, replay


def reference(text: str):
    """(gated, podname, namespace, user) as gpu.py:87-88,184-199 computes them, or the exception."""
    squeezed = text.replace("\n", "").replace(" ", "")
    labels = json.loads(squeezed)[0]["Config"]["Labels"]
    gate = "False"
    if "mlgpu" in labels.keys():
        gate = labels["mlgpu"]
    gated = gate == "mlgpu"
    pod = ns = user = None
    if "io.kubernetes.pod.name" in labels.keys():
        pod = labels["io.kubernetes.pod.name"]
        ns = labels["io.kubernetes.pod.namespace"]
        if gated:
            user = fleet._user_name(pod)
    return gated, pod, ns, user


def check(text: str, must_take=False):
    got = hc.inspect_fields(text.encode("utf-8", "surrogatepass") if isinstance(text, str) else text)
    if got is None:
        assert not must_take, text
        return False
    want = reference(text)                            # raises if the scanner took what the reference dies on
    assert got[0] == want[0], text
    if want[1] is None:
        assert got[1] is None, text
    else:
        assert got[1:3] == want[1:3], text
        if want[0]:
            assert got[3] == want[3], text
    return True


def test_every_container_of_the_replay_is_taken():
    taken = 0
    for node in range(6):
        snap = replay.make_node(5, node, gpus=8, containers=12, procs_per_container=2, idle_every=3, ungated_every=4)
        for cid, text in snap.inspect.items():
            taken += check(text, must_take=True)
    assert taken == 72


def test_pod_name_rules():
    def doc(pod, ns="ns", gate="mlgpu"):
        return json.dumps([{"Config": {"Labels": {"mlgpu": gate, "io.kubernetes.pod.name": pod,
                                                   "io.kubernetes.pod.namespace": ns}}}], indent=2)
    for pod in ["turing-gpu-notebook-alice-5d9f-xk2p", "turing-gpu-notebook-alice-bob-5d9f-xk2p", "turing-gpu-notebook-x",
                "turing-gpu-notebook-", "a-turing-gpu-notebook-b-c-d-turing-gpu-notebook-e", "jupyter-q-0", "a-b", "a-",
                "-", "x-turing-gpu-notebook-a-b"]:
        assert check(doc(pod), must_take=True), pod
    assert hc.inspect_fields(doc("nodash").encode()) is None          # IndexError gpu.py:199: the host's
    assert check(doc("nodash", gate="no"), must_take=True)            # ... but not read for an ungated container
    assert check(doc("a-b", gate="ml gpu"), must_take=True)           # the removed space makes it the gate
    assert hc.inspect_fields(doc("a -b").encode()) is None            # a value with a removed byte: the host's


def test_shapes_the_scanner_leaves_to_the_host():
    base = {"Config": {"Labels": {"mlgpu": "mlgpu"}}}
    for text in ["", "{}", "[]", "[1]", "[{}]", '[{"Config": 1}]', '[{"Config": {}}]', '[{"Config": {"Labels": null}}]',
                 '[{"Config": {"Labels": {"mlgpu": 1}}}]', json.dumps([base]) + "x", json.dumps([base])[:-1],
                 '[{"Config": {"Labels": {"mlgpu": "ml\\u0067pu"}}}]', '[{"Config": {"Labels": {"mlgpu": "mlgp\xfc"}}}]',
                 '[{"Config": {"Labels": {}}, "Config": {"Labels": {}}}]',
                 '[{"Config": {"Labels": {"io.kubernetes.pod.name": "a-b"}}}]', "[NaN]", "[{\"Config\":{\"Labels\":{}}},]"]:
        assert hc.inspect_fields(text.encode()) is None, text
    # valid JSON of every kind around the labels is taken
    rich = [{"Id": "x" * 64, "Args": ["-c", "sleep 1"], "State": {"Running": True, "Pid": 123, "ExitCode": -0, "F": 1.5e-3,
             "Nil": None}, "Config": {"Env": ["A=1"], "Labels": {"mlgpu": "mlgpu", "other": "v"}, "Cmd": None},
             "Mounts": [], "Nested": [[[{"a": [1, 2, {"b": []}]}]]]}, {"second": "element"}, 7]
    assert check(json.dumps(rich, indent=4), must_take=True)
    assert check(json.dumps(rich).replace(",", ",\t").replace(":", " :\r "), must_take=True)


TEXT = st.lists(st.sampled_from(list("abMl-gpu0 \n\t{}[]\":,\\.eE+tfn1") + ["turing-gpu-notebook-", "mlgpu",
                                     "io.kubernetes.pod.name", "io.kubernetes.pod.namespace", "Config", "Labels", "\u00e9"]),
                max_size=12).map("".join)
LABEL_VALUE = st.one_of(TEXT, st.integers(-5, 5), st.none(), st.booleans(), st.lists(st.integers(0, 3), max_size=2))
LABELS = st.one_of(st.none(), st.dictionaries(
    st.one_of(st.sampled_from(["mlgpu", "io.kubernetes.pod.name", "io.kubernetes.pod.namespace", "ml gpu"]), TEXT),
    LABEL_VALUE, max_size=5))
DOC = st.lists(st.one_of(
    st.fixed_dictionaries({"Config": st.one_of(st.none(), st.fixed_dictionaries({"Labels": LABELS}),
                                               st.fixed_dictionaries({"Env": st.lists(TEXT, max_size=2), "Labels": LABELS}))},
                          optional={"Id": TEXT, "N": st.floats(allow_nan=False, allow_infinity=False)}),
    st.integers(0, 9), TEXT), max_size=3)


@settings(max_examples=600, deadline=None)
@given(doc=DOC, indent=st.sampled_from([None, 1, 4]), damage=st.integers(0, 40), seed=st.integers(0, 2 ** 20))
def test_whatever_the_scanner_takes_is_what_the_reference_computes(doc, indent, damage, seed):
    text = json.dumps(doc, indent=indent)
    if damage < 12 and text:                          # a random edit: drop / double / replace a byte
        rng = random.Random(seed)
        i = rng.randrange(len(text))
        text = [text[:i] + text[i + 1:], text[:i] + text[i] + text[i:],
                text[:i] + rng.choice(' \n\t"{}[]:,\\x0-') + text[i + 1:]][damage % 3]
    check(text)


NAME = st.lists(st.sampled_from(list("abz09-_. ") + ["turing-gpu-notebook-", "user00017", "-5d9f", "-xk2p"]),
                max_size=8).map("".join)
CLEAN_LABELS = st.dictionaries(
    st.one_of(st.sampled_from(["mlgpu", "io.kubernetes.pod.name", "io.kubernetes.pod.namespace", "m l g p u",
                               "annotation.io.kubernetes.container.restartCount"]), NAME),
    st.one_of(NAME, st.sampled_from(["mlgpu", "ml gpu", "False"])), max_size=6)
CLEAN_DOC = st.lists(st.fixed_dictionaries(
    {"Id": NAME, "Config": st.fixed_dictionaries({"Hostname": NAME, "Labels": CLEAN_LABELS}, optional={"Env": st.lists(NAME, max_size=2)})},
    optional={"State": st.fixed_dictionaries({"Pid": st.integers(0, 10 ** 7), "Running": st.booleans()})}), min_size=1, max_size=2)


@settings(max_examples=500, deadline=None)
@given(doc=CLEAN_DOC, indent=st.sampled_from([None, 2, 4]), damage=st.integers(0, 40), seed=st.integers(0, 2 ** 20))
def test_mostly_clean_documents(doc, indent, damage, seed):
    text = json.dumps(doc, indent=indent)
    if damage < 6:
        rng = random.Random(seed)
        i = rng.randrange(len(text))
        text = [text[:i] + text[i + 1:], text[:i] + text[i] + text[i:],
                text[:i] + rng.choice(' \n\t"{}[]:,\\x0-') + text[i + 1:]][damage % 3]
    check(text)
