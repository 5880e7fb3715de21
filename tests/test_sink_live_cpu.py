"""CPU: the sink helpers (kgm/sink.py) and the live single-node producer (kgm/live.py)."""
import gzip
import types

import numpy as np

import cases
import golden_io
from kgm import live, sink
from oracle import poll_oracle, window_oracle

GOLD = golden_io.load()


def _host_result(docs_lines):
    """What Monitor.encode returns, built from lines (no GPU here)."""
    body = b"".join(poll_oracle.body_of(ls) for ls in docs_lines)
    lens = [len(l.encode()) + 1 for ls in docs_lines for l in ls]
    r = types.SimpleNamespace()
    r.body = body
    r.line_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    r.doc_line_off = np.concatenate([[0], np.cumsum([len(ls) for ls in docs_lines])]).astype(np.int64)
    return r


def test_timestamp_suffix_and_bodies():
    docs = [g["lines"] for g in GOLD[:12]]
    res = _host_result(docs)
    ts = [1_700_000_000_000_000_000 + 10 ** 9 * d for d in range(len(docs))]
    got = sink.with_timestamps(res, ts)
    want = "".join("%s %d\n" % (l, ts[d]) for d, ls in enumerate(docs) for l in ls).encode()
    assert got == want
    assert sink.with_timestamps(_host_result([[], []]), [1, 2]) == b""
    for compress in (False, True):
        chunks = sink.write_bodies(want, max_bytes=2000, compress=compress)
        raw = [gzip.decompress(p) if compress else p for p, _ in chunks]
        assert b"".join(raw) == want and len(chunks) > 3
        assert all(c.endswith(b"\n") for c in raw)
        assert all(len(c) <= 2000 or c.count(b"\n") == 1 for c in raw)
        assert all((h.get("Content-Encoding") == "gzip") == compress for _, h in chunks)
    assert sink.write_bodies(b"", 100) == []


def test_window_lines_equal_the_oracle_line_form():
    """node_gpu_w / pod_gpu_w lines: the product's encoder over the array form of the groups ==
    the oracle's encoder over its own groups (parity unpinned by the reference, see
    oracle/window_oracle.py)."""
    docs, hosts, cids = [], [], []
    sel = [g for g in GOLD if g["err"] is None and g["lines"]][:10]
    for n, g in enumerate(sel):
        hosts.append(str(g["node_name"]))
        base = len(cids)
        stats_ids = [row.split(",")[0][1:] for row in g["stats_text"].split("\n") if row]
        cids += stats_ids
        index_of = {c: base + i for i, c in enumerate(stats_ids)}
        for tick in range(5):
            docs.append((n, tick, g["lines"], g["gpu_csv"], index_of))
    groups = window_oracle.windows(docs, 2)
    want = window_oracle.lines(groups, hosts, cids, 2, t0_ns=5, tick_ns=10 ** 9)
    keys = list(groups)
    arr = {"n": len(keys), "kind": np.array([k[0] for k in keys], np.uint8),
           "node": np.array([groups[k]["node"] for k in keys], np.uint32),
           "container": np.array([k[1] if k[0] else 0xFFFFFFFF for k in keys], np.uint32),
           "gpu_id": np.array([k[2] for k in keys], np.uint32), "window": np.array([k[3] for k in keys], np.uint32),
           "count": np.array([groups[k]["count"] for k in keys], np.uint32)}
    for m in ("gpu_util", "mem", "power"):
        for stat, dt in (("count", np.uint32), ("min", np.uint32), ("max", np.uint32), ("last", np.uint32),
                         ("sum", np.uint64), ("mean", np.float64)):
            arr["%s_%s" % (m, stat)] = np.array([groups[k][m][stat] or 0 for k in keys], dt)
    got = sink.window_lines(arr, hosts, cids, 2, t0_ns=5, tick_ns=10 ** 9)
    assert got == want and len(got) > 20
    assert any(l.startswith("pod_gpu_w,host=") for l in got) and any(l.startswith("node_gpu_w,host=") for l in got)
    assert all(l.rsplit(" ", 1)[1].isdigit() for l in got)


class _StubMonitor:
    def __init__(self):
        self.calls = []

    def load_fleet(self, entries):
        self.calls.append(("load_fleet", entries))

    def encode(self, g, a, nodes):
        self.calls.append(("encode", g, a, nodes))
        return types.SimpleNamespace(n_bytes=3, body=b"x\n", lines=lambda d: ["x"], error=lambda d: None)


def test_live_poller_issues_the_references_commands():
    """Same argv and shell pipeline as gpu.py:92-116,177,184,221, containers in
    docker stats order, every text handed on unchanged."""
    g = next(c for c in GOLD if c["name"] == "appendix_b") if any(c["name"] == "appendix_b" for c in GOLD) else GOLD[0]
    issued, piped = [], []

    def command(args):
        issued.append(list(args))
        if args[:2] == ["docker", "ps"]:
            return g["ps_text"]
        if args[:2] == ["docker", "stats"]:
            return g["stats_text"]
        if args[:2] == ["docker", "top"]:
            return g["top"][args[2]]
        if args[:2] == ["docker", "inspect"]:
            return g["inspect"][args[2]]
        return g["gpu_csv"] if args[2].startswith("--query-gpu=") else g["apps_csv"]

    def popen(cmd):
        piped.append(cmd)
        return g["smi_list"].get(cmd.split("docker exec -i ", 1)[1].split(" ", 1)[0], "")

    sent = []
    mon = _StubMonitor()
    poller = live.LivePoller(mon, command=command, popen=popen, node_name=g["node_name"], sink=sent.append)
    got = poller.collect()
    assert issued[:4] == [
        ["docker", "ps", "--format", "'{{.ID}},{{.Image}},{{.Ports}}'"],
        ["docker", "stats", "--no-stream", "--format", "'{{.Container}},{{.MemUsage}},{{.CPUPerc}}'"],
        ["nvidia-smi", "--format=csv",
         "--query-gpu=gpu_uuid,utilization.gpu,utilization.memory,memory.total,memory.used,memory.free"],
        ["nvidia-smi", "--format=csv", "--query-compute-apps=pid,process_name,gpu_uuid,used_memory"]]
    cids = [row.split(",")[0][1:] for row in g["stats_text"].split("\n") if row]
    assert [a[2] for a in issued[4:] if a[1] == "top"] == cids
    assert all(a in (["docker", "top", a[2], "-eo", "pid"], ["docker", "inspect", a[2]]) for a in issued[4:])
    assert piped == ["docker exec -i %s nvidia-smi -L |cut -d: -f3|cut -d\\) -f1" % c for c in cids]
    for k in cases.INPUT_KEYS:
        if k in ("top", "inspect", "smi_list"):
            assert got[k] == {c: g[k].get(c, "") for c in cids} or got[k] == {c: g[k][c] for c in cids}
        else:
            assert got[k] == g[k], k
    lines, err = poller.poll()
    assert (lines, err) == (["x"], None) and sent == [b"x\n"]
    assert mon.calls[0][0] == "load_fleet" and mon.calls[1][1:] == ([g["gpu_csv"]], [g["apps_csv"]], [0])
