"""CPU: the kernels' per-thread leaf functions (host build of the same headers) against
Python's csv / re / repr and against the oracle's line templates."""
import csv
import io
import random
import re

import numpy as np
import pytest

import golden_io
import hostcheck as hc
from oracle import poll_oracle

GOLD = golden_io.load()


def test_float_repr_matches_python_repr():
    rng = np.random.default_rng(5)
    bits = rng.integers(1, 0x7FF0000000000000, size=60000, dtype=np.uint64).view(np.float64)
    u = rng.integers(0, 200000, size=60000)
    t = rng.integers(1, 200000, size=60000)
    special = [1.0, 10.0, 1e15, 1e16, 1e17, 1e22, 1e23, 5e-324, 2.2250738585072014e-308,
               1.7976931348623157e308, 1e-4, 1e-5, 1e-3, 9007199254740993.0, 0.1 + 0.2, 2.0 ** 63,
               2.0 ** 64, 123456789012345680.0] + [float(i) for i in range(300)] + \
              [10.0 ** k for k in range(-25, 25)] + [2.0 ** k for k in range(-70, 70)]
    for xs in (bits, u * 100.0 / t, np.asarray(special)):
        got = hc.float_repr_many(xs)
        want = [repr(float(x)) for x in xs]
        assert got == want


def test_mem_util_is_pythons_arithmetic():
    rng = random.Random(3)
    pairs = [(1, 183359), (183359, 183359), (1, 3), (7, 9), (61120, 81920), (1, 1000001),
             (0, 9), (2 ** 63 + 12345, 7), (9007199254740993, 1), (18446744073709551615, 3)]
    pairs += [(rng.randrange(0, 2 ** 64), rng.randrange(1, 2 ** 64)) for _ in range(2000)]
    pairs += [(rng.randrange(0, 200000), rng.randrange(1, 200000)) for _ in range(5000)]
    for used, total in pairs:
        assert hc.mem_util_repr(used, total) == repr(used * 100.0 / total), (used, total)


def _py_cells(row: str):
    rows = list(csv.reader(io.StringIO(row)))
    return rows[0] if rows else []


ROW_POOL = []
for g in GOLD:
    for text in (g["gpu_csv"], g["apps_csv"]):
        ROW_POOL.extend(r for r in text.split("\n")[1:] if "\r" not in r and '"' not in r)


def _mutate(rng, row):
    for _ in range(rng.randrange(3)):
        if not row:
            break
        i = rng.randrange(len(row))
        k = rng.randrange(5)
        if k == 0:
            row = row[:i] + row[i + 1:]
        elif k == 1:
            row = row[:i] + rng.choice(", 7x\t[") + row[i:]
        elif k == 2:
            row = row[:i] + row[i + rng.randrange(1, 12):]
        elif k == 3:
            row = row[:i] + rng.choice("0123456789") * rng.randrange(1, 25) + row[i:]
        else:
            row = " " * rng.randrange(3) + row
    return row


def test_gpu_row_scanner_matches_csv_and_re():
    rng = random.Random(11)
    rows = [r for r in ROW_POOL if r.startswith(("GPU", "\t"))][:200]
    rows += [_mutate(rng, rng.choice(rows)) for _ in range(3000)] + ["", " ", ",", ",,,,,,,", "a" * 300]
    for row in rows:
        cells = _py_cells(row)
        got = hc.scan_gpu_row(row.encode())
        assert got["ncells"] == min(len(cells), 255), row
        assert got["parse_error"] == (1 if not cells else 0)
        lo = 0
        for i, cell in enumerate(cells[:6]):
            assert row[lo:got["cell_end"][i]] == cell
            lo = got["cell_end"][i] + 1
        if cells:
            key = cells[0].lstrip()
            assert row[got["key_lo"]:got["cell_end"][0]] == key
            assert got["key_hash"] == hc.hash_bytes(key.encode())
        for slot, col in enumerate((1, 3, 4, 5)):
            if col < len(cells):
                digits = re.sub(r"\D", "", cells[col])
                assert got["nd"][slot] == min(len(digits), 255)
                if col in (3, 4) and 0 < len(digits) <= 19:
                    assert got["mem_total" if col == 3 else "mem_used"] == int(digits)


def test_proc_row_scanner_matches_csv_and_re():
    rng = random.Random(12)
    rows = [r for r in ROW_POOL if not r.startswith(("GPU", "\t", '"')) and r.isascii()][:300]
    rows += [_mutate(rng, rng.choice(rows)) for _ in range(3000)] + ["", "1", "1,2", "1,2,3", ",,,"]
    for row in rows:
        cells = _py_cells(row)
        got = hc.scan_proc_row(row.encode())
        assert got["ncells"] == min(len(cells), 255), row
        assert got["parse_error"] == (1 if len(cells) < 3 else 0)
        if len(cells) >= 1:
            pid = cells[0].encode()
            assert got["pid_len"] == min(len(pid), 255)
            if len(pid) <= 15:
                packed = got["pid_lo"].to_bytes(8, "little") + got["pid_hi"].to_bytes(8, "little")
                assert packed[:len(pid)] == pid and packed[15] == len(pid)
                assert packed[len(pid):15] == b"\0" * (15 - len(pid))
            else:
                assert got["flags"] & 16
        if len(cells) >= 2:
            assert got["name_len"] == len(cells[1].replace(" ", ""))
        if len(cells) >= 3:
            key = cells[2].lstrip()
            lo = got["cell_end"][1] + 1
            assert row[got["key_lo"]:got["cell_end"][2]] == key
            assert row[lo:got["cell_end"][2]] == cells[2]
            assert got["uuid_hash"] == hc.hash_bytes(key.encode())
        if len(cells) >= 4:
            assert got["nd_mem"] == min(len(re.sub(r"\D", "", cells[3])), 255)


def test_utf8_process_name_and_quoted_uuid():
    """Bytes >= 0x80 pass through in the process name (it is only copied, gpu.py:203) and are
    flagged anywhere else; a first gpu cell in quotes that close in front of the delimiter is
    the text between the quotes (csv.reader), any other leading quote is flagged."""
    row = "34443, /home/ren\u00e9/\u6a21\u578b/py thon, GPU-aaaa, 15524 MiB"
    got = hc.scan_proc_row(row.encode())
    assert got["flags"] == 0 and got["ncells"] == 4
    assert got["name_len"] == len(_py_cells(row)[1].replace(" ", "").encode())
    assert hc.scan_proc_row("34443, python, GPU-\u00e9, 1 MiB".encode())["flags"] & 4
    assert hc.scan_proc_row("3\u0664, python, GPU-a, 1 MiB".encode())["flags"] & 4
    assert hc.scan_gpu_row("GPU-a, 1 %, 2 %, 3 MiB, \u0664 MiB, 5 MiB".encode())["flags"] & 4
    row = '"GPU-aaaa", 49 %, 28 %, 16160 MiB, 15524 MiB, 636 MiB'
    got = hc.scan_gpu_row(row.encode())
    cells = _py_cells(row)
    assert cells[0] == "GPU-aaaa" and got["flags"] == 0 and got["ncells"] == 6
    assert row[got["key_lo"]:got["cell_end"][0]] == "GPU-aaaa"
    assert got["key_hash"] == hc.hash_bytes(b"GPU-aaaa")
    assert got["nd"] == [2, 5, 5, 3] and got["mem_total"] == 16160 and got["mem_used"] == 15524
    assert hc.scan_gpu_row(b'" GPU-a", 1 %')["key_lo"] == 2               # lstrip applies to the content
    for bad in (b'"GPU-a, 1 %', b'"GPU,a", 1 %', b'"GPU-a"x, 1 %', b'"GPU""a", 1 %', b'GPU-a,"1 %"'):
        assert hc.scan_gpu_row(bad)["flags"] & 2, bad
    assert hc.scan_gpu_row(b'GPU-a, 1 " %')["flags"] == 0                 # a quote inside a cell is a character


def test_row_flags():
    assert hc.scan_gpu_row(b"GPU-a, 1 %\r")["flags"] == 0                 # line-terminator CR
    assert hc.scan_gpu_row(b"GPU-a, 1 %\r\r")["ncells"] == 2
    assert hc.scan_gpu_row(b"\r")["ncells"] == 0                          # blank "\r\n" line
    assert hc.scan_gpu_row(b"GPU-a, 1\r %")["parse_error"] == 6           # csv.Error
    assert hc.scan_gpu_row(b'"GPU-a, 1 %')["parse_error"] == 20                # quotes that do not close
    assert hc.scan_gpu_row(b'GPU-a,"1 %"')["parse_error"] == 20
    assert hc.scan_gpu_row(b'GPU-a, "1 %"')["parse_error"] == 0           # quote not at cell start
    assert hc.scan_gpu_row("GPU-é, 1 %".encode())["parse_error"] == 21
    assert hc.scan_gpu_row(b"g,1,2," + b"9" * 20 + b",5,6")["node_error"] == 22
    assert hc.scan_gpu_row(b"g,1,2,0,5,6")["node_error"] == 3
    assert hc.scan_gpu_row(b"g,1,2,x,5,6")["node_error"] == 2
    assert hc.scan_gpu_row(b"g,1,2,3,5")["node_error"] == 1


def _gold_rows(g):
    return g["gpu_csv"].split("\n")[1:], g["apps_csv"].split("\n")[1:]


def test_node_lines_match_the_oracle_template():
    n = 0
    for g in GOLD:
        if g["err"] or "\r" in g["gpu_csv"]:
            continue
        rows = [r for r in _gold_rows(g)[0] if r]
        for sid, row in enumerate(rows):
            cells = _py_cells(row)
            if len(cells) < 6:
                continue
            if max(len(re.sub(r"\D", "", cells[3])), len(re.sub(r"\D", "", cells[4]))) > 19:
                continue                                  # documented limit: KGM_DOC_UNSUPPORTED_RANGE
            rec = dict(zip(["gpu_uuid", "used_gpu", "utilization.memory", "memory.total",
                            "memory.used", "memory.free"], cells))
            try:
                want = poll_oracle.node_line(g["node_name"], str(sid), rec)
            except Exception:
                continue
            got, planned, _ = hc.node_line(str(g["node_name"]), row.encode(), sid)
            assert got == (want + "\n").encode()
            assert planned == len(got)
            n += 1
    assert n > 300


def test_pod_lines_match_the_oracle_template():
    rng = random.Random(2)
    grow = "GPU-aaaa, 4.9 %, 28 %, 16160 MiB, 15524 MiB, 636 MiB"
    prows = ["34443, /opt/conda/envs/py2/bin/python, GPU-aaaa, 15524 MiB", "1, a b  c, GPU-aaaa, [N/A]",
             " 7 ,x, GPU-aaaa,1,5", "123456789012345, n, u, 00 7MiB"]
    for prow in prows:
        for sid in (0, 7, 10, 123):
            host, ns, cid, user, pod = "node-%d" % rng.randrange(99), "ns", "16ee8a182726", "alice-bob", \
                "turing-gpu-notebook-alice-bob-5d9f-xk2p"
            c = _py_cells(prow)
            want = poll_oracle.pod_line(host, ns, str(sid), cid, c[0], re.sub(r"\D", "", " 4.9 %"), user,
                                        c[1].replace(" ", ""), pod, re.sub(r"\D", "", c[3]))
            got, planned, _ = hc.busy_line(host, ns, cid, user, pod, prow.encode(), grow.encode(), sid)
            assert got == (want + "\n").encode()
            assert planned == len(got)
    for sid in (None, 0, 12):
        want = poll_oracle.pod_line("h", "-", sid, "cid", "-", 0, "-", "-", "-", 0)
        got, planned = hc.idle_line("h", "-", "cid", "-", "-", sid)
        assert got == (want + "\n").encode() and planned == len(got)


# ---- fast path of the fused kernel (kgm_fast.cuh) ---------------------------------------------
# The fast path may decline a row ("irregular": the kernel then re-does the document with the
# generic code tested above); whenever it accepts one, its bytes must be the reference's.
def _digits(s):
    return re.sub(r"\D", "", s)


def test_fast_gpu_row_parser_agrees_with_csv_and_re_or_declines():
    rng = random.Random(17)
    base = [r for r in ROW_POOL if r.startswith("GPU")][:200]
    rows = base + [_mutate(rng, rng.choice(base)) for _ in range(4000)]
    accepted = 0
    for row in rows:
        if not row.isascii():
            continue
        for shift in (0, rng.randrange(16)):
            got = hc.fast_gpu_row(row.encode(), shift)
            cells = _py_cells(row) if "\r" not in row and '"' not in row else None
            if got is None:
                continue
            accepted += 1
            assert cells is not None and len(cells) >= 6, row
            assert got["key"].decode() == cells[0].lstrip(), row
            assert got["util"].decode() == _digits(cells[1]), row
            assert got["mt"].decode() == _digits(cells[3]) and got["mu"].decode() == _digits(cells[4]), row
            assert got["mf"].decode() == _digits(cells[5]), row
            assert int(_digits(cells[3])) != 0
            assert got["hash"] == hc.key_hash_bytes(cells[0].lstrip().encode())
    assert accepted > 1500                                   # the clean rows do take the fast path
    typical = re.compile(r"^GPU-[0-9a-f-]+, \d+ %, \d+ %, [1-9]\d{0,8} MiB, \d{1,9} MiB, \d{1,9} MiB$")
    assert sum(1 for row in base if typical.match(row)) > 20
    for row in base:
        if typical.match(row):
            assert hc.fast_gpu_row(row.encode()) is not None, row


def test_fast_proc_row_parser_agrees_with_csv_and_re_or_declines():
    rng = random.Random(19)
    base = [r for r in ROW_POOL if not r.startswith("GPU") and r.count(",") == 3][:200]
    rows = base + [_mutate(rng, rng.choice(base)) for _ in range(4000)]
    accepted = 0
    for row in rows:
        if not row.isascii():
            continue
        shift = rng.randrange(16)
        got = hc.fast_proc_row(row.encode(), shift)
        if got is None:
            continue
        accepted += 1
        assert "\r" not in row and '"' not in row, row
        cells = _py_cells(row)
        assert len(cells) >= 4, row
        assert got["pid"].decode() == cells[0], row
        assert got["name"].decode() == cells[1].replace(" ", ""), row
        assert got["key"].decode() == cells[2].lstrip(), row
        assert got["mem"].decode() == _digits(cells[3]), row
        assert got["hash"] == hc.key_hash_bytes(cells[2].lstrip().encode())
        if len(cells[0]) <= 7:
            assert got["pidkey"] == int.from_bytes(cells[0].encode(), "little") | (len(cells[0]) << 56)
    assert accepted > 1500


def test_fast_lines_are_the_generic_lines_at_every_alignment():
    """node / busy / idle lines composed by the word-wise writer == the byte-wise composer
    (itself pinned to the oracle above), at every stage alignment and text shift, without
    touching a byte outside the line."""
    rng = random.Random(23)
    grows = [r for r in ROW_POOL if r.startswith("GPU")][:60]
    prows = [r for r in ROW_POOL if not r.startswith("GPU") and r.count(",") == 3][:60]
    prows += ["34443, python3 train.py --lr 1, %s, 419 MiB" % "GPU-aaaa", "7, a  b, GPU-x, 1 MiB"]
    hosts = ["node-a", "k8s-node-gpu-00012", "n", "None"]
    checked = 0
    for _ in range(400):
        host = rng.choice(hosts)
        grow, prow = rng.choice(grows), rng.choice(prows)
        sid = rng.choice([0, 1, 7, 9, 10, 15, 99, 100, 255])
        shift, align = rng.randrange(16), rng.randrange(16)
        want, planned_w, n_w = hc.node_line(host, grow.encode(), sid)
        got, planned, guards = hc.fast_node_line(host, grow.encode(), sid, shift, align)
        if got is not None:
            assert want is not None and got == want and planned == len(want) == planned_w and guards, grow
            checked += 1
        else:
            assert want is None or len(want) > 8192 or not hc.fast_gpu_row(grow.encode())
        ns, cid, user, pod = rng.choice([("ns2", "76a082ecc41d", "q", "jupyter-q-0"),
                                         ("10007089-name", "16ee8a182726", "alice-bob",
                                          "turing-gpu-notebook-alice-bob-5d9f-xk2p"), ("-", "c", "-", "-")])
        want, planned_w, n_w = hc.busy_line(host, ns, cid, user, pod, prow.encode(), grow.encode(), sid)
        got, planned, guards = hc.fast_busy_line(host, ns, cid, user, pod, prow.encode(), grow.encode(), sid,
                                                 shift, align)
        if got is not None:
            assert want is not None and got == want and planned == len(want) and guards, (prow, grow)
            checked += 1
        for s in (sid, None):
            want, planned_w = hc.idle_line(host, ns, cid, user, pod, s)
            got, planned, guards = hc.fast_idle_line(host, ns, cid, user, pod, s, align)
            assert got == want and planned == len(want) and guards
            checked += 1
    assert checked > 1200


# ---- k_poll2 (kgm_poll2_leaf.cuh): lines written as 4 or 8 independent jobs -------------------
def test_p2_gpu_row_values_and_fields():
    """parse_gpu_row2: same fields as the v1 fast parser, plus int(digits) of the memory cells
    computed four digits at a time."""
    rng = random.Random(29)
    base = [r for r in ROW_POOL if r.startswith("GPU")][:200]
    rows = base + [_mutate(rng, rng.choice(base)) for _ in range(3000)]
    rows += ["GPU-a, 1 %%, 2 %%, %d MiB, %d MiB, 3 MiB" % (t, u) for t, u in
             [(1, 0), (9, 9), (12, 3), (123, 45), (1234, 999), (12345, 1234), (183359, 99999),
              (10 ** 17 + 7, 10 ** 17 - 1), (99999999, 12345678), (100000001, 1)]]
    accepted = 0
    for row in rows:
        if not row.isascii():
            continue
        shift = rng.randrange(16)
        got, v1 = hc.p2_gpu_row(row.encode(), shift), hc.fast_gpu_row(row.encode(), shift)
        assert (got is None) == (v1 is None), row
        if got is None:
            continue
        accepted += 1
        for k in ("key", "util", "mt", "mu", "mf", "hash"):
            assert got[k] == v1[k], (row, k)
        assert got["total"] == int(got["mt"]) and got["used"] == int(got["mu"]), row
    assert accepted > 1200


@pytest.mark.parametrize("jobs", [4, 8])
def test_p2_lines_are_the_generic_lines_at_every_alignment(jobs):
    """node / busy / idle lines written job by job (last job first) == the byte-wise composer
    (pinned to the oracle above), at every stage alignment and text shift, with the planned
    length equal to the written one and no byte touched outside the line."""
    rng = random.Random(31 + jobs)
    grows = [r for r in ROW_POOL if r.startswith("GPU")][:60]
    prows = [r for r in ROW_POOL if not r.startswith("GPU") and r.count(",") == 3][:60]
    prows += ["34443, python3 train.py --lr 1, %s, 419 MiB" % "GPU-aaaa", "7, a  b, GPU-x, 1 MiB",
              "123456, /a/very/long/path/to/some/binary/that/goes/on/and/on/python3.11, GPU-x, 40000 MiB"]
    hosts = ["node-a", "k8s-node-gpu-00012", "n", "None"]
    checked = 0
    for _ in range(400):
        host = rng.choice(hosts)
        grow, prow = rng.choice(grows), rng.choice(prows)
        sid = rng.choice([0, 1, 7, 9, 10, 15, 99, 100, 255])
        shift, align = rng.randrange(16), rng.randrange(16)
        want, planned_w, n_w = hc.node_line(host, grow.encode(), sid)
        got, planned, guards = hc.p2_node_line(host, grow.encode(), sid, shift, align, jobs)
        if got is not None:
            assert want is not None and got == want and planned == len(want) == planned_w and guards, grow
            checked += 1
        else:
            assert want is None or not hc.fast_gpu_row(grow.encode())
        ns, cid, user, pod = rng.choice([("ns2", "76a082ecc41d", "q", "jupyter-q-0"),
                                         ("10007089-name", "16ee8a182726", "alice-bob",
                                          "turing-gpu-notebook-alice-bob-5d9f-xk2p"), ("-", "c", "-", "-")])
        want, planned_w, n_w = hc.busy_line(host, ns, cid, user, pod, prow.encode(), grow.encode(), sid)
        got, planned, guards = hc.p2_busy_line(host, ns, cid, user, pod, prow.encode(), grow.encode(), sid,
                                               shift, align, jobs)
        if got is not None:
            assert want is not None and got == want and planned == len(want) and guards, (prow, grow)
            checked += 1
        for s in (sid, None):
            want, planned_w = hc.idle_line(host, ns, cid, user, pod, s)
            got, planned, guards = hc.p2_idle_line(host, ns, cid, user, pod, s, align, jobs)
            assert got == want and planned == len(want) and guards
            checked += 1
    assert checked > 1200


def test_p2_power_cell_is_the_staged_scanner_s_value():
    """Stage 3's fused front end reads the 7th --query-gpu cell itself (gpu_row_power); whatever it
    takes is what the staged scanner (GpuRowScan) and k_win_samples make of the same row."""
    rng = random.Random(77)
    WIN_SAT, NONE = 0xFFFFFFFE, 0xFFFFFFFF
    base = "GPU-3a6c7f0e-1c2d-4e5f-8a9b-0c1d2e3f4a5b, 41 %, 17 %, 183359 MiB, 20480 MiB, 162879 MiB"
    cells = ["", ", 415.27 W", ", [N/A]", ", 0.00 W", ", 1000 W", ", 4294967295 W", ", 4294967294 W", ", 42949672.96 W",
             ", 99999999999 W", ", 1.2.3.4 W, extra", ", 7 W,", ",5", ", 12345678901"]
    checked = 0
    for _ in range(300):
        cell = rng.choice(cells) if rng.random() < 0.7 else ", %s W" % "".join(rng.choice("0123456789.") for _ in range(rng.randrange(1, 13)))
        row = (base + cell).encode()
        got = hc.p2_gpu_power(row, rng.randrange(16))
        staged = hc.scan_gpu_row(row)["power"]
        want = min(staged, WIN_SAT) if staged != NONE else NONE      # k_win_samples
        if got is None:
            seventh = cell.split(",")[1] if cell else ""
            assert sum(ch.isdigit() for ch in seventh) > 10, cell
        else:
            assert got == want, (cell, got, want)
            checked += 1
    assert checked > 200
