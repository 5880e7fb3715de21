"""GPU: the allocation-free tokenizer entry point (kgm_tokenize_gpu_noalloc).

Some users of this library cannot allocate device memory themselves -- a collector sharing
a GPU whose memory a training job has fully reserved, or one running under a memory
manager that owns all allocations.  For them the stage-1 tokenizer exists as one extra
entry point whose working storage is a fixed array inside the library.  This test calls it
through the C ABI twice in a row (the second call on a different, smaller batch, so that
anything left in the shared array by the first call would show) and checks what it
decodes -- gpu_util, mem_used, mem_total per sample -- against what the ordinary entry point
(kgm_encode_batch) prints for the same texts.  The values are integers: the comparison is
exact, like every other parity test here.  The same values are also derived from the text
with the oracle's own rules (csv.reader cells, poll_oracle.digits_of), and a batch with a
damaged document and a repeated uuid pins what the entry point reports for those: raw rows in
text order (no last-row-wins dictionary) and a per-document parse status.

The file name makes pytest collect it last; it also passes when run alone.
"""
import csv
import io
import re

import numpy as np
import pytest

from fleet_util import entry_of_snap
from kgm import _ffi, monitor, replay

pytestmark = pytest.mark.gpu

_NODE = re.compile(r"^node_gpu,.* gpu_util=(\d*),mem_util=[^,]*,mem_used=(\d+),mem_total=(\d+),")


def _reference_samples(mon, rp):
    """(util, mem_used, mem_total) per document, from the lines of the existing entry point."""
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    res = mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node)
    out = []
    for d in range(rp.n_docs):
        rows = []
        for line in res.lines(d):
            m = _NODE.match(line)
            if m:
                rows.append((int(m.group(1) or 0), int(m.group(2)), int(m.group(3))))
        out.append(rows)
    assert not res.status.any()
    return out


def _oracle_samples(gpu_texts):
    """The same three integers per --query-gpu row, from the text alone: cells as csv.reader
    splits them (gpu.py:47), digits as gpu.py:133-134,138 strip them."""
    from oracle import poll_oracle
    out = []
    for text in gpu_texts:
        rows = list(csv.reader(io.StringIO(text)))[1:]                   # header dropped blind (gpu.py:67-68)
        dig = lambda row, i: poll_oracle.digits_of(row[i]) if i < len(row) else ""
        out.append([(int(dig(r, 1) or 0), int(dig(r, 4) or 0), int(dig(r, 3) or 0)) for r in rows])
    return out


def _noalloc_samples(lib, rp, max_rows, want_err=None):
    import torch
    dev = torch.device("cuda", 0)
    texts = [t.encode() for t in rp.gpu_texts]
    off = np.zeros(len(texts) + 1, np.int64)
    np.cumsum([len(t) for t in texts], out=off[1:])
    raw = b"".join(texts)
    csv = torch.zeros(len(raw) + 64, dtype=torch.uint8, device=dev)      # padded, torch aligns to 512 B
    csv[:len(raw)] = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    n = len(texts)
    row0 = torch.full((n,), -1, dtype=torch.int32, device=dev)
    nrows = torch.full((n,), -1, dtype=torch.int32, device=dev)
    err = torch.full((n,), -1, dtype=torch.int32, device=dev)
    util = torch.full((max_rows,), -1, dtype=torch.int32, device=dev)
    used = torch.full((max_rows,), -1, dtype=torch.int64, device=dev)
    total = torch.full((max_rows,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)
    rc = lib.kgm_tokenize_gpu_noalloc(csv.data_ptr(), d_off.data_ptr(), n, len(raw), max_rows,
                                      row0.data_ptr(), nrows.data_ptr(), err.data_ptr(),
                                      util.data_ptr(), used.data_ptr(), total.data_ptr(), None)
    torch.cuda.synchronize(dev)
    if rc != 0:
        return rc, None
    row0, nrows, err = row0.cpu().numpy(), nrows.cpu().numpy(), err.cpu().numpy()
    util, used, total = util.cpu().numpy(), used.cpu().numpy(), total.cpu().numpy()
    if want_err is None:
        assert not err.any()
    else:
        assert err.tolist() == want_err
    out = []
    for d in range(n):
        a, k = int(row0[d]), int(nrows[d])
        out.append([(int(util[a + i]), int(used[a + i]), int(total[a + i])) for i in range(k)])
    # the documents' row blocks tile [0, total rows) without overlap
    spans = sorted((int(row0[d]), int(nrows[d])) for d in range(n) if nrows[d])
    pos = 0
    for a, k in spans:
        assert a == pos
        pos += k
    return rc, out


def test_noalloc_tokenizer_matches_the_ordinary_entry_point_twice_in_a_row():
    pytest.importorskip("torch")
    lib = _ffi.load()
    first = replay.make_replay(seed=31, nodes=3, ticks=4, gpus=8, containers=4, procs_per_container=2)
    second = replay.make_replay(seed=32, nodes=2, ticks=3, gpus=3, containers=2, procs_per_container=1)
    with monitor.Monitor(0) as mon:
        want_first = _reference_samples(mon, first)
        want_second = _reference_samples(mon, second)
    rc, got = _noalloc_samples(lib, first, 256)
    assert rc == 0 and got == want_first == _oracle_samples(first.gpu_texts)
    rc, got = _noalloc_samples(lib, second, 64)          # straight after, smaller batch
    assert rc == 0 and got == want_second == _oracle_samples(second.gpu_texts)
    assert sum(len(r) for r in want_first) == 96 and sum(len(r) for r in want_second) == 18


def test_noalloc_tokenizer_reports_raw_rows_and_parse_status():
    """A blank line inside a text is a parse-level failure of that document (KeyError at
    gpu.py:51); a repeated uuid is NOT folded (the entry point returns rows, the dictionary
    semantics of gpu.py:54 belong to the full path); "[N/A]" cells decode to 0."""
    pytest.importorskip("torch")
    lib = _ffi.load()
    rp = replay.make_replay(seed=34, nodes=1, ticks=3, gpus=4, containers=1, procs_per_container=1)
    rows = rp.gpu_texts[1].split("\n")
    rp.gpu_texts[1] = "\n".join(rows[:2] + [""] + rows[2:])             # blank line after the first row
    rows = rp.gpu_texts[2].split("\n")
    uuid = rows[1].split(",")[0]
    rp.gpu_texts[2] = "\n".join(rows[:3] + ["%s, [N/A], [N/A], 100 MiB, 7 MiB, 93 MiB" % uuid] + rows[3:])
    rc, got = _noalloc_samples(lib, rp, 64, want_err=[0, _ffi.DOC_KEY_ERROR, 0])
    assert rc == 0
    want = _oracle_samples(rp.gpu_texts)
    assert got[0] == want[0] and got[2] == want[2]
    assert len(got[2]) == 5 and got[2][2] == (0, 7, 100)                 # the repeated uuid stays a row of its own
    assert [r for r in got[1] if r != (0, 0, 0)] == want[1][:1] + want[1][2:]   # the blank line is a row without cells


def test_noalloc_tokenizer_refuses_what_does_not_fit():
    pytest.importorskip("torch")
    lib = _ffi.load()
    rp = replay.make_replay(seed=33, nodes=1, ticks=2, gpus=8, containers=1, procs_per_container=1)
    rc, _ = _noalloc_samples(lib, rp, _ffi.NOALLOC_MAX_ROWS + 1)    # larger than the fixed array
    assert rc == _ffi.E_LIMIT
    rc, _ = _noalloc_samples(lib, rp, 8)                            # 16 rows in the text, room for 8
    assert rc == _ffi.E_LIMIT
    rc, got = _noalloc_samples(lib, rp, 16)                         # and it still works afterwards
    assert rc == 0 and sum(len(r) for r in got) == 16
