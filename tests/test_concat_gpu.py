"""GPU: kgm_concat_shards -- fixed-stride shard slots packed rank-major from device-side sizes
(the multi-GPU assembly step of kgm/shard.py, here on one GPU with made-up shards)."""
import numpy as np
import pytest

from kgm import _ffi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sizes", [[1000, 333, 0, 4097], [5, 1, 2, 3], [16, 16, 16], [0, 0, 7], [123457, 99991]])
def test_slots_are_packed_in_rank_order(sizes):
    torch = pytest.importorskip("torch")
    lib = _ffi.load()
    dev = torch.device("cuda", 0)
    world = len(sizes)
    stride = (max(sizes) + 64 + 15) // 16 * 16
    rng = np.random.default_rng(sum(sizes))
    host = rng.integers(0, 256, size=world * stride, dtype=np.uint8)
    slots = torch.from_numpy(host).to(dev)
    counts = torch.tensor([[s, 10 * s + 1] for s in sizes], dtype=torch.int64, device=dev)
    for shift in (0, 1, 7):                                  # output not 16-byte aligned either
        out = torch.full((sum(sizes) + 64,), 0xEE, dtype=torch.uint8, device=dev)
        totals = torch.zeros(2, dtype=torch.int64, device=dev)
        rc = lib.kgm_concat_shards(slots.data_ptr(), stride, counts.data_ptr(), world, out.data_ptr() + shift,
                                   out.numel() - shift, totals.data_ptr(), None)
        torch.cuda.synchronize(dev)
        assert rc == 0
        want = b"".join(host[r * stride:r * stride + s].tobytes() for r, s in enumerate(sizes))
        got = out.cpu().numpy()
        assert got[shift:shift + len(want)].tobytes() == want
        assert np.all(got[:shift] == 0xEE) and np.all(got[shift + len(want):] == 0xEE)   # nothing else touched
        assert totals.cpu().tolist() == [sum(sizes), sum(10 * s + 1 for s in sizes)]


def test_slots_that_carry_their_own_sizes():
    torch = pytest.importorskip("torch")
    lib = _ffi.load()
    dev = torch.device("cuda", 0)
    sizes, stride = [700, 0, 4099], 4128
    rng = np.random.default_rng(1)
    host = rng.integers(0, 256, size=len(sizes) * stride, dtype=np.uint8)
    for r, s in enumerate(sizes):
        host[r * stride:r * stride + 16] = np.frombuffer(np.asarray([s, s + 5], np.int64).tobytes(), np.uint8)
    slots = torch.from_numpy(host).to(dev)
    out = torch.zeros(sum(sizes) + 32, dtype=torch.uint8, device=dev)
    totals = torch.zeros(2, dtype=torch.int64, device=dev)
    assert lib.kgm_concat_shards(slots.data_ptr(), stride, None, len(sizes), out.data_ptr(), out.numel(),
                                 totals.data_ptr(), None) == 0
    torch.cuda.synchronize(dev)
    want = b"".join(host[r * stride + 16:r * stride + 16 + s].tobytes() for r, s in enumerate(sizes))
    assert out.cpu().numpy()[:len(want)].tobytes() == want
    assert totals.cpu().tolist() == [sum(sizes), sum(s + 5 for s in sizes)]


def test_a_size_beyond_the_slot_or_the_output_is_reported():
    """A header that claims more than its slot holds, or shards that do not fit the output:
    the totals read {UINT64_MAX, UINT64_MAX} and nothing is written past the buffers."""
    torch = pytest.importorskip("torch")
    lib = _ffi.load()
    dev = torch.device("cuda", 0)
    stride = 4128
    for sizes, out_bytes in (([700, 5000, 64], 8192), ([700, 800, 64], 1000)):
        host = np.zeros(len(sizes) * stride, np.uint8)
        for r, s in enumerate(sizes):
            host[r * stride:r * stride + 16] = np.frombuffer(np.asarray([s, 1], np.int64).tobytes(), np.uint8)
        slots = torch.from_numpy(host).to(dev)
        out = torch.full((out_bytes + 64,), 0xEE, dtype=torch.uint8, device=dev)
        totals = torch.zeros(2, dtype=torch.int64, device=dev)
        assert lib.kgm_concat_shards(slots.data_ptr(), stride, None, len(sizes), out.data_ptr(), out_bytes,
                                     totals.data_ptr(), None) == 0
        torch.cuda.synchronize(dev)
        assert totals.cpu().tolist() == [-1, -1]
        assert out.cpu().numpy()[out_bytes:].tobytes() == b"\xee" * 64
