"""Multi-GPU plumbing: partition of the fleet by host and assembly of the ranks' outputs.

Every series key of the reference starts with ``host=`` (gpu.py:137,207,241) and every join
input is node-local, so whole hosts are the partition unit: all documents of a host go to
one rank and no sample ever crosses GPUs.  The only exchange is (1) an all-gather of each
rank's ``{bytes, lines}`` so that every rank knows where its slice of the fleet-wide body
starts, and (2) the concatenation of the encoded buffers, rank-major.  One process per GPU;
``torch.distributed`` (NCCL on the GPU box, gloo in the CPU tests) carries both.
"""
from __future__ import annotations

import hashlib

import torch
import torch.distributed as dist


def host_hash(name: str) -> int:
    return int.from_bytes(hashlib.blake2b(name.encode(), digest_size=8).digest(), "little")


def assign_hosts(names, weights, world: int):
    """host -> rank.  Size-aware (largest first onto the least loaded rank) so that skewed
    hosts do not unbalance the shards; ties and order are fixed by a hash of the host name,
    so every rank computes the same map without talking."""
    order = sorted(range(len(names)), key=lambda i: (-int(weights[i]), host_hash(names[i]), names[i]))
    load = [0] * world
    out = [0] * len(names)
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[i] = r
        load[r] += int(weights[i])
    return out


def gather_counts(n_bytes: int, n_lines: int, device, group=None) -> torch.Tensor:
    """All-gather of the per-rank result sizes -> int64 [world, 2]."""
    world = dist.get_world_size(group)
    mine = torch.tensor([n_bytes, n_lines], dtype=torch.int64, device=device)
    out = torch.empty((world, 2), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out.view(-1), mine, group=group)
    return out


def slice_offsets(counts: torch.Tensor):
    """Exclusive scan of the gathered counts: where each rank's bytes / lines start."""
    c = counts.cpu()
    off = torch.zeros((c.shape[0] + 1, 2), dtype=torch.int64)
    off[1:] = torch.cumsum(c, 0)
    return off


def concat_encoded(local: torch.Tensor, counts: torch.Tensor, out: torch.Tensor = None, group=None):
    """Rank-major concatenation of every rank's encoded bytes on every rank.

    ``local`` is this rank's uint8 buffer (its first ``counts[rank,0]`` bytes are valid).
    The variable-length all-gather is one padded ``all_gather_into_tensor`` followed by a
    compaction of the padding on the device."""
    world = dist.get_world_size(group)
    off = slice_offsets(counts)
    sizes = counts[:, 0].cpu().tolist()
    maxb = max(max(sizes), 1)
    maxb = (maxb + 15) // 16 * 16
    padded = torch.empty(maxb, dtype=torch.uint8, device=local.device)
    padded[:sizes[dist.get_rank(group)]] = local[:sizes[dist.get_rank(group)]]
    gathered = torch.empty((world, maxb), dtype=torch.uint8, device=local.device)
    dist.all_gather_into_tensor(gathered.view(-1), padded, group=group)
    total = int(off[-1, 0])
    if out is None:
        out = torch.empty(max(total, 1), dtype=torch.uint8, device=local.device)
    for r in range(world):
        out[int(off[r, 0]):int(off[r, 0]) + sizes[r]] = gathered[r, :sizes[r]]
    return out[:total], off


class DevicePointer:
    """Zero-copy torch view of a raw device buffer owned by libkgm (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (max(int(nbytes), 1),), "typestr": "|u1",
                                         "data": (int(ptr), False), "version": 2}


def tensor_of_device_buffer(ptr: int, nbytes: int, device) -> torch.Tensor:
    return torch.as_tensor(DevicePointer(ptr, nbytes), device=device)
