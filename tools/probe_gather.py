"""Scratch probe (not part of the product): what the pieces of ShardGather cost between two ranks."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200")]
import torch, torch.distributed as dist
from kgm import shard
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
dev = torch.device("cuda", lr)
n = 46 << 20
buf = torch.empty(n, dtype=torch.uint8, device=dev)
def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
def p2p():
    ops = [dist.P2POp(dist.isend, buf, 1)] if rank == 0 else [dist.P2POp(dist.irecv, buf, 0)]
    for w in dist.batch_isend_irecv(ops): w.wait()
t = timeit(p2p)
if rank == 0: print("nccl p2p 46 MiB: %.3f ms = %.0f GB/s" % (t, n / t / 1e6))
ctl = dist.new_group(backend="nccl")
def counts():
    c = shard.gather_counts(123, 456, dev, ctl); shard.slice_offsets(c)
t = timeit(counts)
if rank == 0: print("counts all-gather + host copy: %.3f ms" % t)
big = torch.empty(2 * n, dtype=torch.uint8, device=dev)
def ag():
    dist.all_gather_into_tensor(big, buf)
t = timeit(ag)
if rank == 0: print("all_gather_into_tensor 2 x 46 MiB: %.3f ms" % t)
sg = shard.ShardGather(n + 16, dev)
def sgstep():
    sg.start(buf, n, 1000)
t = timeit(sgstep)
sg.finish()
if rank == 0: print("ShardGather.start back to back: %.3f ms" % t)
dist.barrier(); dist.destroy_process_group()
