"""NCCL all-gather of fixed-size shards alone (no other kernels): what the exchange of the
multi-GPU bench costs per step when it has the GPUs to itself.
torchrun --nproc-per-node N tools/probe_allgather.py [MB per rank]"""
import os, sys
import torch, torch.distributed as dist
rank = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
mb = float(sys.argv[1]) if len(sys.argv) > 1 else 48
n = int(mb * 1e6) // 16 * 16
src = torch.empty(n, dtype=torch.uint8, device="cuda")
dst = torch.empty(n * world, dtype=torch.uint8, device="cuda")
for _ in range(5):
    dist.all_gather_into_tensor(dst, src)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    dist.all_gather_into_tensor(dst, src)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
if rank == 0:
    print("all_gather %d ranks x %.1f MB: %.3f ms, ingress per rank %.0f GB/s" % (world, n / 1e6, ms, (world - 1) * n / ms / 1e6))
dist.destroy_process_group()
