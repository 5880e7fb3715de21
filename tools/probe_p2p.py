"""Scratch probe (not part of the product): peer copy bandwidth between two GPUs of one process,
and the timing dict of the config-5 miniature."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200"), os.path.join(ROOT, "tests")]
import torch
n = torch.cuda.device_count()
print("devices", n)
if n >= 2:
    print("can access peer 1->0:", torch.cuda.can_device_access_peer(1, 0))
    a = torch.empty(64 << 20, dtype=torch.uint8, device="cuda:1")
    b = torch.empty(64 << 20, dtype=torch.uint8, device="cuda:0")
    for _ in range(3):
        b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    t0 = time.perf_counter()
    for _ in range(10):
        b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    dt = (time.perf_counter() - t0) / 10
    print("peer copy 64 MiB: %.3f ms = %.1f GB/s" % (dt * 1e3, (64 << 20) / dt / 1e9))
    os.system("nvidia-smi topo -m | head -8")
from kgm import monitor, replay
from fleet_util import entry_of_snap
rp = replay.make_skewed_replay(seed=5, hosts=40, epochs=3, ticks_per_epoch=4)
with monitor.Monitor(0) as mon:
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    res = mon.encode(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick)
    print(mon.timing(), int((res.status != 0).sum()), rp.n_docs)
    print("max texts", max(map(len, rp.gpu_texts)), max(map(len, rp.apps_texts)))
