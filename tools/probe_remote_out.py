"""k_poll writing its lines into a peer GPU's memory (kgm_set_output_buffer) vs its own:
one process, context on cuda:1, buffer on cuda:0."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "k8s-gpu-monitor_b200"))
import torch
from kgm import monitor, replay, shard
from kgm.fleet import node_entry
rp = replay.make_replay(seed=1, **replay.CONFIGS["config2"])
entries = [node_entry(s.name, s.ps_text, s.stats_text, s.top, s.inspect, s.smi_list) for s in rp.nodes]
print("peer access 1->0:", torch.cuda.can_device_access_peer(1, 0))
a = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda:0"); b = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda:1")
b.copy_(a); a.copy_(b); torch.cuda.synchronize(0); torch.cuda.synchronize(1)      # torch enables peer access both ways
with monitor.Monitor(1) as mon:
    mon.load_fleet(entries)
    mon.upload(mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
    r = mon.encode_resident(); nb = int(r.n_bytes)
    want = mon.fetch().body
    def run(tag):
        ts = []
        for _ in range(12):
            mon.encode_resident(); ts.append(mon.timing()["encode_ms"])
        print("%-7s k_poll %.4f ms (min %.4f)" % (tag, sum(ts[2:]) / 10, min(ts)))
    run("local")
    remote = torch.zeros(nb + (1 << 20), dtype=torch.uint8, device="cuda:0")
    mon.set_output_buffer(remote.data_ptr(), remote.numel())
    run("remote")
    torch.cuda.synchronize(1)
    hdr = remote[:16].view(torch.int64).cpu().tolist()
    got = remote[16:16 + nb].cpu().numpy().tobytes()
    print("header", hdr, "bytes equal:", got == want)
    mon.set_output_buffer(0, 0)
    run("local")
