"""Wall clock of kgm_window_reduce on the resident configs[1] batch (development aid)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200"), os.path.join(ROOT, "tests")]
import torch
from kgm import monitor, replay
from fleet_util import entry_of_snap
rp = replay.make_replay(seed=0, **replay.CONFIGS["config2"])
with monitor.Monitor(0) as mon:
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
    mon.upload(batch)
    mon.encode_resident()
    for w in (10, 60):
        for _ in range(3):
            mon.window_reduce(w)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            r = mon.window_reduce(w)
        torch.cuda.synchronize()
        print("window_ticks %d: %d groups, %.3f ms per call, front end %s" % (w, r["n"], (time.perf_counter() - t0) * 50,
                                                                        mon.window_front()), flush=True)
