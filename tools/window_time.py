"""Stage 3 (kgm_window_reduce) on BASELINE configs[1]: wall clock of the call with the batch
resident, window = 10 ticks."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "k8s-gpu-monitor_b200"))
from kgm import monitor, replay
from kgm.fleet import node_entry
rp = replay.make_replay(seed=1, **replay.CONFIGS["config2"])
entries = [node_entry(s.name, s.ps_text, s.stats_text, s.top, s.inspect, s.smi_list) for s in rp.nodes]
with monitor.Monitor(0) as mon:
    mon.load_fleet(entries)
    mon.upload(mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick))
    for w in (10, 60):
        ts = []
        for _ in range(8):
            t0 = time.perf_counter()
            out = mon.window_reduce(w)
            ts.append(1e3 * (time.perf_counter() - t0))
        print("window %2d ticks: %d groups, %.3f ms per call (min %.3f), %d samples" % (
            w, out["n"], sum(ts[2:]) / 6, min(ts), int(out["count"].sum())))
