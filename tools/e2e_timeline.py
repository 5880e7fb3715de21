"""Where the end-to-end call spends its time (config 2, pinned host buffers): wall clock of
kgm_encode_batch next to the device spans kgm_get_timing reports for the chunked path."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "k8s-gpu-monitor_b200"))
import ctypes
from kgm import monitor, replay, _ffi
from kgm.fleet import node_entry

cfg = replay.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "config2"]
rp = replay.make_replay(seed=1, **cfg)
entries = [node_entry(s.name, s.ps_text, s.stats_text, s.top, s.inspect, s.smi_list) for s in rp.nodes]
with monitor.Monitor(0) as mon:
    mon.load_fleet(entries)
    batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
    for _ in range(5):
        mon.encode_raw(batch, want_host_copy=True)
    for _ in range(8):
        t0 = time.perf_counter()
        mon.encode_raw(batch, want_host_copy=True)
        wall = 1e3 * (time.perf_counter() - t0)
        t = mon.timing()
        print("wall %.3f ms  device total %.3f  first-text %.3f  kernels %.3f  copy-tail %.3f  launches %d" % (
            wall, t["total_ms"], t["h2d_ms"], t["encode_ms"], t["d2h_ms"], t["launches"]))
    tin = (ctypes.c_float * 8)(); tk = (ctypes.c_float * 8)()
    n = mon._lib.kgm_debug_chunk_times(mon._ctx, tin, tk, 8)
    print("chunks", n)
    print("text in     ", " ".join("%.3f" % tin[i] for i in range(max(n, 0))))
    print("kernel done ", " ".join("%.3f" % tk[i] for i in range(max(n, 0))))
