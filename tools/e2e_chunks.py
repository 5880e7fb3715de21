"""End-to-end time of kgm_encode_batch (host texts in, host lines out) on the configs[1] batch for a
few chunk counts (development aid).   python tools/e2e_chunks.py 4 6 8"""
import os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] != "--one":
    for c in sys.argv[1:]:
        e = dict(os.environ); e["KGM_CHUNKS"] = c
        out = subprocess.run([sys.executable, __file__, "--one"], env=e, stdout=subprocess.PIPE, text=True).stdout
        print("chunks", c, out.strip(), flush=True)
    sys.exit(0)
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200"), os.path.join(ROOT, "tests")]
from kgm import monitor, replay
from fleet_util import entry_of_snap
rp = replay.make_replay(seed=0, **replay.CONFIGS["config2"])
with monitor.Monitor(0) as mon:
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
    for _ in range(5):
        mon.encode_raw(batch, want_host_copy=True)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter()
        mon.encode_raw(batch, want_host_copy=True)
        ts.append(time.perf_counter() - t0)
    ts.sort()
    print("e2e median %.4f ms  min %.4f ms  kernel %s" % (1e3 * ts[len(ts) // 2], 1e3 * ts[0], mon.last_kernel()))
