"""Per-stage SM clocks of the fused kernel (needs a library built with
KGM_NVCC_DEFINES=-DKGM_PROFILE_STAGES; rebuild without it afterwards)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200"), os.path.join(ROOT, "tests")]
from kgm import _ffi, monitor, replay
from fleet_util import entry_of_snap
CFG = dict(replay.CONFIGS[os.environ.get("KGM_CFG", "config2")])
if "KGM_TICKS" in os.environ: CFG["ticks"] = int(os.environ["KGM_TICKS"])
rp = replay.make_replay(seed=0, **CFG)
if os.environ.get("KGM_TICK_MAJOR"):                 # documents tick by tick instead of host by host
    import numpy as np
    order = np.lexsort((rp.doc_node, rp.doc_tick))
    rp = replay.Replay(rp.seed, rp.nodes, rp.doc_node[order], rp.doc_tick[order],
                       [rp.gpu_texts[d] for d in order], [rp.apps_texts[d] for d in order])
names2 = {1: "windows+rows", 2: "gpu rows", 3: "process rows", 4: "containers+lines", 5: "publish", 6: "slot wait",
          7: "compose", 8: "look-back+wait", 9: "copy-out"}
names = {1: "A work", 2: "A wait", 3: "B work", 4: "B wait", 5: "C work", 6: "C wait", 7: "busy entries+wait",
         8: "D lengths", 9: "segments+compose", 10: "look-back+wait", 11: "line_off+copy-out", 12: "last barrier"}
with monitor.Monitor(0) as mon:
    mon.load_fleet([entry_of_snap(s) for s in rp.nodes])
    batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
    mon.upload(batch)
    reps = 5
    for _ in range(reps):
        mon.encode_resident()
    buf = (C.c_uint64 * 64)()
    n = _ffi.load().kgm_debug_stage_clocks(buf)
    print("values", n, "kernel ms", mon.timing()["encode_ms"])
    docs = rp.n_docs * reps
    if mon.timing()["docs_in_flight_per_sm"] and os.environ.get("KGM_PIPELINE") != "fused-v1":
        tot = 0
        for st in range(1, 10):
            v = buf[st] / docs
            tot += v
            print("%-22s %8.0f clk/doc" % (names2[st], v))
        print("total %8.0f clk/doc  (%d documents in flight per SM)" % (tot, mon.timing()["docs_in_flight_per_sm"]))
        sys.exit(0)
    for half in (0, 1):
        tot = 0
        for st in range(1, 13):
            v = buf[half * 32 + st] / docs
            tot += v
            print("half %d  %-22s %8.0f clk/doc" % (half, names[st], v))
        print("half %d  total %8.0f clk/doc" % (half, tot))
