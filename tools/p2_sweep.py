"""Times the poll kernel on the configs[1] batch for a few (jobs per line, warps per SM) settings."""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def run(env):
    e = dict(os.environ); e.update(env)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "20", "--warmup", "5", "--no-cpu"],
                         env=e, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    d = json.loads(out.strip().splitlines()[-1])
    return d["roofline"]["launch_ms"], d["ms_per_step"], d.get("docs_in_flight_per_sm")
for env in ({}, {"KGM_POLL2_JOBS": "8"}, {"KGM_POLL2_WARPS": "12"}, {"KGM_POLL2_WARPS": "8"}, {"KGM_POLL2_JOBS": "8", "KGM_POLL2_WARPS": "16"},
            {"KGM_PIPELINE": "fused-v1"}):
    print(env, run(env), flush=True)
