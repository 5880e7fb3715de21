"""Times the poll kernels on a bench config for a few settings (development aid).
    python tools/p2_sweep.py "({'KGM_POLL2_LINES': '32'}, ['--config', 'config3'])" ..."""
import ast, os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def run(env, args):
    e = dict(os.environ); e.update(env)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "12", "--warmup", "3", "--no-cpu"] + args,
                         env=e, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    d = json.loads(out.strip().splitlines()[-1])
    return "alone %.4f ms  step %.4f ms  in flight %s  frac %.4f" % (d["roofline"]["launch_ms"], d["ms_per_step"], d.get("docs_in_flight_per_sm"), d["roofline"]["frac"])
for spec in sys.argv[1:]:
    env, args = ast.literal_eval(spec)
    print(env, args, run(env, args), flush=True)
