#!/bin/bash
# Development loop on the GPU box: quick parity on the default kernel, a short bench, optionally an
# ncu capture of the poll kernel (NCU=1).  Everything lands in gpurun_out/<tag>*.
tag=${1:-x}
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "fused and not generic and not v1" 2>&1 | tail -4
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/$tag.json 2> gpurun_out/$tag.err; echo bench rc=$?
python - <<PY
import json
d=json.load(open("gpurun_out/$tag.json"))
print("value %.1fM  step %.4f ms  k_poll alone %.4f ms  frac %.4f  slow %d  e2e %.3f ms" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["launch_ms"], d["roofline"]["frac"], d["slow_docs_per_step"], d["e2e"]["ms_per_step"]))
PY
if [ -n "$NCU" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_poll -s 8 -c 1 -o gpurun_out/$tag python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/$tag.ncu.log 2>&1; echo ncu rc=$?
fi
