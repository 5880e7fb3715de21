#!/bin/bash
# tools/n_sweep.sh N "ENV=.. ENV2=.." steps   -> one torchrun bench line, summarised
N=$1; ENVS=$2; STEPS=${3:-20}
env $ENVS timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps $STEPS --warmup 5 --no-cpu > gpurun_out/ns.json 2> gpurun_out/ns.err || tail -5 gpurun_out/ns.err
python - <<PY
import json
d=json.loads(open("gpurun_out/ns.json").read().strip().splitlines()[-1])
print("N=$N [$ENVS] steps=$STEPS: %.1fM samples/s  step %.4f ms  mode %s verified %s  host %s  e2e %.3f ms" % (d["value"]/1e6, d["ms_per_step"], d.get("gather_mode"), d.get("concat_verified"), d["host_ms_per_step_rank0"], d["e2e"]["ms_per_step"]))
PY
