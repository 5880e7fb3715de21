#!/bin/bash
# tools/n_config.sh N config steps -> bench line under gpurun_out/r2_bench_<config>_<N>gpu.json
N=$1; C=$2; STEPS=${3:-10}
out=gpurun_out/r2_bench_${C}_${N}gpu.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --config $C --steps $STEPS --warmup 3 --no-cpu > $out 2> ${out%.json}.err || tail -5 ${out%.json}.err
python - <<PY
import json
d=json.loads(open("$out").read().strip().splitlines()[-1])
print("N=$N $C: %.1fM samples/s  step %.4f ms  mode %s verified %s slow %s shard %s" % (d["value"]/1e6, d["ms_per_step"], d.get("gather_mode"), d.get("concat_verified"), d["slow_docs_per_step"], d["config"]["shard"]))
PY
