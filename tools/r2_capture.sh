#!/bin/bash
# Round-2 evidence run on one B200: tests, bench lines, launch list, ncu captures (each ncu pass only
# after the same command has exited 0 without ncu).  Output under gpurun_out/r2f_*.
set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/r2f_gputests.txt
timeout 600 python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo bench rc=$?
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2f_bench_reference_arm.json 2> gpurun_out/r2f_ref.err; echo ref rc=$?
timeout 300 python bench.py --pipeline fused-v1 --steps 20 --warmup 5 --no-cpu > gpurun_out/r2f_bench_config2_team_kernel.json 2>/dev/null
timeout 300 python bench.py --config config3 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2f_bench_config3.json 2>/dev/null
timeout 300 python bench.py --config config3 --pipeline fused-v1 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2f_bench_config3_team_kernel.json 2>/dev/null
timeout 300 python bench.py --config config4 --steps 20 --warmup 3 --no-cpu > gpurun_out/r2f_bench_config4.json 2>/dev/null
timeout 300 python bench.py --config config5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2f_bench_config5.json 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2f_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu > /dev/null 2>&1; echo launches rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_poll2 -s 8 -c 1 -o gpurun_out/r2f_poll2 python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1; echo ncu2 rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_poll -s 8 -c 1 -o gpurun_out/r2f_poll python bench.py --pipeline fused-v1 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1; echo ncu1 rc=$?
timeout 600 ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:k_poll2 -s 8 -c 8 --csv --log-file gpurun_out/r2f_poll2_traffic.csv python bench.py --steps 8 --warmup 3 --no-cpu > /dev/null 2>&1; echo traffic rc=$?
KGM_PIPELINE=fused-v2 python tools/stage_clocks.py > gpurun_out/r2f_stage_clocks_note.txt 2>&1 || true
