#!/bin/bash
# round 2: GPU tests, rooflines + bench (tools/gpu_r2_check2.sh), then the ncu captures and the launch list of the bench command
bash tools/gpu_r2_check2.sh
bash tools/gpu_r2_ncu.sh
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context > gpurun_out/r2_bench_short.json 2> gpurun_out/r2_bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context > gpurun_out/r2_ncu_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r2_launches_bench.csv
du -sh gpurun_out
