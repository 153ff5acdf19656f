#!/bin/bash
# round 2 evidence: train-mode parity file, ncu --set full of the roofline kernel classes, compute-sanitizer over the op tests,
# ncu launch list of the bench command
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dropout_parity.py -m gpu -q -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -40 > gpurun_out/r2_dropout_parity.log
tail -3 gpurun_out/r2_dropout_parity.log
bash tools/gpu_r2_ncu.sh
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_ops.py tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider -k "not gemm_shapes or 264" > gpurun_out/r2_sanitizer_memcheck_ops.log 2>&1
echo "memcheck rc=$?"; tail -6 gpurun_out/r2_sanitizer_memcheck_ops.log
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context > gpurun_out/r2_bench_short.json 2> gpurun_out/r2_bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context > gpurun_out/r2_ncu_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r2_launches_bench.csv
