# 1-GPU validation + evidence: GPU tests, smoke, bench line, ncu launch list of the bench command, all-configs report
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench1_full.json 2> gpurun_out/bench1_full.err; echo "bench rc=$?"; tail -c 900 gpurun_out/bench1_full.json
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench1_pre_ncu.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py gpurun_out/launches_bench.csv 12 > gpurun_out/launches_bench_summary.txt 2>&1; head -25 gpurun_out/launches_bench_summary.txt
timeout 900 python tools/config_report.py gpurun_out/configs_b200.json > gpurun_out/config_report.log 2>&1; echo "config_report rc=$?"; tail -3 gpurun_out/config_report.log | cut -c1-300
