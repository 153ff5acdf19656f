set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -2
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
bash tools/gpu_benchN.sh 2 --no-e2e
