set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -x -q -m gpu 2>&1 | tail -2
bash tools/gpu_benchN.sh 2
