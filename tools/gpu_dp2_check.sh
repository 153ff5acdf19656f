set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -x -q -m gpu -s > gpurun_out/r1b_dp_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1b_dp_tests.log
tail -25 gpurun_out/r1b_dp_tests.log
run() { name=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e "$@" > gpurun_out/bench2_$name.json 2> gpurun_out/bench2_$name.err; echo "rc=$?"; tail -c 1500 gpurun_out/bench2_$name.json; tail -5 gpurun_out/bench2_$name.err; }
run peer32 --exchange peer
run peer16 --exchange peer --bucket-mb 16
run nccl128 --exchange nccl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --steps 20 --warmup 5 --exchange peer > gpurun_out/bench2_peer_full.json 2> gpurun_out/bench2_peer_full.err; tail -c 600 gpurun_out/bench2_peer_full.json
