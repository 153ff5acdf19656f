#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_exchange.py tests/test_gpu_model.py -m gpu -q --timeout 900 -p no:cacheprovider 2>&1 | tail -40 > gpurun_out/r2_tail_pytest.log
tail -40 gpurun_out/r2_tail_pytest.log | grep -E "^E|passed|failed|FAILED|Error" | head -30
