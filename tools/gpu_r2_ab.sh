#!/bin/bash
# round 2, GPU call: protocol test with the G15 rule, A/B of tile / build variants on the bf16x2f step, ncu launch list, timeline
mkdir -p gpurun_out
python -m pytest tests/test_gpu_protocol.py -m gpu -q --timeout 1200 -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r2_protocol_pytest.log
tail -3 gpurun_out/r2_protocol_pytest.log
ab() { name=$1; shift; env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context --no-rooflines > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; python - gpurun_out/ab_$name.json $name <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('AB', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']))
PY
}
ab default M2F_NOP=1
ab narrow M2F_NARROW=1
ab cta2_shape M2F_GEMM_CTA2=2
ab cta2_0 M2F_GEMM_CTA2=0
[ -f multimodal-emotion-recognition_b200/build/variants/libm2f_b200_pair96.so ] && ab pair96 M2F_LIB_PATH=$PWD/multimodal-emotion-recognition_b200/build/variants/libm2f_b200_pair96.so
[ -f multimodal-emotion-recognition_b200/build/variants/libm2f_b200_pair96.so ] && ab pair96_narrow M2F_LIB_PATH=$PWD/multimodal-emotion-recognition_b200/build/variants/libm2f_b200_pair96.so M2F_NARROW=1
ab default_again M2F_NOP=1
# timeline of one step (per-op %globaltimer stamps)
M2F_TIMELINE=gpurun_out/r2_step_timeline.tsv timeout 300 python tools/step_loop.py 8 bf16x2f > gpurun_out/r2_step_loop.log 2>&1
tail -3 gpurun_out/r2_step_loop.log
# ncu launch list of a short bench (kernel shares)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context --no-rooflines > gpurun_out/r2_ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r2_launches.csv 30 > gpurun_out/r2_launches_summary.txt 2>&1
head -40 gpurun_out/r2_launches_summary.txt
