#!/bin/bash
# round 2: ncu --set full of the kernel classes the rooflines quote (plain run first, as the profiling recipe requires)
mkdir -p gpurun_out
python tools/kernel_loop.py 2 > gpurun_out/r2_kernel_loop_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_kernel|ln_fwd_staged|ln_bwd_staged|attn_fwd|attn_bwd' -c 16 -f -o gpurun_out/r2_kernels python tools/kernel_loop.py 2 > gpurun_out/r2_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/r2_ncu_full.log; ls -la gpurun_out/r2_kernels.ncu-rep
