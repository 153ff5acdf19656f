#!/bin/bash
# round 2: ncu --set full of the kernel classes the rooflines quote (plain run first, as the profiling recipe requires).  The report is
# turned into CSV pages on the box (the .ncu-rep exceeds what gpurun copies back) and removed.
mkdir -p gpurun_out
python tools/kernel_loop.py 1 > gpurun_out/r2_kernel_loop_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_kernel|gemm_persist|ln_fwd|ln_bwd|attn_fwd|attn_bwd' -c 13 -f -o /tmp/r2_kernels python tools/kernel_loop.py 1 > gpurun_out/r2_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2_ncu_full.log
ncu -i /tmp/r2_kernels.ncu-rep --page raw --csv > gpurun_out/r2_kernels_raw.csv 2> gpurun_out/r2_ncu_export.err
ncu -i /tmp/r2_kernels.ncu-rep --page details --csv > gpurun_out/r2_kernels_details.csv 2>> gpurun_out/r2_ncu_export.err
ls -la /tmp/r2_kernels.ncu-rep gpurun_out/r2_kernels_raw.csv gpurun_out/r2_kernels_details.csv
