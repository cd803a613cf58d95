#!/bin/bash
# ncu capture of the 2D band kernels after the wide-row change (1 GPU, 8192 x 1024 slab = the 8-GPU share of 8192^2).
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
C1="python bench.py --problem poisson2d5 --grid 8192 1024 --steps 3 --warmup 3 --no-e2e --persistent off"
$C1 > gpurun_out/r2k_plain1.log 2>&1 && $NCU -k regex:tma -s 4 -c 2 -f -o gpurun_out/r2k_merged2d_slab $C1 > gpurun_out/r2k_ncu1.log 2>&1; echo "c1 rc=$?"
ls -la gpurun_out/*.ncu-rep
