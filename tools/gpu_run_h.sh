#!/bin/bash
# Round-2 ncu captures (1 GPU): the band kernels after the partial-sum rewrite (7-point and 27-point slabs) and the
# streaming CSR SpMV. Each command first runs plain and must exit 0.
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
C1="python bench.py --grid 512 512 64 --steps 3 --warmup 3 --no-e2e --persistent off"
C2="python bench.py --problem stencil3d27 --grid 384 384 48 --steps 3 --warmup 3 --no-e2e --persistent off"
C3="python bench.py --problem csr7 --grid 256 --steps 3 --warmup 3 --no-e2e"
$C1 > gpurun_out/r2h_plain1.log 2>&1 && $NCU -k regex:tma -s 4 -c 2 -f -o gpurun_out/r2h_merged7_slab $C1 > gpurun_out/r2h_ncu1.log 2>&1; echo "c1 rc=$?"
$C2 > gpurun_out/r2h_plain2.log 2>&1 && $NCU -k regex:tma -s 4 -c 2 -f -o gpurun_out/r2h_merged27_slab $C2 > gpurun_out/r2h_ncu2.log 2>&1; echo "c2 rc=$?"
$C3 > gpurun_out/r2h_plain3.log 2>&1 && $NCU -k regex:k_spmv_csr_stream -s 2 -c 1 -f -o gpurun_out/r2h_csr_stream $C3 > gpurun_out/r2h_ncu3.log 2>&1; echo "c3 rc=$?"
ls -la gpurun_out/*.ncu-rep
