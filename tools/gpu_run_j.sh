#!/bin/bash
# Round-2 experiment (1 GPU): 2D grids with one TMA operation per row tile (wide rows) vs one per 256-cell box.
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
run() { name=$1; shift; timeout 120 "$@" > gpurun_out/r2j_$name.json 2> gpurun_out/r2j_$name.err; echo "$name rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2j_$name.json | head -1)"; }
B="python bench.py --no-e2e --warmup 10 --problem poisson2d5"
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k "poisson2d" > gpurun_out/r2j_tests_ops.log 2>&1; echo "ops tests rc=$?"; tail -3 gpurun_out/r2j_tests_ops.log
run 2d_wide   $B --grid 8192 --steps 100
CGB_TMA_NARROW=1 run 2d_narrow $B --grid 8192 --steps 100
run 2dslab_wide   $B --grid 8192 1024 --steps 400
CGB_TMA_NARROW=1 run 2dslab_narrow $B --grid 8192 1024 --steps 400
run 2d_wide_fused   $B --grid 8192 --steps 100 --variant fused
CGB_TMA_NARROW=1 run 2d_narrow_fused $B --grid 8192 --steps 100 --variant fused
timeout 600 python -m pytest tests/test_gpu_solver.py tests/test_gpu_ops.py -x -q -m gpu > gpurun_out/r2j_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2j_tests.log
