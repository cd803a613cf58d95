#!/bin/bash
# Round-2 (1 GPU): leaner plane loops (ring position without a division, no second row on 2D grids, opaque tap offsets,
# idle threads shadow thread 0). Tests first, then the named configurations and their 8-GPU shares.
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
run() { name=$1; shift; timeout 120 "$@" > gpurun_out/r2l_$name.json 2> gpurun_out/r2l_$name.err; echo "$name rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2l_$name.json | head -1)"; }
timeout 600 python -m pytest tests/test_gpu_solver.py tests/test_gpu_ops.py -x -q -m gpu > gpurun_out/r2l_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2l_tests.log
B="python bench.py --no-e2e --warmup 10"
run 512       $B --grid 512 --steps 100
run 2d        $B --problem poisson2d5 --grid 8192 --steps 100
run 27        $B --problem stencil3d27 --grid 384 --steps 100
run slab      $B --grid 512 512 64 --steps 400
run 2dslab    $B --problem poisson2d5 --grid 8192 1024 --steps 400
run 27slab    $B --problem stencil3d27 --grid 384 384 48 --steps 400
run 256       $B --grid 256 --steps 400
run 1024      $B --grid 1024 --steps 25
run 27pipe    $B --problem stencil3d27 --grid 384 --steps 100 --variant pipelined
run 512fused  $B --grid 512 --steps 100 --variant fused
run 512persist $B --grid 512 --steps 100 --persistent on
