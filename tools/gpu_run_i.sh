#!/bin/bash
# Round-2 z-split experiment (1 GPU): multi-round item decomposition chosen by the cost model in make_geo.
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
run() { name=$1; shift; timeout 120 "$@" > gpurun_out/r2i_$name.json 2> gpurun_out/r2i_$name.err; echo "$name rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2i_$name.json | head -1)"; }
B="python bench.py --no-e2e --warmup 10"
run 27_auto   $B --problem stencil3d27 --grid 384 --steps 100
CGB_ZSPLIT=1 run 27_zs1 $B --problem stencil3d27 --grid 384 --steps 100
CGB_ZSPLIT=6 run 27_zs6 $B --problem stencil3d27 --grid 384 --steps 100
run 27slab_auto $B --problem stencil3d27 --grid 384 384 48 --steps 400
CGB_ZSPLIT=1 run 27slab_zs1 $B --problem stencil3d27 --grid 384 384 48 --steps 400
run 512_auto  $B --grid 512 --steps 100
CGB_ZSPLIT=8 run 512_zs8 $B --grid 512 --steps 100
CGB_ZSPLIT=2 run 512_zs2 $B --grid 512 --steps 100
run 1024_auto $B --grid 1024 --steps 25
CGB_ZSPLIT=1 run 1024_zs1 $B --grid 1024 --steps 25
CGB_ZSPLIT=8 run 1024_zs8 $B --grid 1024 --steps 25
run 384_auto $B --grid 384 --steps 200
CGB_ZSPLIT=1 run 384_zs1 $B --grid 384 --steps 200
run slab_auto $B --grid 512 512 64 --steps 400
CGB_ZSPLIT=8 run slab_zs8 $B --grid 512 512 64 --steps 400
CGB_ZSPLIT=4 run slab_zs4 $B --grid 512 512 64 --steps 400
timeout 600 python -m pytest tests/test_gpu_solver.py tests/test_gpu_ops.py -x -q -m gpu > gpurun_out/r2i_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2i_tests.log
