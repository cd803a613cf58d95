#!/bin/bash
# Round-2 GPU call A (1 GPU): full GPU test tier, then the persistent kernel against the two-kernel driver.
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/r2a_clocks.csv &
SMI=$!
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
b() { name=$1; shift; timeout 300 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(cut -c1-400 gpurun_out/$name.json)"; }
b r2a_512_n1_persist   --steps 100 --warmup 10 --persistent on --trace
b r2a_512_n1_twokernel --steps 100 --warmup 10 --persistent off
b r2a_slab64_persist   --grid 512 512 64 --steps 200 --warmup 20 --persistent on --trace --no-e2e
b r2a_slab64_twokernel --grid 512 512 64 --steps 200 --warmup 20 --persistent off --no-e2e
b r2a_slab256_persist   --grid 256 256 32 --steps 400 --warmup 20 --persistent on --trace --no-e2e
b r2a_slab256_twokernel --grid 256 256 32 --steps 400 --warmup 20 --persistent off --no-e2e
b r2a_27_auto      --problem stencil3d27 --steps 100 --warmup 10 --no-e2e
b r2a_27_fused_tma --problem stencil3d27 --steps 100 --warmup 10 --variant fused --spmv tma --no-e2e
b r2a_27_recompute --problem stencil3d27 --steps 100 --warmup 10 --variant recompute --no-e2e
b r2a_27_merged_persist --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent on --trace --no-e2e
b r2a_27_merged_two     --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent off --no-e2e
b r2a_2d_persist   --problem poisson2d5 --steps 100 --warmup 10 --persistent on --no-e2e
b r2a_2d_twokernel --problem poisson2d5 --steps 100 --warmup 10 --persistent off --no-e2e
kill $SMI
