#!/bin/bash
# Round-2 GPU call B (1 GPU): persistent kernel v2 (service CTA, split barrier, lazy scalars) vs two kernels per iteration.
export CGB_TIMEOUT_MS=300
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_solver.py -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
b() { name=$1; shift; timeout 90 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/$name.json')); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d.get('iteration_budget_us'), d['clocks']['sm_mhz'], d['clocks']['reasons'], 'e2e', d['e2e'] and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-800:])
")"; }
b r2b_512_n1_persist   --steps 100 --warmup 10 --persistent on --trace
b r2b_512_n1_twokernel --steps 100 --warmup 10 --persistent off
b r2b_slab64_persist   --grid 512 512 64 --steps 200 --warmup 20 --persistent on --trace --no-e2e
b r2b_slab64_twokernel --grid 512 512 64 --steps 200 --warmup 20 --persistent off --no-e2e
b r2b_slab256_persist   --grid 256 256 32 --steps 400 --warmup 20 --persistent on --trace --no-e2e
b r2b_slab256_twokernel --grid 256 256 32 --steps 400 --warmup 20 --persistent off --no-e2e
b r2b_27_merged_persist --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent on --trace --no-e2e
b r2b_27_merged_two     --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent off --no-e2e
b r2b_2d_persist   --problem poisson2d5 --steps 100 --warmup 10 --persistent on --trace --no-e2e
b r2b_2d_twokernel --problem poisson2d5 --steps 100 --warmup 10 --persistent off --no-e2e
