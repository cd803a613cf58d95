#!/bin/bash
export CGB_TIMEOUT_MS=300
mkdir -p gpurun_out
b() { name=$1; shift; timeout 120 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/$name.json')); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d.get('iteration_budget_us'), d['clocks']['sm_mhz'], d['clocks']['reasons'])
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-600:])
")"; }
b r2d_27_two     --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent off --no-e2e
b r2d_27_recompute --problem stencil3d27 --steps 100 --warmup 10 --variant recompute --no-e2e
b r2d_27_fused_tma --problem stencil3d27 --steps 100 --warmup 10 --variant fused --spmv tma --no-e2e
b r2d_27_auto --problem stencil3d27 --steps 100 --warmup 10 --no-e2e
b r2d_27_persist --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent on --no-e2e
b r2d_2d_two --problem poisson2d5 --steps 100 --warmup 10 --persistent off --no-e2e
b r2d_512_two --steps 100 --warmup 10 --persistent off --no-e2e
b r2d_slab64_two --grid 512 512 64 --steps 200 --warmup 20 --persistent off --no-e2e
b r2d_slab64_persist --grid 512 512 64 --steps 200 --warmup 20 --persistent on --trace --no-e2e
