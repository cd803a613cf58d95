#!/bin/bash
# Round-2 GPU call C (1 GPU): PDL on the two-kernel path, persistent kernel v2 with register-resident service.
export CGB_TIMEOUT_MS=300
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_solver.py tests/test_gpu_ops.py -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -4 gpurun_out/r2c_pytest.log
b() { name=$1; shift; timeout 120 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/$name.json')); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d.get('iteration_budget_us'), d['clocks']['sm_mhz'], d['clocks']['reasons'], 'e2e', d['e2e'] and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-600:])
")"; }
for g in "512 512 512" "512 512 64" "256 256 32"; do
  n=$(echo $g | tr ' ' x)
  b r2c_${n}_persist --grid $g --steps 200 --warmup 20 --persistent on --trace --no-e2e
  b r2c_${n}_pdl     --grid $g --steps 200 --warmup 20 --persistent off --no-e2e
  CGB_PDL=0 b r2c_${n}_nopdl --grid $g --steps 200 --warmup 20 --persistent off --no-e2e
done
b r2c_27_persist --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent on --trace --no-e2e
b r2c_27_pdl     --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent off --no-e2e
CGB_PDL=0 b r2c_27_nopdl --problem stencil3d27 --steps 100 --warmup 10 --variant merged --persistent off --no-e2e
b r2c_27_recompute --problem stencil3d27 --steps 100 --warmup 10 --variant recompute --no-e2e
b r2c_2d_persist --problem poisson2d5 --steps 100 --warmup 10 --persistent on --trace --no-e2e
b r2c_2d_pdl     --problem poisson2d5 --steps 100 --warmup 10 --persistent off --no-e2e
CGB_PDL=0 b r2c_2d_nopdl --problem poisson2d5 --steps 100 --warmup 10 --persistent off --no-e2e
