#!/bin/bash
# Round-2 GPU call E (1 GPU): full GPU test tier + CSR like-for-like race
export CGB_TIMEOUT_MS=2000
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
tail -4 gpurun_out/r2e_pytest.log
b() { name=$1; to=$2; shift; shift; timeout $to python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/$name.json')); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'rnorm', d.get('rnorm_after_steps'), 'e2e', d.get('e2e') and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-600:])
")"; }
b r2e_csr7_256_ours 200 --problem csr7 --grid 256 --steps 50 --warmup 5 --no-e2e
b r2e_csr7_256_ref  200 --problem csr7 --grid 256 --steps 50 --warmup 5 --impl refstyle
b r2e_csr7_512_ours 500 --problem csr7 --steps 50 --warmup 5 --no-e2e
b r2e_csr7_512_ref  300 --problem csr7 --steps 50 --warmup 5 --impl refstyle
b r2e_512_mf 200 --steps 50 --warmup 5 --no-e2e
./bin/refstyle_b4 --kind 2 --grid 256 256 256 --gpus 1 --steps 50 --warmup 5 > gpurun_out/r2e_b4_n1_256.json 2> gpurun_out/r2e_b4_n1_256.err; echo "b4 rc=$? $(cat gpurun_out/r2e_b4_n1_256.json)"
