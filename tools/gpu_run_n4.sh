#!/bin/bash
N=${1:-4}
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
PORT=29970
b() { name=$1; shift; PORT=$((PORT+1)); timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.loads([l for l in open('gpurun_out/$name.json') if l.startswith('{')][-1]); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'rnorm', d.get('rnorm_after_steps'), 'e2e', d.get('e2e') and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-500:])
")"; }
b r2n_512_n${N}_two  --steps 200 --warmup 20
b r2n_256_n${N}_two  --grid 256 --steps 400 --warmup 20 --no-e2e
b r2n_1024_n${N}_two --grid 1024 --steps 50 --warmup 5 --no-e2e
b r2n_weak_n${N}     --grid 512 512 $((64*N)) --steps 200 --warmup 20 --no-e2e
b r2n_2d_n${N}       --problem poisson2d5 --steps 200 --warmup 20 --no-e2e
b r2n_27_n${N}       --problem stencil3d27 --steps 200 --warmup 20 --no-e2e
b r2n_512_n${N}_b4   --steps 50 --warmup 5 --impl refstyle
b r2n_256_n${N}_b4   --grid 256 --steps 50 --warmup 5 --impl refstyle
