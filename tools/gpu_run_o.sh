#!/bin/bash
# Round-2 late multi-GPU check after the loop/geometry changes: bash tools/gpu_run_o.sh N tests|bench
N=${1:-2}; WHAT=${2:-tests}
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
if [ "$WHAT" = tests ]; then
  timeout 1200 python -m pytest tests -m multigpu -x -q > gpurun_out/r2o_pytest_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2o_pytest_n$N.log
  tail -4 gpurun_out/r2o_pytest_n$N.log
  exit 0
fi
PORT=29700
b() { name=$1; shift; PORT=$((PORT+1)); timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.loads([l for l in open('gpurun_out/$name.json') if l.startswith('{')][-1]); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'rnorm', d.get('rnorm_after_steps'), 'e2e', d.get('e2e') and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-700:])
")"; }
b r2o_512_n${N}   --steps 200 --warmup 20 --no-e2e
b r2o_27_n${N}    --problem stencil3d27 --steps 200 --warmup 20 --no-e2e
b r2o_2d_n${N}    --problem poisson2d5 --steps 200 --warmup 20 --no-e2e
b r2o_1024_n${N}  --grid 1024 --steps 50 --warmup 5 --no-e2e
b r2o_256_n${N}   --grid 256 --steps 400 --warmup 20 --no-e2e
