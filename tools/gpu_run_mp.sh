#!/bin/bash
# Round-2 multi-GPU call: bash tools/gpu_run_mp.sh N [tests: all|wide|none]
N=${1:-2}; TESTS=${2:-all}
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
if [ "$TESTS" = all ]; then
  timeout 1500 python -m pytest tests -m multigpu -x -q > gpurun_out/r2m_pytest_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest_n$N.log
elif [ "$TESTS" = wide ]; then
  timeout 900 python -m pytest tests/test_multiprocess.py -m multigpu -x -q -k "all_gpus" > gpurun_out/r2m_pytest_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest_n$N.log
fi
[ "$TESTS" != none ] && tail -4 gpurun_out/r2m_pytest_n$N.log
PORT=29900
b() { name=$1; shift; PORT=$((PORT+1)); timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.loads([l for l in open('gpurun_out/$name.json') if l.startswith('{')][-1]); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'rnorm', d.get('rnorm_after_steps'), 'e2e', d.get('e2e') and round(d['e2e']['value'],1)); b=d.get('iteration_budget_us'); print('   budget', b) if b else None
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-700:])
")"; }
b r2m_512_n${N}_two     --steps 200 --warmup 20 --persistent off
b r2m_512_n${N}_persist --steps 200 --warmup 20 --persistent on --trace
b r2m_256_n${N}_two     --grid 256 --steps 400 --warmup 20 --persistent off --no-e2e
b r2m_256_n${N}_persist --grid 256 --steps 400 --warmup 20 --persistent on --trace --no-e2e
b r2m_512_n${N}_b4      --steps 50 --warmup 5 --impl refstyle
b r2m_256_n${N}_b4      --grid 256 --steps 50 --warmup 5 --impl refstyle
