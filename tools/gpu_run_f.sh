#!/bin/bash
# Round-2 GPU call F (1 GPU): sweep points at N=1, B3 baselines, B0, per-kernel DRAM traffic of the CSR paths
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
b() { name=$1; shift; timeout 300 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(python -c "
import json,sys
try:
    d=json.loads([l for l in open('gpurun_out/$name.json') if l.startswith('{')][-1]); print(round(d['value'],1),'it/s', round(d['ms_per_step']*1000,2),'us', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'rnorm', d.get('rnorm_after_steps'), 'e2e', d.get('e2e') and round(d['e2e']['value'],1))
except Exception as e: print('ERR', e, open('gpurun_out/$name.err').read()[-500:])
")"; }
b r2f_256_n1   --grid 256 --steps 400 --warmup 20 --no-e2e
b r2f_1024_n1  --grid 1024 --steps 30 --warmup 5 --no-e2e
b r2f_weak64_n1 --grid 512 512 64 --steps 200 --warmup 20 --no-e2e
b r2f_512_n1   --steps 200 --warmup 20
b r2f_2d_n1    --problem poisson2d5 --steps 100 --warmup 10 --no-e2e
b r2f_27_n1    --problem stencil3d27 --steps 100 --warmup 10 --no-e2e
b r2f_27pipe_n1 --problem stencil3d27 --variant pipelined --steps 100 --warmup 10 --no-e2e
b r2f_256_n1_b3 --grid 256 --steps 50 --warmup 5 --impl refstyle
b r2f_512_n1_b3 --steps 50 --warmup 5 --impl refstyle
b r2f_2d_n1_b3 --problem poisson2d5 --steps 30 --warmup 5 --impl refstyle
b r2f_27_n1_b3 --problem stencil3d27 --steps 30 --warmup 5 --impl refstyle
# B0: the unmodified reference binary vs our CLI on the 3x3 system (wall time of the whole process)
if [ -x baseline/_ref/ref_cudacg ]; then
  python - > gpurun_out/r2f_b0.txt <<'PY'
import subprocess, time
for exe in ("baseline/_ref/ref_cudacg", "bin/cgb200"):
    ts = []
    for i in range(4):
        t0 = time.perf_counter(); r = subprocess.run([exe], capture_output=True, text=True); ts.append(time.perf_counter() - t0)
    print(exe, "rc", r.returncode, "stdout", repr(r.stdout), "wall_s", [round(t, 3) for t in ts])
PY
  cat gpurun_out/r2f_b0.txt
fi
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
CMD_OURS="python bench.py --problem csr7 --grid 256 --steps 4 --warmup 3 --no-e2e"
CMD_REF="python bench.py --problem csr7 --grid 256 --steps 4 --warmup 3 --impl refstyle"
$CMD_OURS > gpurun_out/r2f_plain_ours.log 2>&1 && ncu --metrics $M --clock-control none -c 40 --csv --log-file gpurun_out/r2f_csr_ours_launches.csv $CMD_OURS > gpurun_out/r2f_ncu_ours.log 2>&1
$CMD_REF > gpurun_out/r2f_plain_ref.log 2>&1 && ncu --metrics $M --clock-control none -c 120 --csv --log-file gpurun_out/r2f_csr_ref_launches.csv $CMD_REF > gpurun_out/r2f_ncu_ref.log 2>&1
ls -la gpurun_out/r2f_csr_*_launches.csv
