#!/bin/bash
# Round-2 final validation (1 GPU): what the driver runs - GPU test tier, smoke, default bench arms
export CGB_TIMEOUT_MS=3000
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
tail -3 gpurun_out/r2g_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5
python bench.py --gpus 1 --steps 20 --warmup 5 | tee gpurun_out/r2g_bench_driverlike_n1.json | cut -c1-900
python bench.py | tee gpurun_out/r2g_bench_default_n1.json | cut -c1-300
