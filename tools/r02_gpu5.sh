#!/bin/bash
# round-2 GPU call 5: bench alone under a watchdog (it overstayed in call 4), then the full suite with per-test timeouts
mkdir -p gpurun_out
HEAT3D_BENCH_WATCHDOG_S=200 timeout 260 python bench.py > gpurun_out/r02_bench5.json 2> gpurun_out/r02_bench5.err
echo "bench rc=$?"; cat gpurun_out/r02_bench5.json; tail -40 gpurun_out/r02_bench5.err
timeout 900 python -m pytest tests -m gpu -x -q --timeout 240 > gpurun_out/r02_pytest5.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest5.log
tail -15 gpurun_out/r02_pytest5.log
