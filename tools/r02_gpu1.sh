#!/bin/bash
# round-2 GPU call 1: parity suite, depth-2 sweeps, default bench, launch list + full ncu capture of the top kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest1.log
tail -3 gpurun_out/r02_pytest1.log
SWEEP_STRICT_ONLY=1 timeout 600 python tools/sweep_t2.py 512 400 > gpurun_out/r02_sweep_burst.txt 2>&1
SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 timeout 900 python tools/sweep_t2.py 512 1500 > gpurun_out/r02_sweep_sust.txt 2>&1
cat gpurun_out/r02_sweep_burst.txt gpurun_out/r02_sweep_sust.txt
timeout 600 python bench.py > gpurun_out/r02_bench1.json 2> gpurun_out/r02_bench1.err
cat gpurun_out/r02_bench1.json
timeout 300 python tools/prof_t2.py 512 40 > gpurun_out/r02_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 2 -o gpurun_out/r02_prof_t2 -f python tools/prof_t2.py 512 40 > gpurun_out/r02_ncu_f.log 2>&1
tail -2 gpurun_out/r02_ncu_f.log
