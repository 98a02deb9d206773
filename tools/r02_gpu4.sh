#!/bin/bash
# round-2 GPU call 4: full suite, default bench line, P/L sweep, ncu of the default kernel (single loop, elected refill)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest4.log
tail -12 gpurun_out/r02_pytest4.log
SWEEP_STRICT_ONLY=1 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2L;64,32,4,4,2X;64,16,2,4,3P" timeout 600 python tools/sweep_t2.py 512 400 > gpurun_out/r02_sweep4_burst.txt 2>&1
SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2L;64,32,4,4,2X;64,16,2,4,3P" timeout 900 python tools/sweep_t2.py 512 1500 > gpurun_out/r02_sweep4_sust.txt 2>&1
cat gpurun_out/r02_sweep4_burst.txt gpurun_out/r02_sweep4_sust.txt
timeout 900 python bench.py > gpurun_out/r02_bench4.json 2> gpurun_out/r02_bench4.err
cat gpurun_out/r02_bench4.json; tail -3 gpurun_out/r02_bench4.err
timeout 300 python tools/prof_t2.py 512 40 > gpurun_out/r02_plain4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 1 -o gpurun_out/r02_prof_t2_P4 -f python tools/prof_t2.py 512 40 > gpurun_out/r02_ncu_f4.log 2>&1
tail -2 gpurun_out/r02_ncu_f4.log
