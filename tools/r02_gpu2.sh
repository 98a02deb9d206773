#!/bin/bash
# round-2 GPU call 2: parity suite on the trimmed depth-2 kernel, access-mode sweep, ncu of two modes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest2.log
tail -3 gpurun_out/r02_pytest2.log
SWEEP_STRICT_ONLY=1 timeout 600 python tools/sweep_t2.py 512 400 > gpurun_out/r02_sweep2_burst.txt 2>&1
SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 timeout 900 python tools/sweep_t2.py 512 1500 > gpurun_out/r02_sweep2_sust.txt 2>&1
cat gpurun_out/r02_sweep2_burst.txt gpurun_out/r02_sweep2_sust.txt
timeout 300 python tools/prof_t2.py 512 40 > gpurun_out/r02_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 1 -o gpurun_out/r02_prof_t2_P -f python tools/prof_t2.py 512 40 > gpurun_out/r02_ncu_f2.log 2>&1
HEAT3D_T2_TUNE=64,32,4,4,2,0,0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 1 -o gpurun_out/r02_prof_t2_L -f python tools/prof_t2.py 512 40 >> gpurun_out/r02_ncu_f2.log 2>&1
tail -2 gpurun_out/r02_ncu_f2.log
