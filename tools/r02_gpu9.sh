#!/bin/bash
# round-2 GPU call 9 (1 GPU): final suite, FP64 issue-rate probe, default bench line, launch list and full ncu capture of the default kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -rs --timeout 300 > gpurun_out/r02_pytest_1gpu_final.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_1gpu_final.log
tail -6 gpurun_out/r02_pytest_1gpu_final.log
timeout 120 tools/fp64_peak > gpurun_out/r02_fp64_peak.txt 2>&1; cat gpurun_out/r02_fp64_peak.txt
timeout 400 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/r02_bench_final.json
CMD="python bench.py --steps 2 --warmup 3 --tsteps 20 --no-cpu-baseline --no-e2e --no-k1"
timeout 200 $CMD > gpurun_out/r02_plain_l.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_steps2_tsteps20.csv $CMD > gpurun_out/r02_ncu_l.log 2>&1
timeout 200 python tools/prof_t2.py 512 40 > gpurun_out/r02_plain9.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 1 -o gpurun_out/r02_prof_t2_final -f python tools/prof_t2.py 512 40 > gpurun_out/r02_ncu_f9.log 2>&1
tail -2 gpurun_out/r02_ncu_f9.log
