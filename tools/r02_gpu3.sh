#!/bin/bash
# round-2 GPU call 3: full suite (new driver + snapshot tests), ncu of the default depth-2 kernel, power/clock comparison
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest3.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest3.log
tail -15 gpurun_out/r02_pytest3.log
for cfg in "P:64,32,4,4,2,0,2:strict" "L:64,32,4,4,2,0,0:strict" "P:64,32,4,4,2,0,2:fma" "r2:64,16,2,4,3,0,2:strict"; do
  IFS=: read name tune arith <<< "$cfg"
  HEAT3D_T2_TUNE=$tune timeout 300 python bench.py --steps 4 --warmup 3 --arith $arith --no-cpu-baseline --no-e2e --no-k1 > gpurun_out/r02_clk_${name}_${arith}.json 2>gpurun_out/r02_clk.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02_clk_${name}_${arith}.json"))
print("${name} ${arith}", round(d["value"],1), d["clocks"])
PY
done
SWEEP_SHAPE=1024,1024,512 SWEEP_STRICT_ONLY=1 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2L" timeout 600 python tools/sweep_t2.py 512 200 > gpurun_out/r02_sweep3_1024.txt 2>&1
cat gpurun_out/r02_sweep3_1024.txt
timeout 300 python tools/prof_t2.py 512 40 > gpurun_out/r02_plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stencil_tma_t2 -s 10 -c 1 -o gpurun_out/r02_prof_t2_P -f python tools/prof_t2.py 512 40 > gpurun_out/r02_ncu_f3.log 2>&1
tail -2 gpurun_out/r02_ncu_f3.log
ls -la gpurun_out | tail -5
