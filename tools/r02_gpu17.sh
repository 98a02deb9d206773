#!/bin/bash
# round-2 GPU call 17 (1 GPU): final suite + default bench line at the final code
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -rs --timeout 200 > gpurun_out/r02_pytest_1gpu_final.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_1gpu_final.log; tail -3 gpurun_out/r02_pytest_1gpu_final.log
timeout 400 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_final.json"))
print("value",round(d["value"],1),"frac",round(d["roofline"]["frac"],3),"k1",round(d["roofline_k1"]["glups"],1),round(d["roofline_k1"]["frac"],3),"fma",round(d["arith_fma"]["glups"],1),
      "e2e",round(d["e2e"]["value"],1),round(d["e2e"]["blocking"]["value"],1),"launches",d["gpu_launches"],d["gpu_launches_other_blocks"],"clocks",d["clocks"]["sm_mhz"],d["clocks"]["reasons"],"cpu",round(d["cpu_baseline"]["value"],4))
PY
