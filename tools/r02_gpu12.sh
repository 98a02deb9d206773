#!/bin/bash
# round-2 GPU call 12 (2 GPUs): bench.py after its last edits, N=1 and N=2, exactly as the driver launches it
mkdir -p gpurun_out; rm -f gpurun_out/scale_r02_check.jsonl gpurun_out/scale.err
timeout 400 bash tools/scale_run.sh 1 "--steps 5 --warmup 3" r02_check > /dev/null
timeout 500 bash tools/scale_run.sh 2 "--steps 5 --warmup 3" r02_check > /dev/null
python - <<'PY'
import json
for l in open("gpurun_out/scale_r02_check.jsonl"):
    if not l.startswith("{"): print("non-json stdout line:", l[:80].strip()); continue
    d=json.loads(l)
    print(d["n_gpus"],"value",round(d["value"],1),"frac",round(d["roofline"]["frac"],3),"traffic",d["roofline"]["traffic"],"k1",d.get("roofline_k1") and round(d["roofline_k1"]["frac"],3),
          "fma",d.get("arith_fma") and round(d["arith_fma"]["glups"],1),"e2e",round(d["e2e"]["value"],1),"pre",d.get("parity_preflight") and d["parity_preflight"]["bitwise"],
          "eff",d.get("parallel_efficiency"),"strong",d.get("strong_1024") and round(d["strong_1024"]["efficiency"],3),"launches",d["gpu_launches"],"clocks",d["clocks"]["sm_mhz"],d["clocks"]["reasons"])
PY
tail -3 gpurun_out/scale.err
