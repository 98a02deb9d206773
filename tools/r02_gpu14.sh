#!/bin/bash
# round-2 GPU call 14 (8 GPUs): the default bench line on the final bench.py
mkdir -p gpurun_out; rm -f gpurun_out/scale_r02_n8b.jsonl gpurun_out/scale.err
ls /sys/devices/system/node | grep node | tr '\n' ' '; nproc
timeout 500 bash tools/scale_run.sh 8 "--steps 5 --warmup 3" r02_n8b > /dev/null
python - <<'PY'
import json
for l in open("gpurun_out/scale_r02_n8b.jsonl"):
    if not l.startswith("{"): continue
    d=json.loads(l)
    print(d["n_gpus"],"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),round(d["e2e"]["blocking"]["value"],1),d["e2e"]["cpu_affinity"],"pre",d["parity_preflight"]["bitwise"],
          "eff",d["parallel_efficiency"],"strong",round(d["strong_1024"]["value"],1),round(d["strong_1024"]["efficiency"],3),"traffic",d["roofline"]["traffic"],"frac",round(d["roofline"]["frac"],3))
PY
tail -2 gpurun_out/scale.err
