#!/bin/bash
# round-2 GPU call 6 (4 GPUs): the whole GPU suite with nothing skipped for lack of devices, default bench lines at N=2,4
# (parity pre-flight, like-for-like scaling base, strong 1024^3 block), ghost depth 4 vs 2 at N=4
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_mg_gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q -rs --timeout 300 > gpurun_out/r02_pytest_4gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_4gpu.log
tail -12 gpurun_out/r02_pytest_4gpu.log
rm -f gpurun_out/scale_r02_default.jsonl gpurun_out/scale_r02_k4.jsonl gpurun_out/scale.err
for N in 2 4; do
  timeout 600 bash tools/scale_run.sh $N "--steps 5 --warmup 3" r02_default > /dev/null
done
cut -c1-400 gpurun_out/scale_r02_default.jsonl
for K in 2 4; do
  timeout 300 bash tools/scale_run.sh 4 "--steps 4 --warmup 2 --workload strong --temporal-k $K --no-e2e --no-base --no-preflight" r02_k4 > /dev/null
  timeout 300 bash tools/scale_run.sh 4 "--steps 4 --warmup 2 --workload weak --temporal-k $K --no-e2e --no-base --no-preflight" r02_k4 > /dev/null
done
python - <<'PY'
import json
for f in ("gpurun_out/scale_r02_default.jsonl","gpurun_out/scale_r02_k4.jsonl"):
    for l in open(f):
        try: d=json.loads(l)
        except Exception: print("bad line",l[:200]); continue
        print(d["n_gpus"],d["config"]["workload"][:20],"k",d["config"]["temporal_k"],"value",round(d["value"],1),"e2e",d["e2e"] and round(d["e2e"]["value"],1),
              "pre",d.get("parity_preflight") and d["parity_preflight"]["bitwise"],"base",d.get("scaling_base_glups"),"eff",d.get("parallel_efficiency"),
              "strong",d.get("strong_1024") and (round(d["strong_1024"]["value"],1),round(d["strong_1024"]["efficiency"],3)),
              "halo",d.get("exposed_halo_ms_per_update"),d.get("exposed_halo_ms_per_update_min_rank"))
PY
tail -5 gpurun_out/scale.err
