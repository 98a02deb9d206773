#!/bin/bash
# round-2 GPU call 8 (8 GPUs): default bench line as the driver launches it; BASELINE config 5 (2048^3) plain, ghost depth 4,
# with full and with decimated periodic snapshots; the heat3D executable as 8 processes and as one process driving 8 GPUs
mkdir -p gpurun_out
rm -f gpurun_out/scale_r02_n8.jsonl gpurun_out/scale_r02_c5.jsonl gpurun_out/scale.err
timeout 500 bash tools/scale_run.sh 8 "--steps 5 --warmup 3" r02_n8 > /dev/null
C5="--grid 2048,2048,2048 --steps 6 --warmup 2 --no-e2e --no-base --no-preflight"
timeout 240 bash tools/scale_run.sh 8 "$C5" r02_c5 > /dev/null
timeout 240 bash tools/scale_run.sh 8 "$C5 --temporal-k 4" r02_c5 > /dev/null
timeout 300 bash tools/scale_run.sh 8 "$C5 --snapshot-every 3" r02_c5 > /dev/null
timeout 300 bash tools/scale_run.sh 8 "$C5 --snapshot-every 3 --snapshot-stride 2" r02_c5 > /dev/null
python - <<'PY'
import json
for f in ("gpurun_out/scale_r02_n8.jsonl","gpurun_out/scale_r02_c5.jsonl"):
    for l in open(f):
        if not l.startswith("{"): continue
        d=json.loads(l)
        print(d["n_gpus"],d["config"]["workload"][:24],"k",d["config"]["temporal_k"],"value",round(d["value"],1),"wall",round(d["wall_glups_incl_snapshots"],1),
              "e2e",d["e2e"] and (round(d["e2e"]["value"],1),round(d["e2e"]["blocking"]["value"],1)),"pre",d.get("parity_preflight") and d["parity_preflight"]["bitwise"],
              "base",d.get("scaling_base_glups"),"eff",d.get("parallel_efficiency"),"strong",d.get("strong_1024") and (round(d["strong_1024"]["value"],1),round(d["strong_1024"]["efficiency"],3)),
              "halo",d.get("exposed_halo_ms_per_update"),d.get("exposed_halo_ms_per_update_min_rank"),"snap",d.get("snapshots"))
PY
tail -3 gpurun_out/scale.err
# the executable: 8 processes, one per slab (the reference's mpirun -n 8), decimated raw dump
D=cuda-aware-mpi-on-3d-heate-quation_b200/bin/heat3D
W=/tmp/h3_mp; rm -rf $W; mkdir -p $W/output
( cd $W; for r in 0 1 2 3 4 5 6 7; do HEAT3D_RANK=$r HEAT3D_WORLD=8 HEAT3D_LOCAL_RANK=$r HEAT3D_ID_FILE=$W/id HEAT3D_PERF=1 HEAT3D_CHECK_EVERY=100 \
    HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=4 timeout 200 /root/repo/$D 1024 1024 1024 400 1e-30 > out.$r 2> err.$r & done; wait )
( echo "== 8 processes: heat3D 1024 1024 1024 400 1e-30 (HEAT3D_WORLD=8, CHECK_EVERY=100, raw dump, stride 4)"; cat $W/out.0; cat $W/err.0; ls -la $W/output ) > gpurun_out/r02_driver_8gpu.txt 2>&1
W=/tmp/h3_sp; rm -rf $W; mkdir -p $W/output
( cd $W; HEAT3D_NDEV=8 HEAT3D_PERF=1 HEAT3D_CHECK_EVERY=100 HEAT3D_OUT_EVERY=100 HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=4 timeout 200 /root/repo/$D 1024 1024 1024 400 1e-30 > out 2> err )
( echo "== 1 process, 8 GPUs: same run with HEAT3D_OUT_EVERY=100 (decimated snapshots in flight while stepping)"; cat $W/out; cat $W/err; ls -la $W/output ) >> gpurun_out/r02_driver_8gpu.txt 2>&1
cat gpurun_out/r02_driver_8gpu.txt
