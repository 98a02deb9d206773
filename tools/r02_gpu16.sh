#!/bin/bash
# round-2 GPU call 16 (1 GPU): sensitivity of the guided chunk plan to its two model parameters (sustained, 512^3 and 1024x1024x512)
mkdir -p gpurun_out; : > gpurun_out/r02_sweep_plan_params.txt
for OVH in 3 6 10; do for ML in 160 260; do
  echo "# HEAT3D_T2_OVH=$OVH HEAT3D_T2_MAXLEN=$ML" >> gpurun_out/r02_sweep_plan_params.txt
  HEAT3D_T2_OVH=$OVH HEAT3D_T2_MAXLEN=$ML SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 SWEEP_TUNES="64,32,4,4,2P" SWEEP_CHUNKS=0 timeout 100 python tools/sweep_t2.py 512 1200 >> gpurun_out/r02_sweep_plan_params.txt 2>&1
  HEAT3D_T2_OVH=$OVH HEAT3D_T2_MAXLEN=$ML SWEEP_SHAPE=1024,1024,128 SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 SWEEP_TUNES="64,32,4,4,2P" SWEEP_CHUNKS=0 timeout 100 python tools/sweep_t2.py 512 1200 >> gpurun_out/r02_sweep_plan_params.txt 2>&1
done; done
cut -c1-110 gpurun_out/r02_sweep_plan_params.txt
