#!/bin/bash
# round-2 GPU call 11 (1 GPU): smoke(), the heat3D executable on BASELINE configs 1 and 2, DRAM traffic of the depth-2 kernel on a 1024x1024x512 slab
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
D=/root/repo/cuda-aware-mpi-on-3d-heate-quation_b200/bin/heat3D
W=/tmp/h3_cfg; rm -rf $W; mkdir -p $W/output; cd $W
( echo "== config 1: heat3D 64 64 64 1000 1e-30 (reference semantics: residual examined every iteration)"; HEAT3D_PERF=1 timeout 100 $D 64 64 64 1000 1e-30 2>&1 | grep -v "^executable"; md5sum output/out.dat
  echo "== config 2: heat3D 512 512 512 10000 1e-30 (same semantics; raw dump, stride 8 - the text dump would be 8.2 GB)"; HEAT3D_PERF=1 HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=8 timeout 200 $D 512 512 512 10000 1e-30 2>&1 | grep -v "^executable"
  echo "== config 2 with HEAT3D_CHECK_EVERY=1000"; HEAT3D_PERF=1 HEAT3D_CHECK_EVERY=1000 HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=8 timeout 200 $D 512 512 512 10000 1e-30 2>&1 | grep -v "^executable" ) > /root/repo/gpurun_out/r02_driver_runs_configs_1_2.txt 2>&1
cd /root/repo; cat gpurun_out/r02_driver_runs_configs_1_2.txt
PROF_SHAPE=1024,1024,512 timeout 200 python tools/prof_t2.py 512 20 > gpurun_out/r02_plain11.log 2>&1 &&
PROF_SHAPE=1024,1024,512 timeout 600 ncu --set full --clock-control none -k regex:stencil_tma_t2 -s 6 -c 1 -o gpurun_out/r02_prof_t2_1024 -f python tools/prof_t2.py 512 20 > gpurun_out/r02_ncu_f11.log 2>&1
tail -2 gpurun_out/r02_ncu_f11.log
