#!/bin/bash
# round-2 GPU call 15 (1 GPU): the whole suite after the residual variant's uniform path, then heat3D on config 2 in reference semantics
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 > gpurun_out/r02_pytest15.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest15.log; tail -3 gpurun_out/r02_pytest15.log
D=/root/repo/cuda-aware-mpi-on-3d-heate-quation_b200/bin/heat3D
W=/tmp/h3_cfg; rm -rf $W; mkdir -p $W/output; cd $W
HEAT3D_PERF=1 HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=8 timeout 200 $D 512 512 512 10000 1e-30 2>&1 | grep -E "Computational|perf"
