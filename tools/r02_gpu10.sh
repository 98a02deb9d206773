#!/bin/bash
# round-2 GPU call 10 (4 GPUs): multi-process tests after the dual-residual convergence loop learned to run one process per slab
mkdir -p gpurun_out
timeout 800 python -m pytest tests/test_multiprocess.py tests/test_gpu_driver.py tests/test_gpu_temporal.py -m gpu -q -rs --timeout 400 > gpurun_out/r02_pytest_mp.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_mp.log
tail -30 gpurun_out/r02_pytest_mp.log
W=/tmp/h3_conv; rm -rf $W; mkdir -p $W/output
D=/root/repo/cuda-aware-mpi-on-3d-heate-quation_b200/bin/heat3D
( cd $W; for r in 0 1 2 3; do HEAT3D_RANK=$r HEAT3D_WORLD=4 HEAT3D_LOCAL_RANK=$r HEAT3D_ID_FILE=$W/id HEAT3D_PERF=1 HEAT3D_OUT_FORMAT=raw HEAT3D_OUT_STRIDE=8 \
    timeout 300 $D 512 512 512 2000 1e-30 > out.$r 2> err.$r & done; wait; echo "== 4 processes, heat3D 512 512 512 2000 1e-30, residual examined every iteration"; cat out.0; grep perf err.0 ) > gpurun_out/r02_driver_4proc_every_iteration.txt 2>&1
cat gpurun_out/r02_driver_4proc_every_iteration.txt
