#!/bin/bash
# round-2 GPU call 13 (1 GPU): neighbour-to-neighbour barrier variant "_Q" - parity first (short timeouts: a wrong handshake hangs), then P vs Q
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_temporal.py -m gpu -q -x --timeout 60 -k "every_depth2_variant" > gpurun_out/r02_pytest_q.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_q.log; tail -5 gpurun_out/r02_pytest_q.log
grep -q "pytest rc=0" gpurun_out/r02_pytest_q.log || exit 1
SWEEP_STRICT_ONLY=1 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2Q" timeout 200 python tools/sweep_t2.py 512 400 > gpurun_out/r02_sweep_q_burst.txt 2>&1
SWEEP_STRICT_ONLY=1 SWEEP_SUSTAINED=1 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2Q;64,32,4,4,2P;64,32,4,4,2Q" SWEEP_CHUNKS=0 timeout 300 python tools/sweep_t2.py 512 1500 > gpurun_out/r02_sweep_q_sust.txt 2>&1
SWEEP_SHAPE=1024,1024,512 SWEEP_STRICT_ONLY=1 SWEEP_CHUNKS=0 SWEEP_TUNES="64,32,4,4,2P;64,32,4,4,2Q" timeout 300 python tools/sweep_t2.py 512 200 > gpurun_out/r02_sweep_q_1024.txt 2>&1
cat gpurun_out/r02_sweep_q_burst.txt gpurun_out/r02_sweep_q_sust.txt gpurun_out/r02_sweep_q_1024.txt
