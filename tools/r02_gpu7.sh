#!/bin/bash
# round-2 GPU call 7 (4 GPUs): the whole GPU suite, nothing skipped for lack of devices
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_mg_gpus.txt
timeout 1200 python -m pytest tests -m gpu -q -rs --timeout 300 > gpurun_out/r02_pytest_4gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest_4gpu.log
tail -40 gpurun_out/r02_pytest_4gpu.log
