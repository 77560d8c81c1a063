#!/bin/bash
# first GPU trip: tests + quick timing sweep
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q -s 2>&1 | tail -40 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
python tools/sweep.py > gpurun_out/sweep.log 2>&1
cat gpurun_out/sweep.log
