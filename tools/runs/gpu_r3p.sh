#!/bin/bash
mkdir -p gpurun_out/r3p
timeout 300 python -m pytest tests/test_gpu_ba.py -m gpu -q -x -k "without_observations" 2>&1 | tail -15 | tee gpurun_out/r3p/heavy_point.log
