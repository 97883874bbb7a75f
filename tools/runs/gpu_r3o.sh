#!/bin/bash
mkdir -p gpurun_out/r3o
timeout 300 python -m pytest tests/test_gpu_vector2.py tests/test_gpu_dense.py -m gpu -q -x 2>&1 | tail -2
timeout 400 python tools/dense_grid_probe.py 2>&1 | tee gpurun_out/r3o/dense_grid.jsonl
