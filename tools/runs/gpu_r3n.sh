#!/bin/bash
mkdir -p gpurun_out/r3n
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -k "point_sharded or warp_per_camera or three_launch or without_rows" > gpurun_out/r3n/multi.log 2>&1; echo "multi exit $?" >> gpurun_out/r3n/multi.log; tail -4 gpurun_out/r3n/multi.log
