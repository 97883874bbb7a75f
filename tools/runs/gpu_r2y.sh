#!/bin/bash
mkdir -p gpurun_out/r2y
timeout 300 python tools/e2e_breakdown2.py venice1778 2>&1 | tee gpurun_out/r2y/e2e2.txt
