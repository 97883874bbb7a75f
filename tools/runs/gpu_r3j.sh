#!/bin/bash
mkdir -p gpurun_out/r3j
timeout 900 python tools/forcing_sweep.py "0:1e-6" "0.25:1e-6" "0.5:1e-6" "0.5:1e-7" "0.75:1e-6" "1:1e-6" "1:1e-5" 2>&1 | tee gpurun_out/r3j/forcing.jsonl
