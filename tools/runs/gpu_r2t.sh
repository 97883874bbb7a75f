#!/bin/bash
out=gpurun_out/r2t
mkdir -p $out
timeout 300 python tools/e2e_breakdown.py venice1778 > $out/e2e_venice.txt 2>&1; cat $out/e2e_venice.txt | tail -5
timeout 300 python tools/plan_timing.py > $out/plan_timing.txt 2>&1; tail -30 $out/plan_timing.txt
