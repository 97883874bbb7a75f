#!/bin/bash
out=gpurun_out/r2z
mkdir -p $out
LIBS="libb200ba.so libb200ba_r1.so libb200ba_r2.so libb200ba_r3.so libb200ba_r4.so" STEPS=9 bash tools/exp_variants.sh 2>&1 | tee $out/variants.txt
