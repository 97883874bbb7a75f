#!/bin/bash
out=gpurun_out/r3d
mkdir -p $out
LIBS="libb200ba.so libb200ba_p1.so libb200ba_p2.so libb200ba_p3.so libb200ba_p4.so libb200ba_p5.so" STEPS=9 bash tools/exp_variants.sh 2>&1 | tee $out/variants.txt
