#!/bin/bash
out=gpurun_out/r2v
mkdir -p $out
LIBS="libb200ba.so libb200ba_b.so libb200ba_c.so libb200ba_d.so libb200ba_e.so libb200ba_f.so" STEPS=9 bash tools/exp_variants.sh 2>&1 | tee $out/variants.txt
