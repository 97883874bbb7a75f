#!/bin/bash
out=gpurun_out/r2w
mkdir -p $out
LIBS="libb200ba.so libb200ba_g.so" STEPS=9 bash tools/exp_variants.sh 2>&1 | tee $out/variants.txt
export B200BA_LIB=$PWD/jaxopt-3d-reconstruction_b200/libb200ba_g.so
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_resjac" -s 4 -c 4 -f -o $out/prof_resjac $CMD > $out/ncu.log 2>&1
echo "ncu exit $?"
