#!/bin/bash
# Single-GPU profile of the multi-rank kernel sequence (B200BA_FORCE_SPLIT=1) on a one-rank share of Final-13682.
# usage: gpurun --timeout 900 -- 'bash tools/ncu_split.sh final13682_r4 <tag>'
w=${1:-final13682_r4}; tag=${2:-split}
out=gpurun_out/$tag; mkdir -p $out
export B200BA_FORCE_SPLIT=1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose --workload $w"
timeout 300 $CMD > $out/bench.json 2> $out/bench.err || exit 1
python - $out/bench.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1])); k = d["extra"]["kernels"]
print("it/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), {a: (b["launches"], round(b["avg_us"])) for a, b in k.items()})
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_cs_warp|k_cs_vec|k_ps" -s 60 -c 8 -f -o $out/prof $CMD > $out/ncu.log 2>&1
echo "ncu exit $?"
