#!/bin/bash
# usage: gpurun --gpus 4 -- 'bash tools/scale_probe.sh "final13682:2 final13682:4 venice1778:2"'
# Short multi-rank bench runs (3 LM iterations each) with the per-kernel table; JSON lines land in gpurun_out/.
for spec in $1; do
  w=${spec%%:*}; n=${spec##*:}
  out=gpurun_out/scale_${w}_$n
  if [ "$n" = 1 ]; then
    timeout 280 python bench.py --steps 3 --warmup 3 --no-pose --no-cpu-baseline --workload $w 2> $out.err | grep "^{" > $out.json
  else
    timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $n --steps 3 --warmup 3 --no-pose --no-cpu-baseline --workload $w 2> $out.err | grep "^{" > $out.json
  fi
  python - $out.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1])); k = d["extra"]["kernels"]
print(sys.argv[1], "it/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1),
      {a: (b["launches"], round(b["avg_us"])) for a, b in k.items()})
PY
done
