#!/bin/bash
mkdir -p gpurun_out/r3r
for pdl in 2 3 2 3; do
  echo "== PDL=$pdl (ring triggers its dependents when a CTA starts its last camera)"
  B200BA_PDL=$pdl timeout 200 python bench.py --steps 18 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],2), d['extra']['parity']['max_rel_loss_dev'])"
done 2>&1 | tee gpurun_out/r3r/pdl_late.txt
