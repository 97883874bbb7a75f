#!/bin/bash
out=gpurun_out/r3a
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -q -x > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" $out/pytest.log | tail -12
for pdl in 0 1 0 1; do
  echo "== PDL=$pdl"
  B200BA_PDL=$pdl timeout 300 python bench.py --steps 18 --warmup 3 --no-cpu-baseline --no-pose --no-extras 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],2), 'e2e', round(d['e2e']['value'],2), ' '.join(f'{n}={k[n][\"avg_us\"]}' for n in ('mv_pt','mv_cam') if n in k), d['extra']['parity']['max_rel_loss_dev'], [s['pcg_iters'] for s in d['extra']['solves']])"
done 2>&1 | tee $out/pdl.txt
