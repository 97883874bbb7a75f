#!/bin/bash
out=gpurun_out/r3b
mkdir -p $out
run() {
  timeout 300 python bench.py --steps 18 --warmup 3 --no-cpu-baseline --no-pose --no-extras 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],2), 'e2e', round(d['e2e']['value'],2), ' '.join(f'{n}={k[n][\"avg_us\"]}' for n in ('mv_pt','mv_cam') if n in k), d['extra']['parity']['max_rel_loss_dev'])"
}
for lib in libb200ba.so libb200ba_o.so; do
  for pdl in 0 1 2 3; do
    echo "== $lib PDL=$pdl"
    B200BA_LIB=$PWD/jaxopt-3d-reconstruction_b200/$lib B200BA_PDL=$pdl run
  done
done 2>&1 | tee $out/pdl.txt
