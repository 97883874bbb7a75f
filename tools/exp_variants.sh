#!/bin/bash
# timing experiments: run the bench with variant libraries (results are NOT correct solves; kernel times only)
for lib in ${LIBS:-libb200ba.so}; do
  echo "== $lib"
  B200BA_LIB=$PWD/jaxopt-3d-reconstruction_b200/$lib timeout 300 python bench.py --steps ${STEPS:-3} --warmup 3 --no-cpu-baseline --no-pose 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],1), ' '.join(f'{n}={k[n][\"avg_us\"]}' for n in ('mv_pt','mv_cam','backsub','resjac','cost','cam_blocks','cam_schur','pt_prepare','pt_blocks','cam_finish') if n in k), [s['pcg_iters'] for s in d['extra']['solves']])"
done
