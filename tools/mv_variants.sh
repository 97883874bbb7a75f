#!/bin/bash
# matvec variant experiment: w placement (gather in pass 2 | scatter in pass 1), optional extra libs
for lib in ${LIBS:-libb200ba.so}; do for ws in 0 1; do
  echo "== $lib WSCATTER=$ws"
  B200BA_LIB=$PWD/jaxopt-3d-reconstruction_b200/$lib B200BA_WSCATTER=$ws python bench.py --steps 7 --warmup 3 --no-cpu-baseline --no-pose 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],1), 'mv_pt',k['mv_pt']['avg_us'],'mv_cam',k['mv_cam']['avg_us'],'vec',k['pcg_vec']['avg_us'],'backsub',k['backsub']['avg_us'],'resjac',k['resjac']['avg_us'],'cam_blocks',k['cam_blocks']['avg_us'],'cam_schur',k['cam_schur']['avg_us'])"
done; done
