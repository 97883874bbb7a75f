#!/bin/bash
# Round-2 third GPU call: full GPU test suite under both gradient modes, tile-kernel v2 timing, ncu --set full of the hot kernels.
out=gpurun_out/r2c
mkdir -p $out
B="python bench.py --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-extras --no-e2e"
summ() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    k=d["extra"]["kernels"]; p=d["extra"]["parity"]
    print(sys.argv[1].split("/")[-1], "it/s %.1f"%d["value"], " ".join("%s=%.1f"%(n,k[n]["avg_us"]) for n in ("mv_pt","mv_cam","resjac","pt_blocks","pt_prepare","cam_schur","backsub","cost") if n in k),
          "pcg", [s["pcg_iters"] for s in d["extra"]["solves"]], "dev %.2e"%(p.get("max_rel_loss_dev") or -1), d["extra"].get("plan"))
except Exception as e:
    print(sys.argv[1], "ERR", e, open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
for eg in 0 1; do
  B200BA_EXACT_GRAD=$eg timeout 900 python -m pytest tests -m gpu -q -k "not 4096x512" > $out/pytest_eg$eg.log 2>&1; echo "pytest exact_grad=$eg exit $?" >> $out/pytest_eg$eg.log
  grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest_eg$eg.log | tail -15
done
timeout 300 python -m pytest tests -m gpu -q -k "4096x512" > $out/pytest_pose_full.log 2>&1; tail -3 $out/pytest_pose_full.log
for tr in 8192 5632 4096 11264; do
  B200BA_TILE_ROWS=$tr timeout 200 $B > $out/tile$tr.json 2> $out/tile$tr.err; summ $out/tile$tr.json
done
B200BA_EXACT_GRAD=1 timeout 200 $B > $out/tile8192_eg1.json 2> $out/tile8192_eg1.err; summ $out/tile8192_eg1.json
B200BA_PS_TILE=0 timeout 200 $B > $out/chunk.json 2> $out/chunk.err; summ $out/chunk.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_ps_tile|k_cs_ring" -s 40 -c 4 -f -o $out/prof_mv $CMD > $out/ncu_mv.log 2>&1
B200BA_PS_TILE=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_ps" -s 40 -c 1 -f -o $out/prof_chunk $CMD > $out/ncu_chunk.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_resjac|k_point_blocks|k_point_prepare|k_cam_blocks" -s 4 -c 6 -f -o $out/prof_lin $CMD > $out/ncu_lin.log 2>&1
ls -la $out | head -40
