#!/bin/bash
# Round-2 second GPU call: correctness of the new kernels (tiled point side, fused residual/Jacobian/camera blocks, 32 B point
# records), then A/B timing of the point-side variants and an rtol sweep with parity against the Venice golden.
out=gpurun_out/r2b
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
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
tail -12 $out/pytest.log
./tools/dmma_probe > $out/dmma.json 2>&1; cat $out/dmma.json
timeout 200 $B > $out/base_tile8192.json 2> $out/base_tile8192.err; summ $out/base_tile8192.json
B200BA_PS_TILE=0 timeout 200 $B > $out/chunk.json 2> $out/chunk.err; summ $out/chunk.json
for tr in 4096 5632 11264 16384; do
  B200BA_TILE_ROWS=$tr timeout 200 $B > $out/tile$tr.json 2> $out/tile$tr.err; summ $out/tile$tr.json
done
for v in t384m2 t128m6 t256m2; do
  if [ -f variants/libb200ba_$v.so ]; then
    for tr in 8192 16384; do
      B200BA_LIB=$PWD/variants/libb200ba_$v.so B200BA_TILE_ROWS=$tr timeout 200 $B > $out/${v}_$tr.json 2> $out/${v}_$tr.err; summ $out/${v}_$tr.json
    done
  fi
done
for rt in 1e-9 1e-8 1e-7; do
  timeout 200 $B --pcg-rtol $rt > $out/rtol$rt.json 2> $out/rtol$rt.err; summ $out/rtol$rt.json
done
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -c 1500 $out/bench.json
ls $out | head -50
