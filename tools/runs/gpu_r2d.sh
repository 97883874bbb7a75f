#!/bin/bash
# Round-2 GPU call d: full GPU suite on the settled defaults (chunk point kernel, stored-row gradient, DMMA camera blocks in the
# fused residual/Jacobian kernel), rtol sweep against the goldens, short bench.
out=gpurun_out/r2d
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
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest.log | tail -15
timeout 200 $B > $out/base.json 2> $out/base.err; summ $out/base.json
timeout 600 python tools/rtol_sweep.py > $out/rtol_sweep.jsonl 2> $out/rtol_sweep.err; cat $out/rtol_sweep.jsonl
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_resjac" -s 4 -c 3 -f -o $out/prof_resjac $CMD > $out/ncu_resjac.log 2>&1
ls -la $out | head -40
