#!/bin/bash
# 8-GPU call: scaling of the point-sharded solve (Venice-1778 headline + Final-13682 in extra.workloads) with the cost-model
# choice of the camera-side scheme, and Venice with the three-launch fallback forced.
out=gpurun_out/r2n
mkdir -p $out
run() { # name, env..., extra args
  name=$1; shift
  env "$@" timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 18 --warmup 3 --no-cpu-baseline $EXTRA > $out/$name.json 2> $out/$name.err
  echo "$name exit $?"
  python - $out/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    k=d["extra"]["kernels"]
    print("N=8 value", round(d["value"],1), "e2e", d["e2e"] and round(d["e2e"]["value"],1), "parity", d["extra"]["parity"].get("max_rel_loss_dev"), d["extra"]["parity"].get("rms_rel_dev"), {n:k[n]["avg_us"] for n in ("mv_pt","mv_cam","pcg_vec","resjac","other") if n in k}, "split/ringx", d["extra"]["plan"])
    w=d["extra"].get("workloads",{})
    if "final13682" in w:
        f=w["final13682"]; print("final13682", f.get("lm_iterations_per_s"), {n:v["avg_us"] for n,v in f.get("kernels",{}).items() if n in ("mv_pt","mv_cam","pcg_vec")}, f.get("error"))
    if "pose" in d["extra"]: print("pose", {a:b for a,b in d["extra"]["pose"].items() if "hist" not in a and "note" not in a})
except Exception as e: print("ERR", e)
PY
}
EXTRA="" run auto A=1
EXTRA="--no-extras --no-pose --no-e2e" run split B200BA_RINGX=0
EXTRA="--no-extras --no-pose --no-e2e" run fused B200BA_RINGX=1
