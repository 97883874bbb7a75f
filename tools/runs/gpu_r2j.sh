#!/bin/bash
out=gpurun_out/r2m
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -q -k "fused or default or four or sharded" > $out/pytest_multi.log 2>&1; grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest_multi.log | tail -5
B200BA_RINGX=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-e2e > $out/bench2_rx1.json 2> $out/bench2_rx1.err
echo "bench2 exit $?"
python - $out/bench2_rx1.json <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
k=d["extra"]["kernels"]
print("N=2 value", round(d["value"],1), "parity", d["extra"]["parity"].get("max_rel_loss_dev"), {n:k[n]["avg_us"] for n in ("mv_pt","mv_cam","pcg_vec","resjac") if n in k})
f=d["extra"]["workloads"]["final13682"]; print("final13682", f.get("lm_iterations_per_s"), {n:v["avg_us"] for n,v in f.get("kernels",{}).items() if n in ("mv_pt","mv_cam","pcg_vec")}, f.get("error"))
PY
