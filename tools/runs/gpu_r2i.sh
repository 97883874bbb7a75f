#!/bin/bash
# 2-GPU call: fused multi-rank ring kernel (ba_cs_ringx.cuh) — single-rank pipeline tests, multi-GPU parity tests, scaling probe.
out=gpurun_out/r2i
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/smi.txt
timeout 300 python -m pytest tests/test_gpu_noise.py -m gpu -q > $out/pytest_noise.log 2>&1; tail -3 $out/pytest_noise.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $out/pytest_multi.log 2>&1; grep -E "^(FAILED|ERROR)|passed|failed|MULTI" $out/pytest_multi.log | tail -12
for rx in 1 0; do
  B200BA_RINGX=$rx timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-e2e > $out/bench2_rx$rx.json 2> $out/bench2_rx$rx.err
  echo "bench2 ringx=$rx exit $?"; tail -c 400 $out/bench2_rx$rx.err
  python - $out/bench2_rx$rx.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    k=d["extra"]["kernels"]
    print("N=2 value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "parity", d["extra"]["parity"].get("max_rel_loss_dev"), {n:k[n]["avg_us"] for n in ("mv_pt","mv_cam","pcg_vec","resjac") if n in k})
    f=d["extra"]["workloads"]["final13682"]; print("final13682", f.get("lm_iterations_per_s"), {n:v["avg_us"] for n,v in f.get("kernels",{}).items() if n in ("mv_pt","mv_cam","pcg_vec")}, f.get("error"))
except Exception as e: print("ERR", e)
PY
done
