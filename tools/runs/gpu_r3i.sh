#!/bin/bash
# 2 GPUs, default environment: the whole multi-rank test file, then the bench line (Venice) as the driver runs it
out=gpurun_out/r3i
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $out/multi.log 2>&1; echo "multi exit $?" >> $out/multi.log; tail -5 $out/multi.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 18 --warmup 3 --no-extras > $out/bench2.json 2> $out/bench2.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r3i/bench2.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"], d["extra"]["plan"], d["extra"]["parity"]["max_rel_loss_dev"])
print({k:(v["launches"], v["avg_us"]) for k,v in d["extra"]["kernels"].items()})
print({k:v for k,v in d["extra"].get("pose",{}).items() if k in ("n_gpus","solves_per_s","e2e_solves_per_s")})
PY
