#!/bin/bash
# the driver's N=8 line with the default environment
mkdir -p gpurun_out/r3l
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 18 --warmup 3 > gpurun_out/r3l/bench8.json 2> gpurun_out/r3l/bench8.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r3l/bench8.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"], d["extra"]["parity"]["max_rel_loss_dev"], d["extra"]["plan"])
print({k:(v["launches"], v["avg_us"]) for k,v in d["extra"]["kernels"].items()})
for k,v in d["extra"]["workloads"].items():
    if "dense" not in v: print(k, {kk:vv for kk,vv in v.items() if kk not in ("kernels","loss_trajectory","plan")})
print({k:v for k,v in d["extra"].get("pose",{}).items() if k in ("n_gpus","solves_per_s","e2e_solves_per_s")})
PY
