#!/bin/bash
out=gpurun_out/r2l
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_ba.py tests/test_gpu_reproj.py -m gpu -q > $out/pytest.log 2>&1; grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest.log | tail -5
timeout 300 python bench.py --steps 18 --warmup 3 --no-pose --no-cpu-baseline --no-extras > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2l/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], d["extra"]["parity"]["max_rel_loss_dev"])
print({k:(v["launches"], v["avg_us"]) for k,v in d["extra"]["kernels"].items()})
print(d["extra"]["solves"])
PY
timeout 200 python tools/plan_timing.py > $out/plan_timing.txt 2>&1; grep -E "rep " $out/plan_timing.txt | tail -3
