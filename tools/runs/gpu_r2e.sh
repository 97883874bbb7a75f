#!/bin/bash
# Round-2 GPU call e: dense reduced-camera path (first run), full suite, bench with the small-scene extras.
out=gpurun_out/r2e
mkdir -p $out
timeout 300 compute-sanitizer --tool racecheck python -m pytest tests/test_gpu_dense.py -m gpu -q -x -k "tiny or small-False" > $out/racecheck.log 2>&1; tail -5 $out/racecheck.log
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest.log | tail -15
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -c 300 $out/bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2e/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"])
print(d["extra"]["parity"])
for k,v in d["extra"]["workloads"].items():
    if "dense" in v:
        for m in ("dense","pcg"):
            e=v[m]; print(k, m, {kk:vv for kk,vv in e.items() if kk in ("us_per_lm_iteration","launches_per_lm_iteration","pcg_iters_per_lm_iteration","kernels_us","error")})
    else: print(k, {kk:vv for kk,vv in v.items() if kk not in ("kernels","loss_trajectory")})
print({k:v for k,v in d["extra"]["kernels"].items()})
PY
