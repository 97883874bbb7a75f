#!/bin/bash
# the allocation-free entry point alone in a fresh process, then the whole GPU suite (it is collected last), then the bench
out=gpurun_out/r3c
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/noalloc_alone.log 2>&1; echo "alone exit $?" >> $out/noalloc_alone.log
tail -5 $out/noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" $out/pytest.log | tail -12
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -c 300 $out/bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r3c/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"])
print(d["extra"]["parity"])
for k,v in d["extra"]["workloads"].items():
    if "dense" in v:
        for m in ("dense","pcg"):
            e=v[m]; print(k, m, {kk:vv for kk,vv in e.items() if kk in ("us_per_lm_iteration","launches_per_lm_iteration","pcg_iters_per_lm_iteration","error")})
    else: print(k, {kk:vv for kk,vv in v.items() if kk not in ("kernels","loss_trajectory")})
print({k:(v["avg_us"],v.get("frac")) for k,v in d["extra"]["kernels"].items()})
print({k:v for k,v in d["extra"]["pose"].items() if "hist" not in k})
PY
timeout 300 python tools/plan_timing.py > gpurun_out/r3c/plan_timing.txt 2>&1; tail -9 gpurun_out/r3c/plan_timing.txt
