#!/bin/bash
out=gpurun_out/r2k
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest.log | tail -12
timeout 200 python tools/pose_probe.py > $out/pose.json 2> $out/pose.err; echo "pose probe exit $?"; cat $out/pose.json | cut -c1-700

timeout 200 python tools/reproj_probe.py > $out/reproj.json 2> $out/reproj.err; cat $out/reproj.json
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -c 300 $out/bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2k/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"])
print(d["extra"]["parity"])
for k,v in d["extra"]["workloads"].items():
    if "dense" in v:
        for m in ("dense","pcg"):
            e=v[m]; print(k, m, {kk:vv for kk,vv in e.items() if kk in ("us_per_lm_iteration","launches_per_lm_iteration","pcg_iters_per_lm_iteration","kernels_us","error")})
    else: print(k, {kk:vv for kk,vv in v.items() if kk not in ("kernels","loss_trajectory")})
print({k:(v["avg_us"],v.get("frac")) for k,v in d["extra"]["kernels"].items()})
print({k:v for k,v in d["extra"]["pose"].items() if "hist" not in k})
PY
