#!/bin/bash
out=gpurun_out/r3s
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" $out/pytest.log | tail -8
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -c 300 $out/bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r3s/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"]["frac"], d["extra"]["parity"]["max_rel_loss_dev"])
for k,v in d["extra"]["workloads"].items():
    if "dense" in v:
        for m in ("dense","pcg"):
            e=v[m]; print(k, m, {kk:vv for kk,vv in e.items() if kk in ("us_per_lm_iteration","launches_per_lm_iteration")})
PY
