#!/bin/bash
# Final single-GPU run of round 2: tests, bench line, ncu launch list + full capture of the PCG kernels, Final-13682 window,
# and the fused multi-rank kernel profiled on one rank.
out=gpurun_out/r2final
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log; tail -3 $out/pytest.log
timeout 400 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_ref.json 2> $out/bench_ref.err; echo "ref exit $?"; tail -c 400 $out/bench_ref.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose --no-extras"
timeout 300 $CMD > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file $out/launches.csv $CMD > $out/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_cs_ring|k_ps" -s 30 -c 4 -f -o $out/prof_mv $CMD > $out/ncu_full_mv.log 2>&1
echo "ncu mv exit $?"
timeout 300 python bench.py --steps 9 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e --workload final13682 > $out/final13682_1gpu_9.json 2> $out/final.err; echo "final exit $?"
B200BA_FORCE_SPLIT=1 B200BA_WARPX=1 timeout 300 $CMD > $out/plain_warpx.json 2>&1 &&
B200BA_FORCE_SPLIT=1 B200BA_WARPX=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_cs_warpx" -s 10 -c 2 -f -o $out/prof_warpx $CMD > $out/ncu_full_warpx.log 2>&1
echo "ncu warpx exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2final/bench.json') if l.startswith('{')][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["roofline"], "launches", d.get("gpu_launches"))
print(d["extra"]["parity"])
for k,v in d["extra"]["workloads"].items():
    if "dense" in v:
        for m in ("dense","pcg"):
            e=v[m]; print(k, m, {kk:vv for kk,vv in e.items() if kk in ("us_per_lm_iteration","launches_per_lm_iteration")})
    else: print(k, {kk:vv for kk,vv in v.items() if kk not in ("kernels","loss_trajectory","plan")})
print({k:(v["avg_us"],v.get("frac")) for k,v in d["extra"]["kernels"].items()})
f=json.loads([l for l in open('gpurun_out/r2final/final13682_1gpu_9.json') if l.startswith('{')][-1]); print("final13682 1 GPU, 9 its:", f["value"])
PY
ls -la $out
