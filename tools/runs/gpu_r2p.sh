#!/bin/bash
# 1-GPU stand-in for ONE rank of Final-13682 on 8 ranks (shape final13682_r8, multi-rank kernel sequence forced): do L2
# eviction-priority hints keep the rank's 116 MB of Jacobian rows in L2 across the two camera passes?
out=gpurun_out/r2p
mkdir -p $out
B="python bench.py --workload final13682_r8 --steps 8 --warmup 3 --no-pose --no-cpu-baseline --no-extras --no-e2e"
summ() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    k=d["extra"]["kernels"]
    print(sys.argv[1].split("/")[-1], "it/s %.1f"%d["value"], " ".join("%s=%.1f"%(n,k[n]["avg_us"]) for n in ("mv_pt","mv_cam","pcg_vec","resjac","pt_blocks","backsub","cost") if n in k))
except Exception as e:
    print(sys.argv[1], "ERR", e, open(sys.argv[1].replace(".json",".err")).read()[-600:])
PY
}
export B200BA_FORCE_SPLIT=1 B200BA_RINGX=0
timeout 200 $B > $out/base.json 2> $out/base.err; summ $out/base.json
for fr in 0.5 0.75 1.0; do
  B200BA_LIB=$PWD/variants/libb200ba_l2hint_$fr.so timeout 200 $B > $out/l2_$fr.json 2> $out/l2_$fr.err; summ $out/l2_$fr.json
done
timeout 200 $B > $out/base2.json 2> $out/base2.err; summ $out/base2.json
CMD="$B --steps 2"
B200BA_LIB=$PWD/variants/libb200ba_l2hint_0.75.so timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_cs_warp|k_ps" -s 60 -c 6 --csv --log-file $out/ncu_l2.csv $CMD > /dev/null 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_cs_warp|k_ps" -s 60 -c 6 --csv --log-file $out/ncu_base.csv $CMD > /dev/null 2>&1
python - <<'PY'
import csv
for f in ("gpurun_out/r2p/ncu_base.csv","gpurun_out/r2p/ncu_l2.csv"):
    try:
        rows=[r for r in csv.DictReader(l for l in open(f) if not l.startswith("=="))]
        agg={}
        for r in rows: agg.setdefault((r["ID"], r["Kernel Name"][:30]),{})[r["Metric Name"]]=r["Metric Value"]+" "+r["Metric Unit"]
        print(f); [print("  ",k,v) for k,v in agg.items()]
    except Exception as e: print(f, "ERR", e)
PY
