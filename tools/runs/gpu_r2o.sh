#!/bin/bash
# 1-GPU: L2 eviction-hint variant A/B, then the round's ncu evidence (launch list + --set full of the hot kernels) on the final build.
out=gpurun_out/r2o
mkdir -p $out
B="python bench.py --steps 18 --warmup 3 --no-pose --no-cpu-baseline --no-extras --no-e2e"
summ() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    k=d["extra"]["kernels"]; p=d["extra"]["parity"]
    print(sys.argv[1].split("/")[-1], "it/s %.1f"%d["value"], " ".join("%s=%.1f"%(n,k[n]["avg_us"]) for n in ("mv_pt","mv_cam","resjac","pt_blocks","pt_prepare","cam_schur","backsub","cost") if n in k), "dev %.2e"%(p.get("max_rel_loss_dev") or -1))
except Exception as e:
    print(sys.argv[1], "ERR", e, open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
timeout 200 $B > $out/base.json 2> $out/base.err; summ $out/base.json
B200BA_LIB=$PWD/variants/libb200ba_l2hint.so timeout 200 $B > $out/l2hint.json 2> $out/l2hint.err; summ $out/l2hint.json
timeout 200 $B > $out/base2.json 2> $out/base2.err; summ $out/base2.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose --no-extras --no-e2e"
timeout 300 $CMD > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 600 --csv --log-file $out/launches.csv $CMD > $out/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_ps|k_cs_ring" -s 40 -c 4 -f -o $out/prof_mv $CMD > $out/ncu_mv.log 2>&1
B200BA_LIB=$PWD/variants/libb200ba_l2hint.so timeout 600 ncu --set full --clock-control none -k regex:"k_ps|k_cs_ring" -s 40 -c 4 -f -o $out/prof_mv_l2hint $CMD > $out/ncu_mv_l2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_pose_lm" -s 3 -c 1 -f -o $out/prof_pose python tools/pose_probe.py > $out/ncu_pose.log 2>&1
cat > $out/dense_probe.py <<'PY'
import os, sys
sys.path.insert(0, '.')
os.environ["B200BA_DENSE"] = "1"; os.environ["B200BA_GRAPH"] = "0"
import b200ba
sc = b200ba.synth.make_ba_scene("ladybug49", seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
for _ in range(2): ba.optimize(x0, prob, None, k3, None)
PY
timeout 300 ncu --set full --clock-control none -k regex:"k_dn_factor|k_dn_assemble|k_resjac" -s 6 -c 4 -f -o $out/prof_dense python $out/dense_probe.py > $out/ncu_dense.log 2>&1
timeout 200 python tools/e2e_breakdown.py > $out/e2e_breakdown.txt 2>&1; cat $out/e2e_breakdown.txt | tail -4
ls -la $out | head -30
