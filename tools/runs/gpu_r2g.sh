#!/bin/bash
out=gpurun_out/r2g
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_dense.py tests/test_gpu_ba.py -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed" $out/pytest.log | tail -8
python - <<'PY'
import os, sys, json, time
sys.path.insert(0, '.')
import torch, b200ba, ctypes as C
for wl in ("ref15x400", "ladybug49", "trafalgar257"):
    for dense in ("1", "0"):
        os.environ["B200BA_DENSE"] = dense
        sc = b200ba.synth.make_ba_scene(wl, seed=0)
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        for _ in range(2): ba.optimize(x0, prob, None, k3, None)
        torch.cuda.synchronize(); t = time.perf_counter()
        n = 0
        for _ in range(5):
            _, st = ba.optimize(x0, prob, None, k3, None); n += st.iter_num
        torch.cuda.synchronize(); dt = time.perf_counter() - t
        prob.lib.b200ba_profile_enable(prob.plan, 1)
        ba.optimize(x0, prob, None, k3, None)
        ms = (C.c_double * 16)(); cnt = (C.c_int64 * 16)()
        prob.lib.b200ba_profile_get(prob.plan, ms, cnt); prob.lib.b200ba_profile_enable(prob.plan, 0)
        print(wl, "dense" if dense == "1" else "pcg", "us/LM-it %.0f" % (1e6 * dt / n), "asm %.1f chol %.1f solve %.1f" % tuple(1e3 * ms[i] / max(cnt[i], 1) for i in (13, 14, 15)), flush=True)
PY
