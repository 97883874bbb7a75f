#!/bin/bash
# One gpurun call: GPU parity tests, bench, launch list and one full ncu capture of the hot kernels.
# usage (from the repo root):  gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
if [ -z "$SKIP_TESTS" ]; then
  timeout 400 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
  tail -3 $out/pytest.log
fi
timeout 300 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -c 600 $out/bench.json
B200BA_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_vector2.py -m gpu -q > $out/pytest_vector2.log 2>&1; echo "vector2 pytest exit $?"; tail -3 $out/pytest_vector2.log
timeout 200 python tools/reproj_probe.py > $out/reproj.json 2> $out/reproj.err; echo "reproj probe exit $?"; cat $out/reproj.json
if [ -n "$VARIANTS" ]; then bash tools/mv_variants.sh > $out/variants.txt 2>&1; cat $out/variants.txt; fi
if [ -z "$SKIP_NCU" ]; then
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-pose"
  timeout 300 $CMD > $out/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 500 --csv --log-file $out/launches.csv $CMD > $out/ncu_list.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on \
      -k regex:"k_cs_ring|k_ps|k_cs<" -s 30 -c 4 -f -o $out/prof_mv $CMD > $out/ncu_full_mv.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on \
      -k regex:"k_resjac|k_cam_blocks|k_point" -s 5 -c 7 -f -o $out/prof_lin $CMD > $out/ncu_full_lin.log 2>&1
  timeout 200 python tools/pose_probe.py > $out/pose.json 2> $out/pose.err &&
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_pose_lm" -s 3 -c 1 -f -o $out/prof_pose \
      python tools/pose_probe.py > $out/ncu_full_pose.log 2>&1
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_reproj_err" -c 2 -f -o $out/prof_reproj \
      python tools/reproj_probe.py venice1778 1 > $out/ncu_full_reproj.log 2>&1
  echo "ncu exit $?"; ls -la $out
fi
