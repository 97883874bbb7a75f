#!/bin/bash
# Round-2 first GPU call: jaxopt probe, GPU tests (incl. the gated vector2 tests), bench line, pose + evaluator probes with ncu.
out=gpurun_out/r2a
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
python -c "import jax, jaxopt; print('JAXOPT', jaxopt.__version__, jax.__version__)" > $out/jaxopt_probe.txt 2>&1; echo "jaxopt probe exit $?" >> $out/jaxopt_probe.txt
cat $out/jaxopt_probe.txt | tail -2
B200BA_EXPERIMENTAL=1 timeout 600 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
tail -15 $out/pytest.log
timeout 300 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -c 400 $out/bench.json
timeout 200 python tools/reproj_probe.py > $out/reproj.json 2> $out/reproj.err; echo "reproj probe exit $?"; cat $out/reproj.json
timeout 200 python tools/pose_probe.py > $out/pose.json 2> $out/pose.err; echo "pose probe exit $?"; cat $out/pose.json
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_pose_lm" -s 3 -c 1 -f -o $out/prof_pose \
    python tools/pose_probe.py > $out/ncu_full_pose.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_reproj_err" -c 2 -f -o $out/prof_reproj \
    python tools/reproj_probe.py venice1778 1 > $out/ncu_full_reproj.log 2>&1
ls -la $out
