#!/bin/bash
out=gpurun_out/r2x
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_vector2.py -m gpu -q -x > $out/v2.log 2>&1; echo "v2 exit $?" >> $out/v2.log
tail -30 $out/v2.log
timeout 1200 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" $out/pytest.log | tail -12
