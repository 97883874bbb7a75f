#!/bin/bash
# 2-GPU: multi-rank tests, then Venice / Final-13682 with the PDL masks (0 none, 2 ring only, 14 + split kernels, 15 + point kernel)
out=gpurun_out/r3e
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $out/multi.log 2>&1; echo "multi exit $?" >> $out/multi.log; tail -4 $out/multi.log
for w in venice1778 final13682; do
for pdl in 0 14 15 0 14; do
  echo "== $w PDL=$pdl"
  B200BA_PDL=$pdl timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus 2 --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-extras --workload $w 2> $out/err_${w}_$pdl.txt | grep "^{" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],2), 'e2e', round(d['e2e']['value'],2), {a:(b['launches'], round(b['avg_us'],1)) for a,b in k.items() if a in ('mv_pt','mv_cam','pcg_vec')}, d['extra'].get('parity',{}).get('max_rel_loss_dev'))"
done
done 2>&1 | tee $out/pdl2.txt
