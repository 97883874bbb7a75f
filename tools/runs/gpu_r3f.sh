#!/bin/bash
# 2-GPU: warp-per-camera fused kernel: tests, then Venice / Final-13682 three-launch vs warpx
out=gpurun_out/r3f
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_ba.py -m gpu -q -x -k "warp_per_camera" > $out/single.log 2>&1; echo "single exit $?" >> $out/single.log; tail -5 $out/single.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -k "warp_per_camera or point_sharded" > $out/multi.log 2>&1; echo "multi exit $?" >> $out/multi.log; tail -12 $out/multi.log
for w in venice1778 final13682; do
for wx in 0 1 0 1; do
  echo "== $w WARPX=$wx"
  B200BA_WARPX=$wx timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus 2 --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-extras --workload $w 2> $out/err_${w}_$wx.txt | grep "^{" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],2), 'e2e', round(d['e2e']['value'],2), {a:(b['launches'], round(b['avg_us'],1)) for a,b in k.items() if a in ('mv_pt','mv_cam','pcg_vec')}, d['extra'].get('parity',{}).get('max_rel_loss_dev'))"
done
done 2>&1 | tee $out/warpx2.txt
