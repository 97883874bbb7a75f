#!/bin/bash
# 8 GPUs: three-launch sequence vs warp-per-camera fused kernel (pipelined / not)
out=gpurun_out/r3h
mkdir -p $out
run() {  # workload, env...
  w=$1; shift
  env "$@" timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus 8 --steps 9 --warmup 3 --no-pose --no-cpu-baseline --no-extras --no-e2e --workload $w 2> $out/err_$w.txt | grep "^{" | tee $out/last_$w.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['extra']['kernels']; print(round(d['value'],2), {a:(b['launches'], round(b['avg_us'],1)) for a,b in k.items() if a in ('mv_pt','mv_cam','pcg_vec')}, d['extra'].get('parity',{}).get('max_rel_loss_dev'))"
}
for w in venice1778 final13682; do
  echo "== $w three-launch"; run $w B200BA_WARPX=0
  echo "== $w warpx pipe"; run $w B200BA_WARPX=1 B200BA_WARPX_PIPE=1 B200BA_XCHG_TIMEOUT_MS=500
  echo "== $w warpx nopipe"; run $w B200BA_WARPX=1 B200BA_WARPX_PIPE=0 B200BA_XCHG_TIMEOUT_MS=500
done 2>&1 | tee $out/warpx8.txt
