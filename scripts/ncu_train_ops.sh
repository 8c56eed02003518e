#!/bin/bash
# ncu --set full captures (warm cache, one launch each) of the per-step kernels of the training recurrence, driven by
# scripts/train_op_timing.py.  Usage (under gpurun): bash scripts/ncu_train_ops.sh <tag>
TAG=${1:-rXX}
for spec in "attn_bwd_fused:attn_bwd_row_kernel:row" "attn_fwd_fused:attn_fwd_fused_kernel:fwd" "gemm_pair_acc:skinny_acc_kernel:skinny"; do
  IFS=: read only kern tag <<< "$spec"
  TRAIN_OP_ONLY=$only timeout 300 ncu --set full --import-source on --cache-control none --clock-control none -k regex:$kern --launch-skip 8 -c 1 \
      -o gpurun_out/${TAG}_op_$tag -f python scripts/train_op_timing.py 12 > gpurun_out/${TAG}_ncu_op_$tag.log 2>&1
done
ls -la gpurun_out/${TAG}_op_*.ncu-rep
