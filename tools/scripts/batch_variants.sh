#!/bin/bash
# Development aid: parity tests + timing of every ekf_batch_kernel variant (EKF_BATCH_VARIANT bits:
# 1 = 2x2-block downdate, 2 = one-pass association, 4 = 20 CTAs/SM).  Output: gpurun_out/batch_variants.log
mkdir -p gpurun_out
out=gpurun_out/batch_variants.log
: > $out
for v in ${VARIANTS:-0 1 2 3 4 5 6 7}; do
  echo "== variant $v" >> $out
  EKF_BATCH_VARIANT=$v timeout 300 python -m pytest tests/test_gpu_batch.py -m gpu -x -q 2>&1 | tail -2 >> $out
  EKF_BATCH_VARIANT=$v timeout 300 python tools/batch_bench.py --steps 20 --warmup 4 2>&1 | tail -1 >> $out
  EKF_BATCH_VARIANT=$v timeout 300 python tools/batch_bench.py --m 5 --steps 20 --warmup 4 2>&1 | tail -1 >> $out
done
cat $out
