#!/bin/bash
# Batched kernel, one GPU: parity tests + device-resident / host-buffer rates per kernel variant (EKF_BATCH_VARIANT).
#   /usr/local/graft/bin/gpurun -- 'bash tools/scripts/batch_variants.sh TAG 6 14'
T=${1:-rX}; shift; O=gpurun_out
for V in "$@"; do
  EKF_BATCH_VARIANT=$V timeout 300 python -m pytest tests/test_gpu_batch.py -x -q 2>&1 | tail -2
  EKF_BATCH_VARIANT=$V timeout 300 python bench.py --steps 10 --warmup 3 --no-n60k --no-small --no-cpu-baseline --batch-steps 20 \
    > $O/${T}_batch_v$V.json 2> $O/${T}_batch_v$V.err; echo "variant $V bench rc=$?"
  python - <<PY
import json
d = json.load(open("$O/${T}_batch_v$V.json"))["batched"]
print("variant $V:", d["value"], d["e2e"]["value"], {k: d[k] for k in d if k in ("ms_per_step", "roofline")})
PY
done
