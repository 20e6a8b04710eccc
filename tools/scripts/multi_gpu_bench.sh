#!/bin/bash
# Round's multi-GPU set on N GPUs of one box: /usr/local/graft/bin/gpurun --gpus N -- 'bash tools/scripts/multi_gpu_bench.sh N TAG'
N=$1; T=${2:-rX}; O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/shard_peer_check.py 2>&1 | tail -1
EKF_SHARD_DROP_GUESS=5 timeout 300 $TR tools/shard_peer_check.py 2>&1 | tail -1
timeout 600 $TR bench.py --gpus $N > $O/${T}_bench$N.json 2> $O/${T}_bench$N.err; echo bench rc=$?
python - <<PY
import json
d=json.load(open("$O/${T}_bench$N.json"))
print(d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["n60k"]["value"], d["n60k"]["e2e"]["value"], d["batched"]["value"], d["batched"]["e2e"]["value"], d["run"]["association_bit_exact_vs_expected_ids"])
PY
