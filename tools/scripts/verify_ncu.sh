#!/bin/bash
# The two ncu passes of verify_round.sh alone.  Usage: verify_ncu.sh TAG
TAG=${1:-rX}
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_N10k.csv \
  python bench.py --steps 20 --warmup 3 --batch-steps 2 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"ekf_(seq|downdate|batch)_kernel" -s 6 -c 6 -f \
  -o gpurun_out/${TAG}_ncu_full python bench.py --steps 1 --warmup 3 --batch-steps 1 --no-cpu-baseline \
  > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log
