#!/bin/bash
# Round-end verification on one B200 (run under gpurun): GPU test suite, bench (ours + reference arm), ncu launch
# list and one `ncu --set full` capture of the three hot kernels.  Usage: verify_round.sh TAG   -> gpurun_out/TAG_*
TAG=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_gpu.log
python bench.py > gpurun_out/${TAG}_bench1.json 2> gpurun_out/${TAG}_bench1.err || tail -5 gpurun_out/${TAG}_bench1.err
tail -c 600 gpurun_out/${TAG}_bench1.json; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_reference_arm.json 2> gpurun_out/${TAG}_reference_arm.err
tail -c 300 gpurun_out/${TAG}_reference_arm.json; echo
for m in 5 10; do
  python bench.py --m $m --steps 100 --no-cpu-baseline --batch-m $m > gpurun_out/${TAG}_bench1_m$m.json 2>/dev/null
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_N10k.csv \
  python bench.py --steps 20 --warmup 3 --batch-steps 2 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"ekf_(seq|downdate|batch)_kernel" -s 6 -c 6 -f \
  -o gpurun_out/${TAG}_ncu_full python bench.py --steps 1 --warmup 3 --batch-steps 1 --no-cpu-baseline \
  > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log
