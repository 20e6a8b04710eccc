#!/bin/bash
# The round's 1-GPU measurement set; run on the GPU box: /usr/local/graft/bin/gpurun -- 'bash tools/scripts/measure_1gpu.sh r2j'
T=${1:-rX}; O=gpurun_out
python bench.py > $O/${T}_bench1.json 2> $O/${T}_bench1.err; echo bench rc=$?
python bench.py --impl reference --steps 20 --warmup 5 > $O/${T}_reference_arm.json 2>/dev/null; echo ref rc=$?
python bench.py --m 5 --steps 100 --no-n60k --no-batched --no-small --no-cpu-baseline > $O/${T}_bench1_m5.json 2>/dev/null
python bench.py --m 10 --steps 100 --no-n60k --no-batched --no-small --no-cpu-baseline > $O/${T}_bench1_m10.json 2>/dev/null
python tools/seq_soak.py --scans 100000 --out $O/${T}_seq_soak.json > $O/${T}_seq_soak.log 2>&1; echo soak rc=$?; tail -1 $O/${T}_seq_soak.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_bench_N10k.csv \
  python bench.py --steps 20 --warmup 3 --no-n60k --no-batched --no-small --no-cpu-baseline > $O/${T}_ncu.log 2>&1; echo ncu rc=$?
