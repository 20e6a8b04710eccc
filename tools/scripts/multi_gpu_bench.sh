N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/r1f_bench$N.json 2> gpurun_out/r1f_bench$N.err; tail -c 400 gpurun_out/r1f_bench$N.json | head -c 400; echo
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --landmarks 60000 --no-batched --no-cpu-baseline > gpurun_out/r1f_bench${N}_60k.json 2> gpurun_out/r1f_bench${N}_60k.err; tail -c 300 gpurun_out/r1f_bench${N}_60k.json
