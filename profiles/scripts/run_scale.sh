# run_scale.sh N: the driver's launch of bench.py on N GPUs of one box -> gpurun_out/r02_bench_${N}gpu.json
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 16 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err
echo rc=$?; tail -c 600 gpurun_out/r02_bench_${N}gpu.json
