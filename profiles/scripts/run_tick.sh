# run_tick.sh N...: the full-tick leg (BASELINE configs[4]) of bench.py on N GPUs -> gpurun_out/r02_tick_${N}gpu.json
for N in "$@"; do
  if [ $N = 1 ]; then python bench.py --gpus 1 --steps 16 --warmup 3 --tick --no-cpu --no-extras > gpurun_out/r02_tick_${N}gpu.json 2> gpurun_out/r02_tick_${N}gpu.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 16 --warmup 3 --tick --no-cpu --no-extras > gpurun_out/r02_tick_${N}gpu.json 2> gpurun_out/r02_tick_${N}gpu.err; fi
  python -c "
import json; d=json.load(open('gpurun_out/r02_tick_${N}gpu.json')); t=d['tick']; print($N, t['value'], t['ms_per_tick'], {k:v for k,v in t.items() if 'phase' in k or 'ms' in k})"
done
