# usage: r2_run_scale.sh N "configs..."   (one torchrun per config on N GPUs)
N=$1; shift
mkdir -p gpurun_out
port=29700
for c in "$@"; do
  port=$((port+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --config $c --steps 10 --warmup 3 --no-baselines > gpurun_out/r2_bench_cfg${c}_${N}gpu.json 2> gpurun_out/r2_bench_cfg${c}_${N}gpu.err; echo "cfg $c N=$N rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_cfg${c}_${N}gpu.json')); print('cfg $c N=$N', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('dp_check'), d['clocks'])"
done
