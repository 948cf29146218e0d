set -u
N=$1
out=gpurun_out/r2n; mkdir -p $out
LYNX_EX_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 100 --warmup 10 > $out/bench_n$N.json 2> $out/bench_n$N.err; echo "bench n$N rc=$?"; grep "rank 0" $out/bench_n$N.err | tail -4; grep -i "Traceback" -A8 $out/bench_n$N.err | head -20
python - <<PY
import json
d=json.load(open('$out/bench_n$N.json'))
for k in ('value','ms_per_step','e2e','exchange_check','exchange_check_exchanges','env_steps','api_e2e','cfg4_100M_sharded','gpu_launches'):
    print(k, json.dumps(d.get(k))[:420])
PY
