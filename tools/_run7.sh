set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo rc=$?
tail -c 3000 gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_n1.json'))
for k in ('value','ms_per_step','e2e','roofline','env_steps','api_e2e','cfg4_100M_sharded','cpu_baseline','gpu_launches','clocks'):
    print(k, json.dumps(d.get(k))[:900])
PY
