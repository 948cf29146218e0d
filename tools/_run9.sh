set -u
out=gpurun_out/r2m; mkdir -p $out
timeout 900 python bench.py --steps 100 --warmup 10 > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; tail -c 600 $out/bench_n1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2m/bench_n1.json'))
print('ms_per_step', d['ms_per_step'], 'frac', d['roofline']['frac'])
print(json.dumps(d['extras'].get('rotating_inputs_no_flush'), indent=1))
print('env', d['env_steps']['value'], d['env_steps']['ms_per_step'])
PY
