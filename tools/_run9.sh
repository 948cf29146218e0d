set -u
out=gpurun_out/r2k; mkdir -p $out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 240 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_gpu.log
timeout 900 python bench.py --steps 100 --warmup 10 --no-extras > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; tail -c 600 $out/bench_n1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2k/bench_n1.json'))
for k in ('value','ms_per_step','e2e','api_e2e'):
    print(k, json.dumps(d.get(k))[:500])
PY
