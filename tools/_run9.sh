set -u
out=gpurun_out/r2j; mkdir -p $out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 180 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu.log
timeout 300 python tools/quick_times.py > $out/quick.txt 2>&1; cat $out/quick.txt
timeout 900 python bench.py --steps 100 --warmup 10 > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; tail -c 600 $out/bench_n1.err
