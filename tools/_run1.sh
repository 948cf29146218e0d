set -u
out=gpurun_out/r2a; mkdir -p $out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 120 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $out/rc.txt
tail -5 $out/pytest_gpu.log
timeout 200 python tools/quick_times.py step1m scr32m out32m cfg4 > $out/quick_default.txt 2>&1; echo "quick rc=$?" | tee -a $out/rc.txt
cat $out/quick_default.txt
(for a in "32000000 40" "1000000 40" "12500000 15"; do timeout 90 ./tools/micro/hist_bench2 $a; done) > $out/hist_bench_c.txt 2>&1
