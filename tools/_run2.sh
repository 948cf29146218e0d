set -u
out=gpurun_out/r2b; mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $out/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee $out/rc.txt
tail -3 $out/noalloc_alone.log
timeout 600 python -m pytest tests -x -q -m gpu --timeout 120 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/rc.txt
tail -3 $out/pytest_gpu.log
timeout 200 python tools/quick_times.py step1m scr32m cfg4 k6 > $out/quick.txt 2>&1; cat $out/quick.txt
