set -u
out=gpurun_out/r2h; mkdir -p $out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 180 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_gpu.log
timeout 200 python tools/quick_times.py > $out/quick.txt 2>&1; cat $out/quick.txt
python tools/_k1_decomp.py
