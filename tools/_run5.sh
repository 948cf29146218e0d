set -u
out=gpurun_out/r2e; mkdir -p $out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 120 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu.log
timeout 200 python tools/quick_times.py step1m scr32m out32m cfg4 > $out/quick.txt 2>&1; cat $out/quick.txt
python tools/prof_target.py --mode screen32m --steps 3 > $out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:track_pipe -s 2 -c 1 -f -o $out/k2_screen32m python tools/prof_target.py --mode screen32m --steps 3 > $out/ncu.log 2>&1; echo "ncu rc=$?"
