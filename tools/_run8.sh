set -u
out=gpurun_out/r2h; mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_parity_r2.py -x -q -m gpu --timeout 180 -k "light" > $out/pytest_light.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_light.log
python tools/prof_target.py --mode pbboth > $out/plain_pb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"parambeam" -c 4 -f -o $out/k6 python tools/prof_target.py --mode pbboth > $out/ncu_pb.log 2>&1; echo "ncu rc=$?"
