set -u
out=gpurun_out/r2l; mkdir -p $out
python tools/prof_target.py --mode pbboth > $out/plain_pb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"parambeam" -c 4 -f -o $out/k_pbboth python tools/prof_target.py --mode pbboth > $out/ncu_pb.log 2>&1; echo "ncu rc=$?"
