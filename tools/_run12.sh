set -u
out=gpurun_out/r2l; mkdir -p $out
m=$1
python tools/prof_target.py --mode $m --steps 3 > $out/plain_$m.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"track_pipe|compose|parambeam" -s 4 -c 2 -f -o $out/k_$m python tools/prof_target.py --mode $m --steps 3 > $out/ncu_$m.log 2>&1; echo "ncu $m rc=$?"
