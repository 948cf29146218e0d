#!/usr/bin/env bash
# One ncu --set full capture of a tools/prof_target.py mode (one ncu per gpurun call, after a plain run):
#   gpurun --timeout 600 -- 'bash tools/ncu_capture.sh screen1m [launch-skip] [launch-count]'
# -> gpurun_out/ncu/k_<mode>.ncu-rep; summarise here with tools/ncu_summary.py / ncu_stalls.py / ncu_lines.py
set -u
m=$1; skip=${2:-4}; cnt=${3:-2}
out=gpurun_out/ncu; mkdir -p $out
python tools/prof_target.py --mode $m --steps 3 > $out/plain_$m.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"track_pipe|compose|parambeam|adj_|beam_|aperture" -s $skip -c $cnt -f -o $out/k_$m python tools/prof_target.py --mode $m --steps 3 > $out/ncu_$m.log 2>&1; echo "ncu $m rc=$?"
