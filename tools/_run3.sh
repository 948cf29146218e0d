set -u
mkdir -p gpurun_out/r2c
for m in screen1m cfg4screen screen32m; do
  python tools/prof_target.py --mode $m --steps 3 > gpurun_out/r2c/plain_$m.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:track_pipe -s 2 -c 1 -f -o gpurun_out/r2c/k2_$m python tools/prof_target.py --mode $m --steps 3 > gpurun_out/r2c/ncu_$m.log 2>&1
  echo "$m rc=$?"
done
