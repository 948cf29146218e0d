#!/usr/bin/env bash
# One GPU-box call that re-measures everything the docs quote (about 3 min of box time):
#   gpurun --timeout 900 -- 'bash tools/gpu_round_check.sh'
# Outputs land in gpurun_out/round_check/ and are merged back by gpurun.
set -u
out=gpurun_out/round_check
mkdir -p "$out"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > "$out/smoke.log" 2>&1; echo "smoke rc=$?" | tee "$out/rc.txt"
timeout 300 python -m pytest tests -x -q -m gpu --timeout 240 > "$out/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$out/rc.txt"
timeout 300 python bench.py > "$out/bench_n1.json" 2> "$out/bench_n1.err"; echo "bench rc=$?" | tee -a "$out/rc.txt"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > "$out/bench_reference.json" 2> "$out/bench_reference.err"; echo "reference arm rc=$?" | tee -a "$out/rc.txt"
timeout 120 python tools/quick_times.py > "$out/quick_times.txt" 2>&1; echo "quick rc=$?" | tee -a "$out/rc.txt"
timeout 60 python tools/track_call_overhead.py > "$out/track_call_overhead.txt" 2>&1; echo "overhead rc=$?" | tee -a "$out/rc.txt"
timeout 60 python tools/e2e_breakdown.py > "$out/e2e_breakdown.txt" 2>&1; echo "e2e rc=$?" | tee -a "$out/rc.txt"
timeout 60 python tools/determinism_check.py 4200000 > "$out/determinism.txt" 2>&1; echo "determinism rc=$?" | tee -a "$out/rc.txt"
tail -2 "$out/pytest_gpu.log"; cat "$out/quick_times.txt"; tail -1 "$out/track_call_overhead.txt"; tail -2 "$out/determinism.txt"
