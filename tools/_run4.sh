set -u
out=gpurun_out/r2d; mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_r2.py tests/test_zz_no_alloc.py -x -q -m gpu > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $out/pytest.log
timeout 100 python tools/quick_times.py step1m scr32m cfg4 2>&1 | grep -v "cfg4 12.5M out" > $out/quick_default.txt; cat $out/quick_default.txt
export LYNX_B200_LIB=$PWD/lynx_b200/lib/liblynx_b200_exp.so
for ms in 2 3 4; do for mw in 6000 12000 20000 40000; do
  echo "== MIN_STAGES=$ms MAX_WORDS=$mw"; LYNX_TILE_MIN_STAGES=$ms LYNX_TILE_MAX_WORDS=$mw LYNX_TILE_PXPP=8 timeout 100 python tools/quick_times.py step1m scr32m cfg4 2>&1 | grep -v "persistent\|cfg4 12.5M out"
done; done > $out/sweep.txt 2>&1
cat $out/sweep.txt
