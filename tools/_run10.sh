python -m pytest tests -x -q -m gpu --timeout 180 2>&1 | tail -2
echo "== old"; LYNX_B200_LIB=lynx_b200/lib/liblynx_b200_old.so python tools/quick_times.py scr32m
echo "== new"; python tools/quick_times.py step1m scr32m cfg4
