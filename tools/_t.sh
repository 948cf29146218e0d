python -m pytest tests -x -q -m gpu --timeout 240 2>&1 | tail -2
python tools/quick_times.py step1m
LYNX_B200_LIB=lynx_b200/lib/liblynx_b200_exp.so python tools/k2_timeline.py screen1m 147 2>&1 | tail -10
