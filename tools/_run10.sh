export LYNX_B200_LIB=lynx_b200/lib/liblynx_b200_exp.so
python tools/k2_timeline.py screen1m 147 | tail -11
