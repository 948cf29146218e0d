python -m pytest tests -x -q -m gpu --timeout 240 2>&1 | tail -2
python tools/quick_times.py step1m scr32m cfg4
