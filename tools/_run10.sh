python -m pytest tests -x -q -m gpu --timeout 240 2>&1 | tail -4
python tools/quick_times.py k6
