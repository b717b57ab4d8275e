timeout 600 python -m pytest tests/test_metrics.py -x -q -m gpu 2>&1 | tail -5
python tools/bench_metrics.py
