set -x
python -m pytest tests -m gpu -q 2>&1 | tail -30
python bench.py --steps 1 --warmup 1 2>&1 | tail -20
