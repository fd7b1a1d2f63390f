mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -4
timeout 300 python tools/sweep_exp.py 0 1 6 14 2>&1 | tail -16
timeout 600 python bench.py --only c3 2>&1 | tail -1 | cut -c1-300
