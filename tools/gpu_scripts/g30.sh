timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
ASB_FRONT_DBG=32 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -4
