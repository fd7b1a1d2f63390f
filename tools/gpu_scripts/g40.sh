timeout 300 python tools/sweep_exp.py 0 1 2 3 2>&1 | grep "^config"
ASB_SWEEP_CONFIG=3 timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
c3() { timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200; }
echo dual; c3
echo tri; ASB_SWEEP_CONFIG=3 c3
echo tri tail3; ASB_SWEEP_CONFIG=3 ASB_TAIL=3 c3
echo tri tail1; ASB_SWEEP_CONFIG=3 ASB_TAIL=1 c3
