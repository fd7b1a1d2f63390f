timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
c3() { timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200; }
echo ctl; ASB_LIB_PATH=$PWD/tools/ctl/libasb_head.so c3
echo new; c3
echo new ASB_VT=1; ASB_VT=1 c3
echo new; c3
timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -4 | cut -c1-330
