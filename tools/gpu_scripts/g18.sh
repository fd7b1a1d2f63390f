mkdir -p gpurun_out
c3() { timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200; }
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -3
echo "HEAD lib"; ASB_LIB_PATH=$PWD/tools/ctl/libasb_head.so c3
echo "new"; c3
echo "new"; c3
timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -5 | tee gpurun_out/g18_pipe.log
