mkdir -p gpurun_out
c3() { timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200; }
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
echo "HEAD lib"; ASB_LIB_PATH=$PWD/tools/ctl/libasb_head.so c3
for t in 0 1 2 3 4; do echo "new TAIL=$t"; ASB_TAIL=$t c3; done
echo "new TAIL=0 PDL=0"; ASB_TAIL=0 ASB_PDL=0 c3
echo "new TAIL=2 PDL=0"; ASB_TAIL=2 ASB_PDL=0 c3
ASB_CTA_PROF=1 ASB_TAIL=2 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -14 | tee gpurun_out/g17_pipe.log
