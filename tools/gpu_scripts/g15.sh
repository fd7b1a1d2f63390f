mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
ASB_PDL=1 timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c1-230
ASB_CTA_PROF=1 ASB_PDL=1 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -14 | tee gpurun_out/g15_pipe1.log
ASB_CTA_PROF=1 ASB_PDL=0 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -14 | tee gpurun_out/g15_pipe0.log
