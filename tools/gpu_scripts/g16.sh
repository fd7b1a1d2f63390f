mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
for t in 0 2 0 2 3; do echo "ASB_TAIL=$t"; ASB_TAIL=$t timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c90-200; done
ASB_CTA_PROF=1 ASB_TAIL=2 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -14 | tee gpurun_out/g16_pipe.log
