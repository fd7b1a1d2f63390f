mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -8 > gpurun_out/g14_pytest.log; cat gpurun_out/g14_pytest.log
ASB_PIPELINE=0 timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c1-500 | tee gpurun_out/g14_c3_p0.json
ASB_PIPELINE=1 timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c1-500 | tee gpurun_out/g14_c3_p1.json
timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -8 | tee gpurun_out/g14_pipe.log
timeout 900 python bench.py > gpurun_out/g14_bench.json 2> gpurun_out/g14_bench.err; tail -c 2500 gpurun_out/g14_bench.json; tail -3 gpurun_out/g14_bench.err
