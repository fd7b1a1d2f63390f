mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py tests/test_gpu_batch_ippe.py -m gpu -q -x 2>&1 | tail -3
ASB_FULL_P=1 timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200
timeout 300 python bench.py --only c3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 us/frame', d['us_per_frame'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'])"
ASB_CTA_PROF=1 timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -8 | cut -c1-330
