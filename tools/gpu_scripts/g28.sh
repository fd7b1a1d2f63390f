mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --only c3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 us/frame', d['us_per_frame'], 'e2e', d['e2e'])"
bash tools/gpu_scripts/g24.sh 2>&1 | tail -2
