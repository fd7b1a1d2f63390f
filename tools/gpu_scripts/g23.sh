mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_batch_ippe.py -m gpu -q -x 2>&1 | tail -3
timeout 900 python bench.py --no-extras --cpu-seconds 1 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value',d['value'],'e2e',d['e2e']['value'],'valid',d['valid'])"
ASB_BATCH_GROUPS=1 timeout 900 python bench.py --no-extras --cpu-seconds 1 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('groups=1 value',d['value'],'e2e',d['e2e']['value'],'valid',d['valid'])"
ASB_BATCH_GROUPS=4 timeout 900 python bench.py --no-extras --cpu-seconds 1 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('groups=4 value',d['value'],'e2e',d['e2e']['value'],'valid',d['valid'])"
timeout 300 python tools/batch_exp.py 2>&1 | tail -12
