timeout 900 python -m pytest tests/test_gpu_batch_ippe.py -m gpu -q -x 2>&1 | tail -3
run() { timeout 900 python bench.py --no-extras --cpu-seconds 1 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value',round(d['value']/1e6,2),'e2e',round(d['e2e']['value']/1e6,2),'valid',d['valid'])"; }
echo fused; run
echo nofuse; ASB_BATCH_NOFUSE=1 run
timeout 300 python tools/batch_exp.py 2>&1 | tail -11
