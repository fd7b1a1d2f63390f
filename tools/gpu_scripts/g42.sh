timeout 900 python -m pytest tests/test_gpu_batch_ippe.py -m gpu -q -x 2>&1 | tail -2
ASB_BATCH_VERBOSE=1 timeout 900 python bench.py --no-extras --cpu-seconds 1 2>&1 | tail -3 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value',round(d['value']/1e6,2),'e2e',round(d['e2e']['value']/1e6,2),'valid',d['valid'])
    elif 'k_batch' in l: print(l.strip())"
timeout 300 python tools/batch_exp.py 2>&1 | tail -11
