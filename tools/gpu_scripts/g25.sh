mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_batch_ippe.py -m gpu -q -x --durations=5 2>&1 | tail -12
