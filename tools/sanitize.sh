#!/bin/bash
# compute-sanitizer over the smallest cases that cover every kernel family: run under gpurun,
# ONE tool per call (B200_PROFILING.md).  usage: tools/sanitize.sh memcheck|racecheck
#   smoke()            single-filter engine (k_f1 / k_f2 / k_sweep with PDL), resident batch (k_batch), IPPE
#   wide + 2-chunk     tests/test_gpu_wide.py -k "rot-scalar_first-32" (two chunks per frame, all four wide kernels)
#                      and a k_batch case with more than one 4-marker chunk
set -e
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanitize_plain.log 2>&1
compute-sanitizer --tool $TOOL --error-exitcode 1 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanitize_${TOOL}_smoke.log 2>&1 || echo "smoke: sanitizer reported errors"
compute-sanitizer --tool $TOOL --error-exitcode 1 python -m pytest tests/test_gpu_wide.py tests/test_gpu_batch_ippe.py -q -x -k "test_wide_equals_resident_engine or test_batch_capacity_flag or (test_wide_matches_oracle and rot-scalar_first-6)" > gpurun_out/sanitize_${TOOL}_batch.log 2>&1 || echo "batch: sanitizer reported errors"
tail -n 12 gpurun_out/sanitize_${TOOL}_smoke.log gpurun_out/sanitize_${TOOL}_batch.log
