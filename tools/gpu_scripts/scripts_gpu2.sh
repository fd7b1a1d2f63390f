set -x
mkdir -p gpurun_out
python - <<'PY'
import ctypes as C
from aruco_slam_b200 import _lib
lib=_lib.lib()
for w in (0,1,2):
    v=C.c_double(); _lib.check(lib.asb_fp64_peak(0,w,C.byref(v))); print("fp64_peak which",w,v.value,"TFLOP/s",flush=True)
v=C.c_double(); _lib.check(lib.asb_copy_bandwidth(0,1<<30,C.byref(v))); print("copy GB/s",v.value)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 700 -c 150 --csv --log-file gpurun_out/launches_c3_full.csv python bench.py --only c3 > gpurun_out/ncu_c3b.log 2>&1
tail -2 gpurun_out/ncu_c3b.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep --launch-skip 130 -c 1 -o gpurun_out/sweep_r1a -f python bench.py --only c3 > gpurun_out/ncu_c3c.log 2>&1
tail -2 gpurun_out/ncu_c3c.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_batch -c 1 -o gpurun_out/batch_r1a -f python bench.py --steps 1 --warmup 0 --no-extras --seqs 1024 --frames 200 > gpurun_out/ncu_c2.log 2>&1
tail -2 gpurun_out/ncu_c2.log
ls -la gpurun_out
