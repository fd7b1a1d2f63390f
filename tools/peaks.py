import ctypes as C, sys
sys.path.insert(0, ".")
from aruco_slam_b200 import _lib
lib = _lib.lib()
for w, name in ((0, "DFMA"), (1, "DMMA"), (2, "mixed")):
    v = C.c_double()
    _lib.check(lib.asb_fp64_peak(0, w, C.byref(v)))
    print(name, v.value, "TFLOP/s", flush=True)
g = C.c_double()
_lib.check(lib.asb_copy_bandwidth(0, 1 << 30, C.byref(g)))
print("copy", g.value, "GB/s")
lat = (C.c_double * 4)()
_lib.check(lib.asb_latency_probe(0, lat))
print("latency clocks [DFMA, rsqrt+add, LDS+cvt, BAR.SYNC(256)]:", [round(x, 1) for x in lat])
