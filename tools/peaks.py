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
for mode, name in ((0, "write-only"), (1, "read-only"), (2, "1 read : 2 writes")):
    g = C.c_double()
    _lib.check(lib.asb_stream_bandwidth(0, 1 << 30, mode, C.byref(g)))
    print(f"stream {name}: {g.value:.0f} GB/s")
for tile in (64, 128):
    for cps in (1, 2, 4):
        u = C.c_double()
        _lib.check(lib.asb_tilepattern_bandwidth(0, 6010, tile, cps, C.byref(u)))
        print(f"sweep memory pattern, no math, tile {tile}, {cps} x 512 threads / SM: {u.value:.1f} us  "
              f"({12 * 6010 ** 2 / u.value / 1e3:.0f} GB/s of 12 N^2 bytes)")
