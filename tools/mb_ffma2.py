import ctypes, sys
sys.path.insert(0, '.')
from deepmetv2_b200._lib import lib
l = lib()
for name in ("gm_bench_ffma", "gm_bench_ffma2"):
    ms, gf = ctypes.c_float(), ctypes.c_double()
    getattr(l, name)(200, ctypes.byref(ms), ctypes.byref(gf))
    print(name, ms.value, "ms", gf.value / ms.value, "GFLOP/ms = TFLOP/s:", gf.value / ms.value / 1e3 * 1e3 / 1e3)
