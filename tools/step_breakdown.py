"""Where a resident somatic-standard step spends its host time: the C call alone against the Python wrapper around it."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.getcwd())
from guacamole_b200 import native
from guacamole_b200.native import Context

ctx = Context(0)
L = 50_000_000
ctx.synth_generate(L, (30, 60), (False, True), 3000)
params = native.ss_params()
r = native.make_ranges([(0, 0, L)])
for i in range(5):
    calls, n, arena_p, alen, nv = C.POINTER(native.SSCallC)(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_int64()
    t0 = time.perf_counter()
    st = ctx.lib.guac_somatic_standard(ctx.ctx, C.byref(params), r, C.c_int32(1), C.byref(calls), C.byref(n), C.byref(arena_p),
                                       C.byref(alen), None, C.byref(nv))
    t1 = time.perf_counter()
    ctx.lib.guac_free(ctx.ctx, arena_p)
    ctx.lib.guac_free(ctx.ctx, C.cast(calls, C.c_void_p))
    t2 = time.perf_counter()
    out = ctx.somatic_standard([(0, 0, L)])
    t3 = time.perf_counter()
    print(f"C call {1e3 * (t1 - t0):.3f} ms, frees {1e3 * (t2 - t1):.3f} ms, wrapper call {1e3 * (t3 - t2):.3f} ms, kernel {ctx.last_kernel_ms()[0]:.3f} ms, st {st}")
