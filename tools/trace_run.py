"""Development timing of guac_somatic_standard on the bench workload:  python tools/trace_run.py [ss_impl] [profile]"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
from guacamole_b200.native import Context  # noqa: E402

ctx = Context(0)
L = 50_000_000
profile = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ctx.synth_generate(L, (30, 60), (False, True), 3000, profile=profile)
if len(sys.argv) > 1:
    ctx.set_option("ss_impl", int(sys.argv[1]))
for i in range(4):
    t = time.perf_counter()
    out = ctx.somatic_standard([(0, 0, L)])
    dt = time.perf_counter() - t
    print("step", i, dt * 1e3, "ms kernel", ctx.last_kernel_ms()[0], "heavy", ctx.last_stat("heavy_loci"), "fallback", ctx.last_stat("fallback_tiles"),
          "calls", len(out["calls"]), file=sys.stderr)
