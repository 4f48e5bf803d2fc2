"""Development aid: where ss_tile_kernel's SM clocks go, phase by phase (build with -DGUAC_TK_PHASE_CLOCKS, see tile_kernels.cuh).
    GUAC_B200_LIB=guacamole_b200/csrc/variants/phase.so python tools/phase_clocks.py [joint|ss|ss1] [mbp]
Prints the share of every phase in the summed per-CTA clocks of one call (thread 0 of each CTA, barrier to barrier)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from guacamole_b200.commands import somatic_joint as sj  # noqa: E402
from guacamole_b200.native import Context  # noqa: E402

NAMES = ["init + tile/ref loads", "A: reads -> coverage differences + events", "slow reads (exact cursor)", "B: scans", "bucket events by locus",
         "C: one thread per locus", "joint: dense event scan", "joint: prune pass (incl. barrier)" , "ss: candidate pass"]


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "joint"
    n = int(float(sys.argv[2]) * 1e6) if len(sys.argv) > 2 else 20_000_000
    ctx = Context(0)
    out = (C.c_ulonglong * 32)()
    if what == "joint":
        ctx.synth_generate(n, (30, 30, 60, 60), (False, False, True, True), 4000)
        p = sj.sj_params([sj.NORMAL_DNA, sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA])
        call = lambda: ctx.somatic_joint_records([(0, 0, n)], p, copy=False)
    else:
        ctx.synth_generate(n, (30, 60), (False, True), 3000, profile=1 if what == "ss1" else 0)
        call = lambda: ctx.somatic_standard([(0, 0, n)])
    call()
    ctx.lib.guac_debug_phase_clocks(out)
    call()
    ctx.lib.guac_debug_phase_clocks(out)
    tot = sum(out[:9])
    if out[9]:
        print('  per tile: scanned loci %.1f, prune candidates %.1f, events %.1f, reads %.1f' % tuple(out[k] / out[9] for k in (10, 11, 12, 13)))
    print(what, n, "kernel_ms", ctx.last_kernel_ms()[0])
    print('  inside the prune pass (warp 0): scan %.1f %%, event loop %.1f %% of all phases' % (out[14] / tot * 100, out[15] / tot * 100))
    for k, name in enumerate(NAMES):
        if k == 7 and what != "joint":
            name = "(end of C)"
        print(f"  {out[k] / tot * 100:5.1f} %  {name}")
    site_report(out)


SITE = ["read-range searches + reference", "span loop (one lane per read)", "slow reads + bail tests", "lane totals + cells",
        "discovery + slot sums", "allele evaluation + records"]


def site_report(out):
    tot = sum(out[16:22])
    if tot:
        print("  site kernel (warp 0 of every CTA; sites that end early count only the phases they reach):")
        for k, name in enumerate(SITE):
            print(f"    {out[16 + k] / tot * 100:5.1f} %  {name}")


if __name__ == "__main__":
    main()
