"""Development timing of guac_somatic_joint on BASELINE configs[3]'s per-GPU shape (2 normal + 2 tumor, 30/30/60/60).
    python tools/bench_joint.py [--mbp 50] [--steps 3] [--impl 2]     (GUAC_TRACE=1 prints the host-side phases)
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from guacamole_b200.commands import somatic_joint as sj  # noqa: E402
from guacamole_b200.native import Context  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--mbp", type=float, default=50.0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--impl", type=int, default=2)
    ap.add_argument("--depths", default="30,30,60,60")
    a = ap.parse_args()
    depths = tuple(int(x) for x in a.depths.split(","))
    n = int(a.mbp * 1e6)
    tumor = tuple(i >= len(depths) // 2 for i in range(len(depths)))
    kinds = [sj.TUMOR_DNA if t else sj.NORMAL_DNA for t in tumor]
    ctx = Context(0)
    ctx.synth_generate(n, depths, tumor, 4000)
    ctx.set_option("sj_impl", a.impl)
    p = sj.sj_params(kinds)
    out = ctx.somatic_joint_records([(0, 0, n)], p, copy=False)
    torch.cuda.synchronize()
    ks = []
    t0 = time.perf_counter()
    for _ in range(a.steps):
        out = ctx.somatic_joint_records([(0, 0, n)], p, copy=False)
        ks.append(ctx.last_kernel_ms()[0])
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / a.steps
    D, S = sum(depths), len(depths)
    bytes_per_locus = 1.615 * D + 1 + 8 * S
    k = sum(ks) / len(ks)
    print(json.dumps({"path": "somatic-joint", "depths": depths, "loci": n, "ms_per_step": dt * 1e3, "kernel_ms": k,
                      "loci_per_s": n / dt, "alleles": int(len(out[0])), "visited": int(out[3]),
                      "frac_of_6552": bytes_per_locus * out[3] / (k / 1e3) / 1e9 / 6552.6,
                      "stats": {s: ctx.last_stat(s) for s in ("fallback_tiles", "heavy_loci", "sj_simple", "sj_bail", "sj_sites", "sj_emitted_sites")}}))


if __name__ == "__main__":
    main()
