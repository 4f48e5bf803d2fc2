"""Secondary measurement: the germline-assembly path (BASELINE configs[1] shape: one 30x sample) on one GPU,
with the oracle timed beside it on a bounded sample.  Not the headline bench (bench.py); prints one JSON line.

    python tools/bench_assembly.py [--mbp 2] [--steps 3] [--cpu-kbp 40]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from guacamole_b200 import synth  # noqa: E402
from guacamole_b200.native import Context, ga_params  # noqa: E402

SUITE = dict(kmer_size=47, assembly_window_range=120, min_occurrence=5, min_mean_kmer_quality=20, min_area_vaf_percent=10)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mbp", type=float, default=2.0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--cpu-kbp", type=int, default=40)
    ap.add_argument("--shortcut", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    import torch
    n = int(a.mbp * 1e6)
    seed = 2000
    ctx = Context(0)
    ctx.synth_generate(n, (30,), (False,), seed)
    p = ga_params(shortcut_assembly=bool(a.shortcut), **SUITE)
    ranges = [(0, 200, n - 200)]
    out = ctx.germline_assembly(ranges, p)  # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        out = ctx.germline_assembly(ranges, p)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / a.steps
    asm_ms, n_win = ctx.last_kernel_ms()
    line = {"metric": "loci/s through germline-assembly (one 30x sample)", "value": (n - 400) / dt, "unit": "loci/s",
            "ms_per_step": dt * 1e3, "windows_assembled": int(len(out["windows"])),
            "windows_consumed": int(out["windows"]["consumed"].sum()), "calls": int(len(out["calls"])),
            "assembly_stage_ms": asm_ms, "windows_per_s_assembly_stage": n_win / (asm_ms * 1e-3) if asm_ms else None,
            "config": {"workload": f"synthetic 30x, {a.mbp} Mbp, k=47 range=120 minOccurrence=5 shortcut={a.shortcut}"}}
    if not a.no_cpu:
        from oracle import oracle
        m = a.cpu_kbp * 1000
        ref = synth.make_reference(m, seed)
        batch = synth.generate_reads(m, (30,), (False,), seed)
        t0 = time.perf_counter()
        want = oracle.germline_assembly(batch, ref, [(0, 200, m - 200)], p)
        cdt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": (m - 400) / cdt, "unit": "loci/s", "cores": 1, "kind": "port",
                                "sample": f"first {a.cpu_kbp} kbp of the same seeded workload",
                                "windows": sum(1 for w in want if w[0] == "W")}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
