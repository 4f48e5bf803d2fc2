"""Secondary measurements: the other operators behind the C ABI on one GPU, each with the oracle timed beside it on a
bounded sample of the same seeded workload.  The headline bench is bench.py (somatic-standard); this prints one JSON
line per path.

    python tools/bench_paths.py [--mbp 5] [--steps 3] [--cpu-kbp 100] [--paths counts,germline,joint,assembly]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from guacamole_b200 import synth  # noqa: E402
from guacamole_b200.commands import somatic_joint as sj  # noqa: E402
from guacamole_b200.native import Context, ga_params, gs_params  # noqa: E402


def timed(fn, steps):
    import torch
    fn()  # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mbp", type=float, default=5.0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--cpu-kbp", type=int, default=100)
    ap.add_argument("--paths", default="counts,germline,joint,assembly")
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    n, m, seed = int(a.mbp * 1e6), a.cpu_kbp * 1000, 2000
    paths = a.paths.split(",")
    ctx = Context(0)
    if not a.no_cpu:
        from oracle import oracle

    def report(name, reference, dt, n_loci, extra, cpu):
        try:
            kernel_ms = ctx.last_kernel_ms()[0]
        except Exception:  # this entry point does not bracket its kernels with events
            kernel_ms = None
        line = {"path": name, "reference": reference, "value": n_loci / dt, "unit": "loci/s", "ms_per_step": dt * 1e3,
                "kernel_ms": kernel_ms, "loci_per_step": n_loci, **extra}
        if cpu is not None:
            line["cpu_baseline"] = {"value": cpu[0], "unit": "loci/s", "cores": 1, "kind": "port",
                                    "sample": f"oracle on the first {a.cpu_kbp} kbp of the same seeded workload ({cpu[1]:.1f} s)"}
            line["speedup_vs_one_core"] = line["value"] / cpu[0]
        print(json.dumps(line), flush=True)

    def cpu_time(fn, loci):
        if a.no_cpu:
            return None
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        return loci / dt, dt

    # one 30x sample (BASELINE configs[1] shape)
    if any(p in paths for p in ("counts", "germline", "assembly")):
        ctx.synth_generate(n, (30,), (False,), seed)
        if not a.no_cpu:
            ref1 = synth.make_reference(m, seed)
            batch1 = synth.generate_reads(m, (30,), (False,), seed)
    if "counts" in paths:
        dt, out = timed(lambda: ctx.pileup_counts([(0, 0, n)]), a.steps)
        report("pileup counts (pileupFlatMapOneSample, counting closure)", "distributed/PileupFlatMapUtils.scala:52", dt, n,
               {"loci_visited": int(len(out[0]))}, cpu_time(lambda: oracle.pileup_counts(batch1, ref1, [(0, 0, m)]), m))
    if "germline" in paths:
        p = gs_params()
        dt, out = timed(lambda: ctx.germline_standard([(0, 0, n)], p), a.steps)
        report("germline toy caller (callPileupVariant)", "commands/GermlineAssemblyCaller.scala:135-153,233-261", dt, n,
               {"calls": int(len(out["calls"]))}, cpu_time(lambda: oracle.germline_standard(batch1, ref1, [(0, 0, m)], p), m))
    if "assembly" in paths:
        p = ga_params(kmer_size=47, assembly_window_range=120, min_occurrence=5, min_mean_kmer_quality=20, min_area_vaf_percent=10)
        dt, out = timed(lambda: ctx.germline_assembly([(0, 200, n - 200)], p), a.steps)
        report("germline-assembly (discoverGermlineVariants)", "commands/GermlineAssemblyCaller.scala:84-219", dt, n - 400,
               {"windows_assembled": int(len(out["windows"])), "windows_consumed": int(out["windows"]["consumed"].sum()),
                "calls": int(len(out["calls"]))},
               cpu_time(lambda: oracle.germline_assembly(batch1, ref1, [(0, 200, m - 200)], p), m - 400))
    if "joint" in paths:  # normal 30x + two tumors 60x / 30x
        depths, tumor, kinds = (30, 60, 30), (False, True, True), [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
        ctx.synth_generate(n, depths, tumor, seed)
        p = sj.sj_params(kinds)
        dt, out = timed(lambda: ctx.somatic_joint_records([(0, 0, n)], p), a.steps)  # records as numpy arrays, not Python dicts
        cpu = None
        if not a.no_cpu:
            ref3 = synth.make_reference(m, seed)
            batch3 = synth.generate_reads(m, depths, tumor, seed)
            cpu = cpu_time(lambda: oracle.somatic_joint(batch3, ref3, [(0, 0, m)], p), m)
        report("somatic-joint (makeCalls / callPileups), 3 samples 30x/60x/30x", "commands/SomaticJointCaller.scala:154-203", dt, n,
               {"alleles_emitted": int(len(out[0]))}, cpu)


if __name__ == "__main__":
    main()
