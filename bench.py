#!/usr/bin/env python
"""bench.py — loci/sec through pileup + likelihood (BASELINE.json metric).

Workload (N=1 and per rank for N>1, weak scaling): BASELINE.json configs[2] —
somatic-standard tumor/normal on synthetic 60x/30x reads with planted SNVs over a 50 Mbp
chromosome-scale contig (the largest configuration of `configs` that fits one GPU; configs[3],
the 3 Gbp 4-sample genome, is this same path sharded by LociPartitioning ranges, one range per
rank).  Reads are generated directly in HBM by the seeded generator (guac_synth_generate;
guacamole_b200/synth.py is its specification) with seed 3000 + rank.

A "step" is one pass of the hot path (guac_somatic_standard: pileup + likelihood + call records)
over the rank's resident batch.  `value` times K steps with inputs resident in HBM; `e2e` times
the same call fed from pinned HOST buffers (guac_reads_upload inside the timed region, call
records read back).  `--impl reference` times the CPU restatement of the reference (oracle/)
on the host cores, on a bounded sample of the same seeded workload.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEPTHS = (30, 60)  # sample 0 = normal 30x, sample 1 = tumor 60x
TUMOR = (False, True)
READ_BYTES = 161.5  # 24 B header + 25 B bases + 12.5 B mask + 100 B quals
METRIC = "loci/sec (pileup+likelihood)"


def algorithmic_bytes_per_locus(total_depth: int, n_samples: int) -> float:
    # 1.615*D + 1 (reference) + 8*S (dense per-sample summary)
    return READ_BYTES / 100.0 * total_depth + 1 + 8 * n_samples


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period: float = 0.05):
        # NVML queries serialise in the driver: with one sampler per rank a short period steals host time from the
        # synchronous C calls of every rank (measured at N = 8).  50 ms by default; the timed region is stretched to
        # hold a few samples by repeating the steps, never by sampling faster.
        self.period = period
        self.gpu, self.rows, self.proc = gpu_index, [], None
        self.nvml, self.stop_flag, self.thread = None, threading.Event(), None

    def _nvml_loop(self):
        n = self.nvml
        names = (("hw_slowdown", n.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", n.nvmlClocksEventReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", n.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", n.nvmlClocksEventReasonSwPowerCap))
        while True:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                self.rows.append([str(self.gpu), str(sm), str(self.max_sm), "0"] + ["Active" if mask & bit else "Not Active" for _, bit in names])
            except Exception:
                pass
            if self.stop_flag.wait(self.period):
                return

    def start(self):
        # in-process NVML first (nvidia-smi needs ~100 ms to produce its first line, longer than a short timed region)
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            index = int(visible.split(",")[self.gpu]) if visible and all(v.strip().isdigit() for v in visible.split(",")) else self.gpu
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._nvml_loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self.stop_flag.set()
            self.thread.join(timeout=5)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_sample(contig_len: int, seed: int, n_threads: int, target_s: float = 12.0):
    """Times the oracle (CPU restatement of the reference) on a bounded window of the workload."""
    import numpy as np
    from guacamole_b200 import synth
    from guacamole_b200.io import ContigWindow, Reference
    from oracle import oracle

    def run(lo, hi):
        j_lo, j_hi = synth.read_window_for_loci(contig_len, DEPTHS, lo, hi)
        j_lo -= j_lo % 8
        batch = oracle.synth_generate(contig_len, DEPTHS, TUMOR, seed, j_lo, j_hi)
        pad = 256
        rlo, rhi = max(0, lo - pad), min(contig_len, hi + pad)
        ref = Reference({"synth1": ContigWindow("synth1", rlo, oracle.synth_reference(seed, rlo, rhi), contig_len)})
        t = time.perf_counter()
        out = oracle.somatic_standard(batch, ref, [(0, lo, hi)], n_threads=n_threads)
        return out["n_visited"], time.perf_counter() - t, out

    base = min(contig_len // 4, 10_000_000)
    n, dt, _ = run(base, min(contig_len - 200, base + 50_000 * max(1, n_threads // 4)))  # calibration
    rate = n / dt
    span = int(min(contig_len - 200 - base, max(100_000, rate * target_s)))
    n, dt, out = run(base, base + span)
    return {"loci": n, "seconds": dt, "value": n / dt, "calls": len(out["calls"]), "window": [base, base + span], "out": out}


def run_reference(args, rank, world):
    if rank != 0:
        return
    n_threads = os.cpu_count() or 1
    contig_len = args.contig_mbp * 1_000_000
    vals = []
    sample = None
    for i in range(args.warmup + args.steps):
        s = cpu_sample(contig_len, 3000, n_threads, target_s=args.reference_step_seconds)
        if i >= args.warmup:
            vals.append(s)
        sample = s
    loci = sum(s["loci"] for s in vals)
    secs = sum(s["seconds"] for s in vals)
    v = loci / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "loci/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * secs / max(1, len(vals)), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"window of BASELINE configs[2] per step (loci {sample['window']} of the somatic-standard tumor/normal "
                               f"synthetic 60x/30x {args.contig_mbp} Mbp contig, planted SNVs): a bounded sample, rates are per locus",
                   "parallelism": f"cpu threads x{n_threads}"},
        "cpu_baseline": {"value": v, "unit": "loci/s", "cores": n_threads, "kind": "port",
                         "sample": f"oracle (C++ restatement; the Scala/Spark reference cannot run: no JVM) on loci "
                                   f"{sample['window']} of the seeded workload per step"},
        "e2e": {"value": v, "unit": "loci/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)

def oracle_threads_rate(prepare, n_threads, base, contig_len, target_s, calib=20_000):
    """The oracle over n_threads disjoint windows at once (the ctypes calls release the GIL; this mirrors Spark local[N]):
    prepare(lo, hi) builds one window's inputs (untimed) and returns a callable that runs the oracle and returns the loci
    it visited."""
    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    n = prepare(base, base + calib)()
    rate = n / max(time.perf_counter() - t0, 1e-9)
    w = int(max(calib, min((contig_len - 400 - base) // n_threads, rate * target_s)))
    jobs = [prepare(base + k * w, base + (k + 1) * w) for k in range(n_threads)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(n_threads) as ex:
        loci = sum(ex.map(lambda j: j(), jobs))
    wall = time.perf_counter() - t0
    return {"value": loci / wall, "loci": int(loci), "seconds": wall, "window": [base, base + n_threads * w]}


def e2e_with_packer(ctx, timed, args, mbp: int = 5):
    """The end-to-end leg from where a shim really starts: decoded SAM records (ASCII bases, one quality byte per base, SAM
    flags) -> guac_pack_reads (Read.apply, the ingest filters, the compact wire form; all host cores) -> guac_reads_upload_compact
    -> guac_somatic_standard, on a shard of the workload small enough to build its records on the host in seconds."""
    import numpy as np
    from guacamole_b200 import host_stage as hs
    from oracle import oracle
    n = mbp * 1_000_000
    batch = oracle.synth_generate(n, DEPTHS, TUMOR, 3000)
    records, keepalive = hs.sam_records_of(batch)
    from guacamole_b200.io import ContigWindow, Reference
    host_ref = Reference({"synth1": ContigWindow("synth1", 0, oracle.synth_reference(3000, 0, n), n)})
    ctx.upload_reference(host_ref, ["synth1"])
    pack_s = []

    def step():
        t = time.perf_counter()
        packed, kept = hs.pack_reads(records, ctx=ctx, non_duplicate=True, passed_vendor_quality_checks=True, is_paired=True,
                                     ref=None if args.e2e_explicit_bases else host_ref, contig_names=["synth1"])
        pack_s.append(time.perf_counter() - t)
        ctx.upload_reads_compact(packed)
        out = ctx.somatic_standard([(0, 0, n)])
        hs.free_packed(packed, ctx=ctx)
        return out["n_visited"], len(out["calls"])
    ms, _, outs = timed(step, 2, 1)
    raw_bytes = sum(int(a.nbytes) for a in keepalive)
    return {"form": f"from SAM-record columns through the host packer (guac_pack_reads{'' if args.e2e_explicit_bases else '_ref'}, {os.cpu_count()} threads), "
                    f"{mbp} Mbp shard of the workload",
            "value": sum(o[0] for o in outs) / (ms / 1000.0), "unit": "loci/s", "host_pack_ms": 1000.0 * sum(pack_s[-2:]) / 2,
            "sam_record_bytes": raw_bytes, "reads": int(batch.n_reads)}


def run_paths(ctx, args, rank, local_rank, world, timed, peak, peak_src):
    """The other configurations of BASELINE.json through their own callers, each with the headline's blocks: configs[3]'s
    per-GPU shape (somatic-joint, 2 normal + 2 tumor at 30/30/60/60), configs[1] (germline toy caller, one 30x sample) and
    configs[4] (germline-assembly, k = 31).  Every rank runs them on its own seeded contig (weak scaling)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from guacamole_b200 import synth
    from guacamole_b200.commands import somatic_joint as sj
    from guacamole_b200.io import ContigWindow, Reference
    from guacamole_b200.native import ga_params, gs_params
    steps = max(1, min(args.steps, 3))

    def total(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def measure(name, reference, depths, tumor, contig_len, seed, call, visited_of, abytes, extra, oracle_job, roof=True, profile=0):
        ctx.synth_generate(contig_len, depths, tumor, seed, profile=profile)
        kms = []

        def step():
            out = call()
            try:
                kms.append(ctx.last_kernel_ms()[0])
            except Exception:
                pass
            return out
        ms, _, outs = timed(step, steps, 1)
        loci = visited_of(outs[0])
        value = total(loci * steps) / (ms / 1000.0)
        entry = {"reference": reference, "metric": METRIC, "value": value, "unit": "loci/s", "ms_per_step": ms / steps,
                 "config": {"workload": name, "loci_per_step_per_gpu": int(loci), "reads_per_gpu": int(sum(depths) * contig_len // 100)}}
        entry["config"].update(extra(outs[0]))
        k_ms = sum(kms[-steps:]) / max(1, len(kms[-steps:])) if kms else None
        if roof and k_ms:
            achieved = abytes * loci / (k_ms / 1000.0) / 1e9
            entry["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                 "traffic": None, "kernel_ms": k_ms, "algorithmic_bytes_per_locus": abytes,
                                 "algorithmic_bytes_per_launch": abytes * loci, "peak_source": peak_src,
                                 "kernel": "every kernel of one call, first launch to last (CUDA events on the library's stream)"}
        elif k_ms:
            entry["kernel_ms"] = k_ms
        if not args.no_e2e:
            ctx.set_option("compact_ref_relative", 0 if args.e2e_explicit_bases else 1)
            hb = ctx.export_reads_compact_struct()
            ctx.set_option("compact_ref_relative", 0)
            n = int(hb.n_reads)

            def column(p, ctype, dtype):
                return np.frombuffer((ctype * n).from_address(C.cast(p, C.c_void_p).value), dtype=dtype)
            n_bases = n * int(hb.uniform_read_len) if hb.uniform_read_len > 0 else int(column(hb.read_len, C.c_uint16, np.uint16).sum(dtype=np.int64))
            n_cigar = int(column(hb.n_cigar_ops8, C.c_uint8, np.uint8).sum(dtype=np.int64)) if hb.n_cigar_ops8 else \
                int(column(hb.n_cigar_ops, C.c_uint16, np.uint16).sum(dtype=np.int64))
            h2d = hb.wire_bytes(n_bases, n_cigar)
            d2h = [0]

            def step_e2e():
                ctx.upload_reads_compact(hb)
                out = call()
                d2h[0] = sum(int(getattr(x, "nbytes", len(x) if isinstance(x, (bytes, bytearray)) else 0)) for x in
                             (out.values() if isinstance(out, dict) else out))
                return out
            e_ms, _, e_outs = timed(step_e2e, steps, 1)
            entry["e2e"] = {"value": total(visited_of(e_outs[0]) * steps) / (e_ms / 1000.0), "unit": "loci/s",
                            "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h[0]),
                            "note": f"reads in the compact wire form ({hb.qual_bits}-bit qualities, "
                                    + ("bases as differences from the resident reference where a read allows it" if hb.mm_count else "every read's bases explicit")
                                    + f": {h2d / max(n, 1):.1f} B per read) from pinned host memory each step"}
            ctx.free_reads_compact_struct(hb)
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            from oracle import oracle
            nthreads = os.cpu_count() or 1

            def prepare(lo, hi):
                j_lo, j_hi = synth.read_window_for_loci(contig_len, depths, lo, hi)
                j_lo -= j_lo % 8
                batch = oracle.synth_generate(contig_len, depths, tumor, seed, j_lo, j_hi, profile=profile)
                rlo, rhi = max(0, lo - 512), min(contig_len, hi + 512)
                ref = Reference({"synth1": ContigWindow("synth1", rlo, oracle.synth_reference(seed, rlo, rhi), contig_len)})
                return lambda: oracle_job(oracle, batch, ref, lo, hi)
            c = oracle_threads_rate(prepare, nthreads, min(contig_len // 4, 1_000_000), contig_len, target_s=6.0)
            entry["cpu_baseline"] = {"value": c["value"], "unit": "loci/s", "cores": nthreads, "kind": "port",
                                     "sample": f"oracle, {nthreads} threads each over its own window, loci {c['window']} of the same seeded "
                                               f"workload: {c['loci']} loci in {c['seconds']:.1f} s"}
        return entry

    paths = {}
    # configs[2] again, on the "realistic" variant of the read model (guacamole_b200/synth.py profile 1): does the fast path
    # survive reads with quality-2 tails, depth that varies, 10 % multi-operator reads and 1 % reads with 10..20 mismatches?
    n2 = args.contig_mbp * 1_000_000
    paths["somatic_standard_realistic_reads"] = measure(
        f"somatic-standard tumor/normal synthetic 60x/30x, {args.contig_mbp} Mbp contig per GPU, realistic read model: read density varying "
        "like a Poisson pileup's, 15 % of reads with a quality-2 tail of 5..30 bases, 10 % multi-operator reads, 1 % of reads with 10..20 "
        "mismatches (BASELINE configs[2] shape)", "commands/SomaticStandardCaller.scala:100-117,155-253", DEPTHS, TUMOR, n2, 3000 + rank,
        lambda: ctx.somatic_standard([(0, 0, n2)]), lambda o: o["n_visited"], algorithmic_bytes_per_locus(sum(DEPTHS), len(DEPTHS)),
        lambda o: {"calls": int(len(o["calls"])), "fallback_tiles": ctx.last_stat("fallback_tiles"), "heavy_loci": ctx.last_stat("heavy_loci")},
        lambda oracle, b, r, lo, hi: oracle.somatic_standard(b, r, [(0, lo, hi)])["n_visited"], profile=1)
    # configs[3]: somatic-joint, 2 normal + 2 tumor (one rank's shard of the sharded genome)
    depths, tumor = (30, 30, 60, 60), (False, False, True, True)
    kinds = [sj.NORMAL_DNA, sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
    n = args.joint_mbp * 1_000_000
    pj = sj.sj_params(kinds)
    paths["somatic_joint_2n2t"] = measure(
        f"somatic-joint 2 normal + 2 tumor synthetic 30x/30x/60x/60x, {args.joint_mbp} Mbp contig per GPU, planted SNVs (one GPU's shard of "
        "BASELINE configs[3])", "commands/SomaticJointCaller.scala:154-203", depths, tumor, n, 4000 + rank,
        lambda: ctx.somatic_joint_records([(0, 0, n)], pj, copy=False), lambda o: o[3], algorithmic_bytes_per_locus(sum(depths), len(depths)),
        lambda o: {"alleles_emitted": int(len(o[0])), "records": "read in place in the library's pinned output buffers (no host copy)",
                   **{k: ctx.last_stat(k) for k in ("fallback_tiles", "heavy_loci", "sj_simple", "sj_bail",
                                                                                       "sj_sites", "sj_emitted_sites")}},
        lambda oracle, b, r, lo, hi: oracle.somatic_joint(b, r, [(0, lo, hi)], pj)[1])
    # configs[1]: the germline toy caller on one 30x sample
    n1 = args.germline_mbp * 1_000_000
    pg = gs_params()
    paths["germline_standard_30x"] = measure(
        f"germline toy caller (callPileupVariant) on one synthetic 30x sample, {args.germline_mbp} Mbp contig (BASELINE configs[1])",
        "commands/GermlineAssemblyCaller.scala:135-153,233-261", (30,), (False,), n1, 1000 + rank,
        lambda: ctx.germline_standard([(0, 0, n1)], pg), lambda o: o["n_visited"], algorithmic_bytes_per_locus(30, 1),
        lambda o: {"calls": int(len(o["calls"])), "fallback_tiles": ctx.last_stat("fallback_tiles"), "heavy_loci": ctx.last_stat("heavy_loci")},
        lambda oracle, b, r, lo, hi: oracle.germline_standard(b, r, [(0, lo, hi)], pg)["n_visited"])
    # configs[4]: germline-assembly, k = 31 windows (hash-table / latency bound: windows and k-mers per second, no HBM roofline)
    n4 = args.assembly_mbp * 1_000_000
    pa = ga_params(kmer_size=31, assembly_window_range=41, min_occurrence=3, min_mean_kmer_quality=0)

    def asm_extra(o):
        wins = int(len(o["windows"]))
        kmers = wins * int(30 * (82 + 100) / 100) * (100 - 31 + 1)  # reads per window x k-mers per read
        d = {"windows_assembled": wins, "windows_consumed": int(o["windows"]["consumed"].sum()), "calls": int(len(o["calls"])),
             "kmer_insertions_per_step_estimate": kmers,
             "note": "the generator's indels are per read (1 % insertions, 1 % deletions of 1..10 bases); it plants SNVs only"}
        try:
            ms, nw = ctx.last_kernel_ms()
            d.update({"assembly_stage_ms": ms, "windows_per_s": nw / (ms * 1e-3), "kmers_per_s_estimate": kmers / (ms * 1e-3),
                      "gb_per_s_estimate": wins * (54.6 * READ_BYTES + 82 + 3822 * 16) / (ms * 1e-3) / 1e9})
        except Exception:
            pass
        return d
    paths["germline_assembly_k31"] = measure(
        f"germline-assembly (discoverGermlineVariants) k = 31, window range 41, one synthetic 30x sample, {args.assembly_mbp} Mbp (BASELINE "
        "configs[4])", "commands/GermlineAssemblyCaller.scala:84-219", (30,), (False,), n4, 5000 + rank,
        lambda: ctx.germline_assembly([(0, 200, n4 - 200)], pa), lambda o: n4 - 400, 0.0, asm_extra,
        lambda oracle, b, r, lo, hi: (oracle.germline_assembly(b, r, [(0, lo, hi)], pa), hi - lo)[1], roof=False)
    return paths


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--contig-mbp", type=int, default=50)
    ap.add_argument("--reference-step-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-explicit-bases", action="store_true", help="e2e: every read's bases on the wire (no reference-relative column)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default, the headline): every rank has its own contig; strong: ONE seeded contig cut by "
                         "UniformPartitioner ranges, every rank calls its range, the calls are gathered and checked against the uncut run")
    ap.add_argument("--no-pin", action="store_true", help="do not pin the ranks to disjoint core sets")
    ap.add_argument("--clock-period-ms", type=float, default=None,
                    help="clock sampler period (default 5 ms on one GPU, 50 ms with several ranks: NVML queries of all ranks serialise "
                         "in the driver); 0 disables the sampler (A/B)")
    ap.add_argument("--no-paths", action="store_true", help="skip the `paths` block (the other callers / configurations)")
    ap.add_argument("--joint-mbp", type=int, default=50)
    ap.add_argument("--germline-mbp", type=int, default=10)
    ap.add_argument("--assembly-mbp", type=int, default=2)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.clock_period_ms is None:
        args.clock_period_ms = 5.0 if world == 1 else 50.0

    # ranks on disjoint core sets: eight synchronous host loops plus their driver threads otherwise migrate over each other
    pinned_cores = None
    if world > 1 and not args.no_pin and hasattr(os, "sched_setaffinity"):
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = max(1, len(cores) // world)
            mine = cores[local_rank * per:(local_rank + 1) * per] or cores
            os.sched_setaffinity(0, mine)
            pinned_cores = len(mine)
        except OSError:
            pinned_cores = None

    import __graft_entry__ as graft
    if rank == 0:
        graft.build()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    if rank != 0:
        graft.build()  # no-op once rank 0 has built
    from guacamole_b200.native import Context, ReadBatchC, SS_CALL_DTYPE

    contig_len = args.contig_mbp * 1_000_000
    ctx = Context(local_rank)
    strong = args.scaling == "strong"
    seed = 3000 if strong else 3000 + rank
    ctx.synth_generate(contig_len, DEPTHS, TUMOR, seed)
    ranges = [(0, 0, contig_len)]
    if strong:
        # LociPartitioning ranges of UniformPartitioner (loci/partitioning/UniformPartitioner.scala:48-86), one per rank.  Every
        # rank holds the whole contig's reads in HBM — a superset of its partition's reads plus halo
        # (readsets/rdd/PartitionedRegions.scala:186-208) — and the planner only ever touches the reads over its own range.
        from guacamole_b200 import partitioning
        ranges = partitioning.ranges_of_partition(partitioning.uniform_partition([(0, 0, contig_len)], world), rank)
        args.no_paths = args.no_e2e = True  # (the end-to-end and per-path legs belong to the weak-scaling headline)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    local_ms = [0.0]

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        outs = [step_fn() for _ in range(steps)]
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        ms = max(e0.elapsed_time(e1), 0.0)
        # the C call is synchronous, so host time between the events is part of the step; use the larger
        ms = max(ms, 0.0)
        local_ms[0] = ms  # this rank's own time (the returned one is the maximum over the ranks)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), wall, outs

    kernel_ms = []

    counters = [0, 0]

    def step_resident():
        out = ctx.somatic_standard(ranges)
        kernel_ms.append(ctx.last_kernel_ms()[0])
        counters[0], counters[1] = ctx.last_stat("fallback_tiles"), ctx.last_stat("heavy_loci")
        return out["n_visited"], len(out["calls"])

    sampler = ClockSampler(local_rank, period=max(args.clock_period_ms, 1.0) / 1000.0)
    launches0 = None

    def step_counted():
        return step_resident()

    # warm-up happens inside timed(); count launches over the timed region only
    for _ in range(args.warmup):
        step_resident()
    kernel_ms.clear()
    launches0 = ctx.kernel_launches
    if args.clock_period_ms > 0:
        sampler.start()
    ms_total, wall, outs = timed(step_counted, args.steps, 0)
    clocks = sampler.stop() if args.clock_period_ms > 0 else {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["sampler off"], "samples": 0}
    # every rank's own step: its time over the timed region, the kernels' share (CUDA events inside the library), the rest
    # (launches, the one wait of the call, the records' copy, Python) — what an N-process box adds shows in the last
    mine = torch.tensor([local_ms[0] / args.steps, sum(kernel_ms) / max(1, len(kernel_ms))], dtype=torch.float64, device="cuda")
    per_rank = [mine.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, mine)
    per_rank = [{"rank": r, "ms_per_step": float(t[0]), "kernel_ms": float(t[1]), "host_ms": float(t[0] - t[1])} for r, t in enumerate(per_rank)]
    if args.clock_period_ms > 0 and clocks.get("samples", 0) < 3:
        # the timed region was shorter than a few sampler periods: sample the clocks over a second, untimed run of the
        # same steps under load (the timed numbers above are not touched)
        sampler = ClockSampler(local_rank, period=0.01)
        sampler.start()
        t_end = time.perf_counter() + 0.25
        while time.perf_counter() < t_end:
            step_resident()
        clocks = sampler.stop()
        clocks["note"] = "sampled over 0.25 s of the same steps right after the timed region (the region itself is shorter than the sampler period)"
    launches = ctx.kernel_launches - launches0
    loci_step = outs[0][0]
    total_loci = torch.tensor([float(sum(o[0] for o in outs))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(total_loci, op=dist.ReduceOp.SUM)
    value = float(total_loci.item()) / (ms_total / 1000.0)

    # roofline of the dominant kernel (CUDA events on the library's stream, averaged over the timed steps)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    k_ms = sum(kernel_ms) / len(kernel_ms)
    abytes = algorithmic_bytes_per_locus(sum(DEPTHS), len(DEPTHS))
    achieved = abytes * loci_step / (k_ms / 1000.0) / 1e9
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.exists(tpath) and args.contig_mbp == 50:
        tj = json.load(open(tpath))
        traffic, traffic_src = tj["traffic_bytes_per_launch"], tj["source"]
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_src, "kernel": "ss_read_events_kernel + ss_tile_kernel<TileCfg<512>, 0, STREAM> (+ heavy-locus gather kernel): the per-locus kernels of one step",
                "kernel_ms": k_ms, "algorithmic_bytes_per_launch": abytes * loci_step, "algorithmic_bytes_per_locus": abytes,
                "peak_source": peak_src}

    # end to end: pinned host buffers in the compact wire form -> guac_reads_upload_compact (PCIe copy + on-device
    # expansion) -> guac_somatic_standard -> call records on host
    e2e = None
    if not args.no_e2e:
        # the wire form with reference-relative bases (the reference genome is resident, as it is for a Spark executor that
        # calls partition after partition): a read of one aligned operator travels as its differences from the reference
        ctx.set_option("compact_ref_relative", 0 if args.e2e_explicit_bases else 1)
        hb = ctx.export_reads_compact_struct()
        ctx.set_option("compact_ref_relative", 0)
        n = int(hb.n_reads)
        def column(p, ctype, dtype):
            return np.frombuffer((ctype * n).from_address(C.cast(p, C.c_void_p).value), dtype=dtype)

        # bytes on the wire, counted from the columns the struct holds (narrow start / operator-count / read-length columns
        # where the exporter could use them)
        n_bases = n * int(hb.uniform_read_len) if hb.uniform_read_len > 0 else int(column(hb.read_len, C.c_uint16, np.uint16).sum(dtype=np.int64))
        n_cigar = int(column(hb.n_cigar_ops8, C.c_uint8, np.uint8).sum(dtype=np.int64)) if hb.n_cigar_ops8 else \
            int(column(hb.n_cigar_ops, C.c_uint16, np.uint16).sum(dtype=np.int64))
        h2d = hb.wire_bytes(n_bases, n_cigar)
        d2h_box = [0]

        def step_e2e():
            ctx.upload_reads_compact(hb)
            out = ctx.somatic_standard(ranges)
            d2h_box[0] = out["calls"].nbytes + len(out["arena"]) + 8
            return out["n_visited"], len(out["calls"])

        e_ms, _, e_outs = timed(step_e2e, max(1, min(args.steps, 3)), 1)
        if any(tuple(o) != tuple(outs[0][:2]) for o in e_outs):
            sys.exit(f"e2e: the batch that went through the wire form gives {e_outs[0]}, the resident one {outs[0][:2]} (loci visited, calls)")
        e_loci = torch.tensor([float(sum(o[0] for o in e_outs))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(e_loci, op=dist.ReduceOp.SUM)
        e2e = {"value": float(e_loci.item()) / (e_ms / 1000.0), "unit": "loci/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h_box[0]),
               "note": f"reads in the compact wire form (guac_read_batch_compact, {hb.qual_bits}-bit qualities, "
                       + ("bases as differences from the resident reference where a read allows it" if hb.mm_count else "every read's bases explicit")
                       + f": {h2d / max(n, 1):.1f} B per read) from pinned host memory each step; reference genome resident"}
        had_ref_relative = bool(hb.mm_count)
        ctx.free_reads_compact_struct(hb)
        # the same leg with the wider quality forms of the wire format (real, unbinned qualities travel as 4-bit codes when
        # a batch holds at most 16 distinct values, one byte per base otherwise), and with the host packer inside
        variants = []
        for bits in ((2, 4, 8) if had_ref_relative else (4, 8)):  # (all with every read's bases explicit)
            ctx.set_option("compact_min_qual_bits", bits)
            hv = ctx.export_reads_compact_struct()
            ctx.set_option("compact_min_qual_bits", 2)

            def step_v():
                ctx.upload_reads_compact(hv)
                out = ctx.somatic_standard(ranges)
                return out["n_visited"], len(out["calls"])
            v_ms, _, v_outs = timed(step_v, 2, 1)
            v_loci = torch.tensor([float(sum(o[0] for o in v_outs))], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(v_loci, op=dist.ReduceOp.SUM)
            variants.append({"form": f"{hv.qual_bits}-bit qualities, every read's bases explicit", "value": float(v_loci.item()) / (v_ms / 1000.0), "unit": "loci/s",
                             "h2d_bytes_per_step": int(hv.wire_bytes(n_bases, n_cigar))})
            ctx.free_reads_compact_struct(hv)
        if rank == 0 and world == 1:
            variants.append(e2e_with_packer(ctx, timed, args))
            ctx.synth_generate(contig_len, DEPTHS, TUMOR, seed)  # (the packer leg uploaded its own shard: restore the workload)
        e2e["variants"] = variants

    strong_check = None
    if strong:
        mine = ctx.somatic_standard(ranges)
        gathered = [None] * world
        if world > 1:
            dist.all_gather_object(gathered, {"start": mine["calls"]["start"].copy(), "odds": mine["calls"]["somatic_log_odds"].copy(),
                                              "n_visited": mine["n_visited"]})
        else:
            gathered = [{"start": mine["calls"]["start"], "odds": mine["calls"]["somatic_log_odds"], "n_visited": mine["n_visited"]}]
        if rank == 0:  # host-side gather of the sparse records, against the uncut run of the same contig
            whole = ctx.somatic_standard([(0, 0, contig_len)])
            starts = np.concatenate([g["start"] for g in gathered])
            odds = np.concatenate([g["odds"] for g in gathered])
            order = np.argsort(starts, kind="stable")
            ok = (len(starts) == len(whole["calls"]) and np.array_equal(starts[order], whole["calls"]["start"]) and
                  bool(np.allclose(odds[order], whole["calls"]["somatic_log_odds"], rtol=1e-9, atol=0)) and
                  sum(g["n_visited"] for g in gathered) == whole["n_visited"])
            strong_check = {"calls": int(len(starts)), "uncut_calls": int(len(whole["calls"])), "ok": bool(ok)}
    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not strong:
        nthreads = os.cpu_count() or 1
        s = cpu_sample(contig_len, 3000, nthreads)
        cpu = {"value": s["value"], "unit": "loci/s", "cores": nthreads, "kind": "port",
               "sample": f"oracle (C++ restatement of the Scala reference, {nthreads} threads over loci ranges) on loci "
                         f"{s['window']} of the same seeded workload: {s['loci']} loci in {s['seconds']:.1f} s"}
        # parity where the number is made: the GPU's calls inside the oracle's window of this very batch
        lo, hi = s["window"]
        got = ctx.somatic_standard(ranges)
        g, w = got["calls"], s["out"]["calls"]
        g = g[(g["start"] >= lo) & (g["start"] < hi)]

        def alleles(calls, arena):
            return [(int(c["start"]), arena[c["ref_off"]:c["ref_off"] + c["ref_len"]], arena[c["alt_off"]:c["alt_off"] + c["alt_len"]])
                    for c in calls]
        ok = len(g) == len(w) and alleles(g, got["arena"]) == alleles(w, s["out"]["arena"]) and bool(
            np.allclose(g["somatic_log_odds"], w["somatic_log_odds"], rtol=1e-9, atol=0))
        parity = {"loci": int(s["loci"]), "calls": int(len(w)), "gpu_calls": int(len(g)), "ok": bool(ok),
                  "compared": "start, ref / alt alleles exactly, somatic log odds at 1e-9 relative, oracle vs GPU inside the cpu_baseline window"}

    paths = None
    if not args.no_paths:
        paths = run_paths(ctx, args, rank, local_rank, world, timed, peak, peak_src)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "loci/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": (f"somatic-standard tumor/normal synthetic 60x/30x, ONE {args.contig_mbp} Mbp contig cut into {world} "
                                    "UniformPartitioner ranges, one per GPU, calls gathered on the host (BASELINE configs[2], strong scaling)"
                                    if strong else
                                    f"somatic-standard tumor/normal synthetic 60x/30x, {args.contig_mbp} Mbp contig per GPU, planted "
                                    "SNVs (BASELINE configs[2]); configs[3] is this path sharded by loci ranges"),
                       "loci_per_step_per_gpu": loci_step, "reads_per_gpu": int(sum(DEPTHS) * contig_len // 100),
                       "calls_per_step": outs[0][1], "fallback_tiles": counters[0], "heavy_loci": counters[1],
                       "parallelism": f"loci-range shards x{world}, no collective", "cores_per_rank": pinned_cores,
                       "l2": "inputs (7.3 GB per GPU) are far larger than L2 (126 MB); no flush needed"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "wall_s_timed": wall, "parity": parity, "per_rank": per_rank, "paths": paths,
        }
        if strong:
            line["strong_scaling_check"] = strong_check
        print(json.dumps(line), flush=True)
        if strong_check is not None and not strong_check["ok"]:
            sys.exit("strong-scaling check failed: the gathered calls differ from the uncut run")
        if parity is not None and not parity["ok"]:
            sys.exit("parity check failed: GPU calls differ from the oracle's inside the cpu_baseline window")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
