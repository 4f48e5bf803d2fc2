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

    def __init__(self, gpu_index: int):
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
            if self.stop_flag.wait(0.002):  # a timed region of a few steps lasts tens of ms: sample every 2 ms
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
        return out["n_visited"], time.perf_counter() - t, len(out["calls"])

    base = min(contig_len // 4, 10_000_000)
    n, dt, _ = run(base, min(contig_len - 200, base + 50_000 * max(1, n_threads // 4)))  # calibration
    rate = n / dt
    span = int(min(contig_len - 200 - base, max(100_000, rate * target_s)))
    n, dt, calls = run(base, base + span)
    return {"loci": n, "seconds": dt, "value": n / dt, "calls": calls, "window": [base, base + span]}


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
        "config": {"workload": f"somatic-standard tumor/normal synthetic 60x/30x, {args.contig_mbp} Mbp contig, planted SNVs "
                               "(BASELINE configs[2])", "parallelism": f"cpu threads x{n_threads}"},
        "cpu_baseline": {"value": v, "unit": "loci/s", "cores": n_threads, "kind": "port",
                         "sample": f"oracle (C++ restatement; the Scala/Spark reference cannot run: no JVM) on loci "
                                   f"{sample['window']} of the seeded workload per step"},
        "e2e": {"value": v, "unit": "loci/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

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
    seed = 3000 + rank
    ctx.synth_generate(contig_len, DEPTHS, TUMOR, seed)
    ranges = [(0, 0, contig_len)]
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), wall, outs

    kernel_ms = []

    def step_resident():
        out = ctx.somatic_standard(ranges)
        kernel_ms.append(ctx.last_kernel_ms()[0])
        return out["n_visited"], len(out["calls"])

    sampler = ClockSampler(local_rank)
    launches0 = None

    def step_counted():
        return step_resident()

    # warm-up happens inside timed(); count launches over the timed region only
    for _ in range(args.warmup):
        step_resident()
    kernel_ms.clear()
    launches0 = ctx.kernel_launches
    sampler.start()
    ms_total, wall, outs = timed(step_counted, args.steps, 0)
    clocks = sampler.stop()
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
    tpath = os.path.join(ROOT, "profiles", "r1_traffic.json")
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
        hb = ctx.export_reads_compact_struct()
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
        e_loci = torch.tensor([float(sum(o[0] for o in e_outs))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(e_loci, op=dist.ReduceOp.SUM)
        e2e = {"value": float(e_loci.item()) / (e_ms / 1000.0), "unit": "loci/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h_box[0]),
               "note": f"reads in the compact wire form (guac_read_batch_compact, {hb.qual_bits}-bit qualities) from pinned host "
                       "memory each step; reference genome resident"}
        ctx.free_reads_compact_struct(hb)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        nthreads = os.cpu_count() or 1
        s = cpu_sample(contig_len, 3000, nthreads)
        cpu = {"value": s["value"], "unit": "loci/s", "cores": nthreads, "kind": "port",
               "sample": f"oracle (C++ restatement of the Scala reference, {nthreads} threads over loci ranges) on loci "
                         f"{s['window']} of the same seeded workload: {s['loci']} loci in {s['seconds']:.1f} s"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "loci/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"somatic-standard tumor/normal synthetic 60x/30x, {args.contig_mbp} Mbp contig per GPU, planted "
                                   "SNVs (BASELINE configs[2]); configs[3] is this path sharded by loci ranges",
                       "loci_per_step_per_gpu": loci_step, "reads_per_gpu": int(sum(DEPTHS) * contig_len // 100),
                       "calls_per_step": outs[0][1], "parallelism": f"loci-range shards x{world}, no collective",
                       "l2": "inputs (7.3 GB per GPU) are far larger than L2 (126 MB); no flush needed"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "wall_s_timed": wall,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
