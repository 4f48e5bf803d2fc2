"""Loci partitioning and the multi-GPU sharding contract.

Mirrors `UniformPartitionerBase.partitionsMap`
(reference: src/main/scala/org/hammerlab/guacamole/loci/partitioning/UniformPartitioner.scala:48-86) and
the read-to-partition assignment of `PartitionedRegions.compute`
(readsets/rdd/PartitionedRegions.scala:186-208: every read goes to each partition whose loci it
overlaps, +- halfWindowSize — the "halo").

Multi-GPU: the genome shards by LociPartitioning ranges; every GPU runs the same hot path over its own
ranges with its own (halo-including) reads; there is no data-path collective.  Results (sparse call
records) are gathered on the host and sorted by (contig, start), as the reference does at output
(jointcaller/VCFOutput.scala:97-100, variants/GenotypeOutputArgs.scala:96-98).
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np

from .reads import ReadBatch

Range = Tuple[int, int, int]  # (contig_id, start, end)


def java_round(x: float) -> int:
    return int(math.floor(x + 0.5))


def uniform_partition(loci: Sequence[Range], num_partitions: int) -> List[Tuple[int, int, int, int]]:
    """[(contig_id, start, end, partition)] — UniformPartitioner.scala:48-86.  `loci` must be sorted,
    disjoint ranges in contig order."""
    assert num_partitions >= 1, "`numPartitions` (--parallelism) should be >= 1"
    count = sum(e - s for _, s, e in loci)
    loci_per_partition = max(1.0, count / float(num_partitions))
    assigned, partition = 0, 0
    out = []
    for contig, s, e in loci:
        start = s
        while start < e:
            remaining = java_round((partition + 1) * loci_per_partition - assigned)
            length = min(remaining, e - start)
            if length > 0:
                # LociMap.Builder merges adjacent ranges with the same value
                if out and out[-1][0] == contig and out[-1][2] == start and out[-1][3] == partition:
                    out[-1] = (contig, out[-1][1], start + length, partition)
                else:
                    out.append((contig, start, start + length, partition))
            start += length
            assigned += length
            if java_round((partition + 1) * loci_per_partition - assigned) == 0:
                partition += 1
    assert assigned == count
    return out


def partitioning_to_string(parts, contig_names: Sequence[str]) -> str:
    """LociPartitioning text form `contig:start-end=partition` (loci/partitioning/LociPartitioning.scala:112-138)."""
    return ",".join(f"{contig_names[c]}:{s}-{e}={p}" for c, s, e, p in parts)


def ranges_of_partition(parts, partition: int) -> List[Range]:
    return [(c, s, e) for c, s, e, p in parts if p == partition]


def reads_for_ranges(batch: ReadBatch, ranges: Sequence[Range], half_window: int = 0) -> np.ndarray:
    """Indices (ascending, i.e. still sorted by (contig, start)) of the reads that overlap any of the
    ranges +- half_window: the partition's reads including its halo (PartitionedRegions.scala:192)."""
    ends = batch.ends()
    keep = np.zeros(batch.n_reads, bool)
    for c, s, e in ranges:
        keep |= (batch.contig_id == c) & (batch.start < e + half_window) & (ends > s - half_window)
    return np.flatnonzero(keep)


def take_reads(batch: ReadBatch, idx: np.ndarray) -> ReadBatch:
    """Sub-batch with the given reads (order preserved)."""
    s0 = batch.seq_off[:-1][idx].astype(np.int64)
    s1 = batch.seq_off[1:][idx].astype(np.int64)
    c0 = batch.cigar_off[:-1][idx].astype(np.int64)
    c1 = batch.cigar_off[1:][idx].astype(np.int64)
    seq_idx = np.concatenate([np.arange(a, b) for a, b in zip(s0, s1)]) if len(idx) else np.zeros(0, np.int64)
    cig_idx = np.concatenate([np.arange(a, b) for a, b in zip(c0, c1)]) if len(idx) else np.zeros(0, np.int64)
    return ReadBatch(
        n_samples=batch.n_samples, contig_names=batch.contig_names, contig_id=batch.contig_id[idx], start=batch.start[idx],
        cigar_off=np.concatenate([[0], np.cumsum(c1 - c0)]).astype(np.uint32), cigar=batch.cigar[cig_idx],
        seq_off=np.concatenate([[0], np.cumsum(s1 - s0)]).astype(np.uint64), bases_ascii=batch.bases_ascii[seq_idx],
        qual=batch.qual[seq_idx], mapq=batch.mapq[idx], flags=batch.flags[idx], sample_id=batch.sample_id[idx])


def shard(batch: ReadBatch, loci: Sequence[Range], world_size: int, half_window: int = 0) -> Dict[int, Tuple[List[Range], ReadBatch]]:
    """rank -> (its loci ranges, its reads incl. halo).  One LociPartitioning entry per rank."""
    parts = uniform_partition(loci, world_size)
    out = {}
    for rank in range(world_size):
        rs = ranges_of_partition(parts, rank)
        out[rank] = (rs, take_reads(batch, reads_for_ranges(batch, rs, half_window)))
    return out


def assign_partitions(intervals: Sequence[Tuple[int, int, int, int]], n_partitions: int, world_size: int) -> Dict[int, List[Range]]:
    """rank -> loci ranges: the loci sets of the capped partitioner (`Context.capped_partitioning`,
    loci/partitioning/CappedRegionsPartitioner.scala:66-88 — each holds at most maxReadsPerPartition reads) dealt to the
    ranks in contiguous blocks, so every rank gets about the same number of reads whatever the coverage looks like."""
    out: Dict[int, List[Range]] = {r: [] for r in range(world_size)}
    for c, s, e, p in intervals:
        rank = min(world_size - 1, p * world_size // max(1, n_partitions))
        if out[rank] and out[rank][-1][0] == c and out[rank][-1][2] == s:
            out[rank][-1] = (c, out[rank][-1][1], e)
        else:
            out[rank].append((c, s, e))
    return out


def gather_calls(per_rank_results: Sequence[dict]) -> dict:
    """Host-side gather of the per-rank sparse results: concatenate, re-base the allele arena offsets,
    sort by (contig, start)."""
    arena = b""
    calls = []
    for r in per_rank_results:
        c = r["calls"].copy()
        c["ref_off"] += len(arena)
        c["alt_off"] += len(arena)
        arena += r["arena"]
        calls.append(c)
    allc = np.concatenate(calls) if calls else np.zeros(0)
    order = np.lexsort((allc["start"], allc["contig_id"]))
    return dict(calls=allc[order], arena=arena, n_visited=sum(r["n_visited"] for r in per_rank_results))
