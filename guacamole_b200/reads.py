"""Host-side read model and the columnar (SoA) batch that crosses the C ABI.

Mirrors the data contract of the reference's ``MappedRead``
(reference: src/main/scala/org/hammerlab/guacamole/reads/MappedRead.scala:21-33) and the
read filters of ``ReadSets.filterRDD`` (readsets/ReadSets.scala:272-304).  BAM/SAM decoding
stays on the host (north star); this module only shapes decoded reads into the layout of
``guac_read_batch`` (include/guac_b200.h).
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import Iterable, List, Optional, Sequence

import numpy as np

CIGAR_OPS = "MIDNSHP=X"
_CIGAR_RE = re.compile(r"(\d+)([MIDNSHP=X])")

FLAG_POSITIVE_STRAND = 1
FLAG_PAIRED = 2
FLAG_DUPLICATE = 4
FLAG_FAILED_QC = 8


def parse_cigar(text: str) -> List[int]:
    """TextCigarCodec.decode -> BAM-encoded words (len << 4 | op)."""
    if text == "*" or text == "":
        return []
    out = []
    pos = 0
    for m in _CIGAR_RE.finditer(text):
        if m.start() != pos:
            raise ValueError(f"bad CIGAR {text!r}")
        pos = m.end()
        out.append((int(m.group(1)) << 4) | CIGAR_OPS.index(m.group(2)))
    if pos != len(text):
        raise ValueError(f"bad CIGAR {text!r}")
    return out


def cigar_to_string(words: Sequence[int]) -> str:
    return "".join(f"{w >> 4}{CIGAR_OPS[w & 15]}" for w in words)


@dataclass
class MappedRead:
    """reads/MappedRead.scala:21-33 (0-based start)."""

    sequence: bytes
    base_qualities: bytes  # Phred values, not +33
    cigar: List[int]  # BAM-encoded
    start: int
    contig: str = "chr1"
    sample_id: int = 0
    alignment_quality: int = 30
    is_positive_strand: bool = True
    is_paired: bool = True
    is_duplicate: bool = False
    failed_vendor_quality_checks: bool = False
    name: str = "read1"

    @property
    def end(self) -> int:
        # MappedRead.scala:75  start + cigar.getPaddedReferenceLength (M D N = X P)
        return self.start + sum(w >> 4 for w in self.cigar if (w & 15) in (0, 2, 3, 6, 7, 8))


def make_read(sequence: str, cigar: str, start: int = 1, chr: str = "chr1", sample_id: int = 0,
              quality_scores: Optional[Sequence[int]] = None, alignment_quality: int = 30,
              is_positive_strand: bool = True) -> MappedRead:
    """Test helper with the defaults of the reference's ReadsUtil.makeRead
    (src/test/scala/org/hammerlab/guacamole/reads/ReadsUtil.scala:62-86): base quality '@' = Phred 31."""
    quals = bytes(quality_scores) if quality_scores is not None else bytes([31] * len(sequence))
    assert len(quals) == len(sequence)
    return MappedRead(sequence.upper().encode(), quals, parse_cigar(cigar), start, chr, sample_id,
                      alignment_quality, is_positive_strand)


@dataclass
class ReadBatch:
    """Columnar batch = one partition of PartitionedRegions (readsets/rdd/PartitionedRegions.scala:56),
    sorted by (contig_id, start).  ``bases_ascii`` is kept for host-side consumers (materialising
    Pileup objects, the test oracle); the device receives the packed form built lazily by
    ``guacamole_b200.native.pack_batch``."""

    n_samples: int
    contig_names: List[str]
    contig_id: np.ndarray
    start: np.ndarray
    cigar_off: np.ndarray
    cigar: np.ndarray
    seq_off: np.ndarray
    bases_ascii: np.ndarray
    qual: np.ndarray
    mapq: np.ndarray
    flags: np.ndarray
    sample_id: np.ndarray
    packed: Optional[dict] = field(default=None, repr=False)

    @property
    def n_reads(self) -> int:
        return int(self.start.shape[0])

    def read(self, i: int) -> MappedRead:
        s0, s1 = int(self.seq_off[i]), int(self.seq_off[i + 1])
        c0, c1 = int(self.cigar_off[i]), int(self.cigar_off[i + 1])
        f = int(self.flags[i])
        return MappedRead(self.bases_ascii[s0:s1].tobytes(), self.qual[s0:s1].tobytes(),
                          [int(w) for w in self.cigar[c0:c1]], int(self.start[i]),
                          self.contig_names[int(self.contig_id[i])], int(self.sample_id[i]), int(self.mapq[i]),
                          bool(f & FLAG_POSITIVE_STRAND), bool(f & FLAG_PAIRED), bool(f & FLAG_DUPLICATE),
                          bool(f & FLAG_FAILED_QC))

    @staticmethod
    def from_reads(reads: Iterable[MappedRead], contig_names: Optional[List[str]] = None,
                   n_samples: Optional[int] = None) -> "ReadBatch":
        reads = list(reads)
        if contig_names is None:
            contig_names = sorted({r.contig for r in reads})
        cid = {c: i for i, c in enumerate(contig_names)}
        # stable sort by (contig, start): PartitionedRegions.scala:207 repartitionAndSortWithinPartitions
        reads.sort(key=lambda r: (cid[r.contig], r.start))
        n = len(reads)
        cigar_off = np.zeros(n + 1, np.uint32)
        seq_off = np.zeros(n + 1, np.uint64)
        for i, r in enumerate(reads):
            assert len(r.sequence) == len(r.base_qualities), "Base qualities / sequence length mismatch"
            cigar_off[i + 1] = cigar_off[i] + len(r.cigar)
            seq_off[i + 1] = seq_off[i] + np.uint64(len(r.sequence))
        flags = np.array([(FLAG_POSITIVE_STRAND if r.is_positive_strand else 0) | (FLAG_PAIRED if r.is_paired else 0) |
                          (FLAG_DUPLICATE if r.is_duplicate else 0) |
                          (FLAG_FAILED_QC if r.failed_vendor_quality_checks else 0) for r in reads], np.uint8)
        if n_samples is None:
            n_samples = (max((r.sample_id for r in reads), default=0) + 1)
        return ReadBatch(
            n_samples=n_samples,
            contig_names=list(contig_names),
            contig_id=np.array([cid[r.contig] for r in reads], np.int32),
            start=np.array([r.start for r in reads], np.int64),
            cigar_off=cigar_off,
            cigar=np.array([w for r in reads for w in r.cigar], np.uint32),
            seq_off=seq_off,
            bases_ascii=np.frombuffer(b"".join(r.sequence for r in reads), np.uint8).copy(),
            qual=np.frombuffer(b"".join(r.base_qualities for r in reads), np.uint8).copy(),
            mapq=np.array([min(r.alignment_quality, 255) for r in reads], np.uint8),
            flags=flags,
            sample_id=np.array([r.sample_id for r in reads], np.uint8),
        )

    def ends(self) -> np.ndarray:
        """read.end for every read (vectorised MappedRead.scala:75)."""
        ops = self.cigar & 15
        lens = (self.cigar >> 4).astype(np.int64)
        consumes = np.isin(ops, (0, 2, 3, 6, 7, 8))
        csum = np.concatenate([[0], np.cumsum(np.where(consumes, lens, 0))])
        return self.start + (csum[self.cigar_off[1:].astype(np.int64)] - csum[self.cigar_off[:-1].astype(np.int64)])


def filter_reads(reads: Iterable[MappedRead], non_duplicate: bool = True, passed_vendor_quality_checks: bool = True,
                 is_paired: bool = True, min_alignment_quality: Optional[int] = None) -> List[MappedRead]:
    """ReadSets.filterRDD (readsets/ReadSets.scala:272-304) with the defaults of
    ReadFilterArgs.parseConfig (readsets/io/ReadFilterArgs.scala:50-63)."""
    out = []
    for r in reads:
        if non_duplicate and r.is_duplicate:
            continue
        if passed_vendor_quality_checks and r.failed_vendor_quality_checks:
            continue
        if is_paired and not r.is_paired:
            continue
        if min_alignment_quality is not None and r.alignment_quality < min_alignment_quality:
            continue
        out.append(r)
    return out
