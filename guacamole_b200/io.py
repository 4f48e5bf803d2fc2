"""Host-side decoding of SAM / BAM / FASTA fixtures into the read and reference model.

Decoding stays on the host (north star; reference: reads/Read.scala:63-116 for the SAMRecord ->
MappedRead mapping, reference/ReferenceBroadcast.scala:90-198 for FASTA and "partial FASTA").
Only what the hot path consumes is decoded: mapped reads, their flags, and upper-cased contigs.
"""
from __future__ import annotations

import gzip
import struct
from dataclasses import dataclass
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

from .reads import CIGAR_OPS, MappedRead, parse_cigar

_SEQ_NT16 = "=ACMGRSVTWYHKDBN"


def _read_from_fields(name: str, flag: int, rname: str, pos1: int, mapq: int, cigar_words: List[int], seq: bytes,
                      quals: bytes, sample_id: int) -> Optional[MappedRead]:
    # Read.apply(SAMRecord) (reads/Read.scala:63-96): mapped iff !unmapped flag, has a reference name,
    # alignment start >= 0 ...; start = getAlignmentStart - 1.
    unmapped = bool(flag & 0x4) or rname == "*" or pos1 <= 0
    if unmapped:
        return None
    if len(quals) == 0:
        quals = bytes(len(seq))
    return MappedRead(
        sequence=seq.upper(),
        base_qualities=quals,
        cigar=cigar_words,
        start=pos1 - 1,
        contig=rname,
        sample_id=sample_id,
        alignment_quality=mapq,
        is_positive_strand=not (flag & 0x10),
        is_paired=bool(flag & 0x1),
        is_duplicate=bool(flag & 0x400),
        failed_vendor_quality_checks=bool(flag & 0x200),
        name=name,
    )


def read_sam(path: str, sample_id: int = 0) -> Tuple[List[MappedRead], List[Tuple[str, int]]]:
    """Returns (mapped reads in file order, [(contig name, length)] from @SQ)."""
    reads: List[MappedRead] = []
    contigs: List[Tuple[str, int]] = []
    with open(path, "rt") as fh:
        for line in fh:
            if line.startswith("@"):
                if line.startswith("@SQ"):
                    f = dict(x.split(":", 1) for x in line.rstrip("\n").split("\t")[1:] if ":" in x)
                    contigs.append((f["SN"], int(f["LN"])))
                continue
            f = line.rstrip("\n").split("\t")
            if len(f) < 11:
                continue
            seq = b"" if f[9] == "*" else f[9].encode()
            quals = b"" if f[10] == "*" else bytes(c - 33 for c in f[10].encode())
            r = _read_from_fields(f[0], int(f[1]), f[2], int(f[3]), int(f[4]), parse_cigar(f[5]), seq, quals, sample_id)
            if r is not None:
                reads.append(r)
    return reads, contigs


def _bgzf_decompress(path: str) -> bytes:
    # BGZF is a series of gzip members; Python's gzip reads multi-member files.
    with gzip.open(path, "rb") as fh:
        return fh.read()


def read_bam(path: str, sample_id: int = 0) -> Tuple[List[MappedRead], List[Tuple[str, int]]]:
    data = _bgzf_decompress(path)
    assert data[:4] == b"BAM\1", "not a BAM file"
    (l_text,) = struct.unpack_from("<i", data, 4)
    p = 8 + l_text
    (n_ref,) = struct.unpack_from("<i", data, p)
    p += 4
    contigs = []
    for _ in range(n_ref):
        (l_name,) = struct.unpack_from("<i", data, p)
        p += 4
        name = data[p:p + l_name - 1].decode()
        p += l_name
        (l_ref,) = struct.unpack_from("<i", data, p)
        p += 4
        contigs.append((name, l_ref))
    reads = []
    n = len(data)
    while p + 4 <= n:
        (block_size,) = struct.unpack_from("<i", data, p)
        p += 4
        q = p
        ref_id, pos, l_read_name, mapq, _bin, n_cigar, flag, l_seq, _nref, _npos, _tlen = struct.unpack_from(
            "<iiBBHHHiiii", data, q)
        q += 32
        name = data[q:q + l_read_name - 1].decode()
        q += l_read_name
        cigar = list(struct.unpack_from(f"<{n_cigar}I", data, q))
        q += 4 * n_cigar
        packed = np.frombuffer(data, np.uint8, (l_seq + 1) // 2, q)
        q += (l_seq + 1) // 2
        seq = np.empty(2 * packed.shape[0], np.uint8)
        seq[0::2] = packed >> 4
        seq[1::2] = packed & 15
        seq_str = "".join(_SEQ_NT16[c] for c in seq[:l_seq]).encode()
        quals = data[q:q + l_seq]
        if l_seq and quals[0] == 0xFF:
            quals = bytes(l_seq)
        p += block_size
        rname = contigs[ref_id][0] if ref_id >= 0 else "*"
        r = _read_from_fields(name, flag, rname, pos + 1, mapq, cigar, seq_str, bytes(quals), sample_id)
        if r is not None:
            reads.append(r)
    return reads, contigs


def read_fasta(path: str) -> Iterator[Tuple[str, bytes]]:
    """(name = first word of header, upper-cased bases): ReferenceBroadcast.readFasta :90-104 (unmaskBases)."""
    opener = gzip.open if path.endswith(".gz") else open
    name, chunks = None, []
    with opener(path, "rt") as fh:
        for line in fh:
            line = line.strip()
            if line.startswith(">"):
                if name is not None:
                    yield name, "".join(chunks).upper().encode()
                name, chunks = line[1:].split()[0], []
            elif line:
                chunks.append(line)
    if name is not None:
        yield name, "".join(chunks).upper().encode()


@dataclass
class ContigWindow:
    """A dense window [offset, offset+len) of one contig; byte 0 = position absent
    (MapBackedReferenceSequence, reference/ReferenceBroadcast.scala:70-83)."""

    name: str
    offset: int
    bases: np.ndarray  # uint8
    contig_length: int

    def __getitem__(self, locus: int) -> int:
        i = locus - self.offset
        if i < 0 or i >= self.bases.shape[0] or self.bases[i] == 0:
            raise KeyError(f"Position {self.name}:{locus} missing from reference")
        return int(self.bases[i])

    def slice(self, start: int, end: int) -> bytes:
        return bytes(self[i] for i in range(start, end))


class Reference:
    """ReferenceGenome (reference/ReferenceGenome.scala:5) backed by dense per-contig windows."""

    def __init__(self, contigs: Dict[str, ContigWindow]):
        self.contigs = contigs

    def get_contig(self, name: str) -> ContigWindow:
        if name not in self.contigs:
            raise KeyError(f"Contig {name} does not exist in the current reference")
        return self.contigs[name]

    @staticmethod
    def from_fasta(path: str) -> "Reference":
        return Reference({n: ContigWindow(n, 0, np.frombuffer(s, np.uint8).copy(), len(s)) for n, s in read_fasta(path)})

    @staticmethod
    def from_partial_fasta(path: str) -> "Reference":
        """readPartialFasta (:119-198): headers look like ``chr1:9242255-9242454/249250621``."""
        pieces: Dict[str, List[Tuple[int, bytes]]] = {}
        lengths: Dict[str, int] = {}
        for desc, seq in read_fasta(path):
            loci_str, length_str = [x.strip() for x in desc.split("/")]
            contig, rng = loci_str.rsplit(":", 1)
            start, end = (int(x) for x in rng.split("-"))
            if end - start != len(seq):
                raise ValueError(f"In partial fasta, region {loci_str} is length {end - start} but its sequence is length {len(seq)}")
            if lengths.setdefault(contig, int(length_str)) != int(length_str):
                raise ValueError(f"In partial fasta, contig lengths for {contig} are inconsistent")
            pieces.setdefault(contig, []).append((start, seq))
        return Reference.from_pieces(pieces, lengths)

    @staticmethod
    def from_pieces(pieces: Dict[str, List[Tuple[int, bytes]]], lengths: Optional[Dict[str, int]] = None) -> "Reference":
        contigs = {}
        for contig, ps in pieces.items():
            lo = min(s for s, _ in ps)
            hi = max(s + len(q) for s, q in ps)
            arr = np.zeros(hi - lo, np.uint8)
            for s, q in ps:
                arr[s - lo:s - lo + len(q)] = np.frombuffer(q.upper() if isinstance(q, bytes) else q.upper().encode(), np.uint8)
            contigs[contig] = ContigWindow(contig, lo, arr, (lengths or {}).get(contig, hi))
        return Reference(contigs)


def make_reference(*contig_start_sequences: Tuple[str, int, str], contig_length: int = 1000) -> Reference:
    """Test helper mirroring ReferenceUtil.makeReference
    (src/test/scala/org/hammerlab/guacamole/reference/ReferenceUtil.scala:19-57)."""
    pieces: Dict[str, List[Tuple[int, bytes]]] = {}
    for contig, start, seq in contig_start_sequences:
        pieces.setdefault(contig, []).append((start, seq.upper().encode()))
    return Reference.from_pieces(pieces, {c: contig_length for c in pieces})
