"""Loads the committed golden fixtures (see make_fixtures.py for how they were produced)."""
import os

import numpy as np

from guacamole_b200.io import ContigWindow, Reference
from guacamole_b200.reads import ReadBatch

HERE = os.path.dirname(os.path.abspath(__file__))
BATCH_FIELDS = ["contig_id", "start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id"]


def load(name: str):
    """-> (ReadBatch, Reference or None, dict of extras)"""
    z = np.load(os.path.join(HERE, name + ".npz"), allow_pickle=False)
    batch = ReadBatch(n_samples=int(z["n_samples"]), contig_names=[str(x) for x in z["contig_names"]],
                      **{f: np.ascontiguousarray(z[f]) for f in BATCH_FIELDS})
    ref = None
    if "ref_names" in z.files:
        contigs = {}
        for i, n in enumerate(z["ref_names"]):
            starts, lens, bases = z[f"ref_{i}_starts"], z[f"ref_{i}_lens"], z[f"ref_{i}_bases"]
            lo, hi = int(starts.min()), int((starts + lens).max())
            arr = np.zeros(hi - lo, np.uint8)
            o = 0
            for s, l in zip(starts, lens):
                arr[int(s) - lo:int(s) - lo + int(l)] = bases[o:o + int(l)]
                o += int(l)
            contigs[str(n)] = ContigWindow(str(n), lo, arr, int(z[f"ref_{i}_length"]))
        ref = Reference(contigs)
    extras = {k: z[k] for k in z.files if k in ("golden_vcf", "force_call_vcf", "positions")}
    return batch, ref, extras
