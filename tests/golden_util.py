"""Access to tests/golden/* (outputs of the reference itself, see tests/golden/make_golden.py)."""
import glob
import gzip
import hashlib
import json
import os

import numpy as np

from bucketmap_b200 import datagen

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(GOLDEN, "*.json")))
ALIGN_CASES = [c for c in CASES if os.path.exists(os.path.join(GOLDEN, c + ".align.sam.gz"))]


class Golden:
    def __init__(self, name):
        self.name = name
        self.meta = json.load(open(os.path.join(GOLDEN, name + ".json")))
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.packed, self.rec_len = z["packed"], z["rec_len"]
        self.names = [str(s) for s in z["names"]]
        self.seq, self.qual, self.read_off = z["seq"], z["qual"], z["read_off"]
        self.read_names = [str(s) for s in z["read_names"]]
        self.NB, self.L = self.meta["NB"], self.meta["L"]
        self.x, self.y = self.meta["x"], self.meta["y"]
        self.flags = self.meta["flags"]
        self.n_bases = int(self.rec_len.sum())

    def sam(self, align=False):
        p = os.path.join(GOLDEN, self.name + (".align.sam.gz" if align else ".sam.gz"))
        return gzip.open(p, "rb").read().decode()

    def write_inputs(self, d):
        fa, fq = os.path.join(d, "genome.fa"), os.path.join(d, "reads.fq")
        datagen.write_fasta(fa, self.packed, self.rec_len, self.names)
        datagen.write_fastq(fq, self.seq, self.qual, self.read_off, self.read_names)
        return fa, fq

    def params_kw(self):
        """reference CLI letters -> keyword names of Oracle.params / bucketmap_b200.default_params
        (CLI -k = index seed q, -l = query seed k)"""
        m = {"k": "q", "l": "k", "r": "r", "s": "s", "d": "d", "b": "b", "e": "e", "n": "n", "p": "p", "u": "u"}
        return {m[k]: v for k, v in self.flags.items() if k in m}

    @property
    def frac(self):
        return float(self.flags.get("f", 0.25))


def sha256_file(p):
    return hashlib.sha256(open(p, "rb").read()).hexdigest()


def sam_records(text):
    """(header lines, record tuples of the 11 mandatory fields)"""
    hdr, recs = [], []
    for line in text.splitlines():
        if line.startswith("@"):
            hdr.append(line)
        elif line:
            recs.append(tuple(line.split("\t")[:11]))
    return hdr, recs
