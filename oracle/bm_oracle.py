"""ctypes binding of oracle/libbm_oracle.so -- TEST INFRASTRUCTURE ONLY.

Allowed importers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
Nothing under bucket-map_b200/ imports this."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbm_oracle.so")


class Params(C.Structure):
    _fields_ = [("bucket_len", C.c_uint32), ("read_len", C.c_uint32), ("q", C.c_uint32), ("k", C.c_uint32),
                ("mapper_samples", C.c_uint32), ("mapper_fault", C.c_uint32), ("distinguishability", C.c_float),
                ("avg_base_quality", C.c_uint32), ("max_candidates", C.c_uint32),
                ("num_segment_samples", C.c_uint32), ("locator_samples", C.c_uint32),
                ("allowed_mismatch", C.c_uint32), ("allowed_indel", C.c_uint32), ("indel_rate", C.c_float),
                ("quality_threshold", C.c_uint32)]


LOC_DTYPE = np.dtype([("bucket", "<u4"), ("offset", "<i4"), ("seg_offset", "<u4"), ("votes", "<u4"),
                      ("is_original", "<u4")])
ALN_DTYPE = np.dtype([("score", "<i4"), ("begin", "<u4"), ("map_qual", "<u4"), ("kept", "<u4"),
                      ("cigar_off", "<u4"), ("cigar_len", "<u4")])
_vp = C.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def build():
    """compile the restatement (gcc only)"""
    import subprocess
    src = os.path.join(_HERE, "bm_oracle.cpp")
    if (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "bm_oracle.h"))):
        subprocess.check_call(["g++", "-O3", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", LIB_PATH, src])


class Oracle:
    def __init__(self, threads=None):
        if not os.path.exists(LIB_PATH):
            build()
        self.lib = C.CDLL(LIB_PATH)
        self.threads = threads or min(os.cpu_count() or 1, 64)
        L = self.lib
        L.orc_fracminhash.restype = C.c_uint32
        L.orc_fracminhash.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64, C.c_float, _vp]
        L.orc_num_buckets.restype = C.c_uint32
        L.orc_num_buckets.argtypes = [_vp, C.c_uint32, C.c_uint32]
        L.orc_bucketize.restype = C.c_uint32
        L.orc_bucketize.argtypes = [_vp, C.c_uint32, C.c_uint32, C.c_uint32, _vp, _vp, _vp]
        L.orc_index_build.restype = None
        L.orc_index_build.argtypes = [_vp, _vp, _vp, C.c_uint32, C.c_uint32, C.c_uint32, _vp, C.c_uint32, _vp,
                                      C.c_uint32]
        L.orc_zeros.restype = C.c_uint32
        L.orc_zeros.argtypes = [_vp, C.c_uint32, C.c_uint32, C.c_float, _vp]
        L.orc_segments.restype = C.c_uint32
        L.orc_segments.argtypes = [C.POINTER(Params), _vp, C.c_uint32, _vp, _vp, _vp]
        L.orc_sample.restype = None
        L.orc_sample.argtypes = [C.c_uint32, C.c_uint32, _vp]
        L.orc_map_segments.restype = None
        L.orc_map_segments.argtypes = [C.POINTER(Params), C.c_uint32, _vp, _vp, C.c_uint32, _vp, _vp, _vp, _vp, _vp,
                                       _vp, C.c_uint32, _vp, _vp, _vp, C.c_uint32]
        L.orc_index_open.restype = _vp
        L.orc_index_open.argtypes = [C.POINTER(Params), C.c_uint32, _vp, _vp, C.c_uint32]
        L.orc_index_close.restype = None
        L.orc_index_close.argtypes = [_vp]
        L.orc_map_segments_h.restype = None
        L.orc_map_segments_h.argtypes = [_vp, C.POINTER(Params), _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint32, _vp, _vp,
                                         _vp, C.c_uint32]
        L.orc_prepare_query.restype = None
        L.orc_prepare_query.argtypes = [C.POINTER(Params), _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint32, _vp, _vp]
        L.orc_locate.restype = None
        L.orc_locate.argtypes = [C.POINTER(Params), _vp, _vp, _vp, C.c_uint32, _vp, _vp, _vp, C.c_uint32, _vp, _vp,
                                 _vp, C.c_uint32, _vp, _vp, _vp, _vp, C.c_uint32]
        L.orc_filter_best.restype = None
        L.orc_filter_best.argtypes = [C.POINTER(Params), _vp, C.c_uint32, _vp, _vp, _vp, _vp]
        L.orc_align.restype = C.c_uint64
        L.orc_align.argtypes = [C.POINTER(Params), _vp, _vp, _vp, _vp, _vp, C.c_uint32, _vp, _vp, _vp, _vp,
                                C.c_uint64, C.c_uint32]
        L.orc_derive_params.restype = None
        L.orc_derive_params.argtypes = [C.POINTER(Params), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                        C.c_uint32, C.c_float, C.c_uint32, C.c_float, C.c_float, C.c_uint32,
                                        C.c_uint32]
        L.orc_cli_index.restype = C.c_int
        L.orc_cli_index.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_uint64,
                                    C.c_uint64, C.c_char_p, C.c_char_p, C.c_uint32]
        L.orc_cli_map.restype = C.c_int
        L.orc_cli_map.argtypes = [C.c_char_p, C.POINTER(Params), C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p,
                                  C.c_int, C.c_uint32]

    def params(self, bucket_len, **kw):
        a = dict(r=300, q=9, k=12, s=15, d=0.5, b=25, e=0.4, n=0.02, p=10, u=40)
        a.update(kw)
        P = Params()
        self.lib.orc_derive_params(C.byref(P), bucket_len, a["r"], a["q"], a["k"], a["s"], a["d"], a["b"], a["e"],
                                   a["n"], a["p"], a["u"])
        return P

    def fracminhash(self, q, x, y, frac):
        k2i = np.empty(1 << (2 * q), np.int32)
        n = self.lib.orc_fracminhash(q, x, y, frac, _p(k2i))
        return k2i, n

    def bucketize(self, rec_len, L, r):
        rec_len = np.ascontiguousarray(rec_len, np.uint64)
        NB = self.lib.orc_num_buckets(_p(rec_len), len(rec_len), L)
        bs, bl, br = np.empty(NB + 1, np.uint64), np.empty(NB + 1, np.uint32), np.empty(NB + 1, np.uint32)
        n = self.lib.orc_bucketize(_p(rec_len), len(rec_len), L, r, _p(bs), _p(bl), _p(br))
        return NB, bs[:n].copy(), bl[:n].copy(), br[:n].copy()

    def index_build(self, packed, b_start, b_len, NB, q, k2i, n_rows):
        rows = np.empty((n_rows, (NB + 7) >> 3), np.uint8)
        self.lib.orc_index_build(_p(packed), _p(b_start), _p(b_len), len(b_start), NB, q, _p(k2i), n_rows, _p(rows),
                                 self.threads)
        return rows

    def zeros(self, rows, NB, d):
        z = np.empty(rows.shape[0], np.uint32)
        v = self.lib.orc_zeros(_p(rows), rows.shape[0], NB, d, _p(z))
        return z, v

    def segments(self, P, read_off):
        n = len(read_off) - 1
        cap = n * max(1, P.num_segment_samples)
        a, b, c = (np.empty(cap, np.uint32) for _ in range(3))
        ns = self.lib.orc_segments(C.byref(P), _p(read_off), n, _p(a), _p(b), _p(c))
        return a[:ns].copy(), b[:ns].copy(), c[:ns].copy()

    def sample(self, n, ub):
        out = np.empty(n, np.uint32)
        self.lib.orc_sample(n, ub, _p(out))
        return out

    def map_segments(self, P, NB, k2i, rows, seq, qual, read_off, segs, threads=None):
        sr, ss, sl = segs
        n_seg = len(sr)
        off = np.empty(2 * n_seg + 1, np.uint32)
        cand = np.empty(2 * n_seg * P.max_candidates + 1, np.uint32)
        stats = np.zeros(4, np.uint64)
        self.lib.orc_map_segments(C.byref(P), NB, _p(k2i), _p(rows), rows.shape[0], _p(seq), _p(qual), _p(read_off),
                                  _p(sr), _p(ss), _p(sl), n_seg, _p(off), _p(cand), _p(stats),
                                  threads or self.threads)
        return off, cand[:int(off[-1])].copy(), stats

    def index_open(self, P, NB, k2i, rows):
        return self.lib.orc_index_open(C.byref(P), NB, _p(k2i), _p(rows), rows.shape[0])

    def index_close(self, ix):
        self.lib.orc_index_close(ix)

    def map_segments_h(self, ix, P, seq, qual, read_off, segs, threads=None):
        sr, ss, sl = segs
        n_seg = len(sr)
        off = np.empty(2 * n_seg + 1, np.uint32)
        cand = np.empty(2 * n_seg * P.max_candidates + 1, np.uint32)
        stats = np.zeros(4, np.uint64)
        self.lib.orc_map_segments_h(ix, C.byref(P), _p(seq), _p(qual), _p(read_off), _p(sr), _p(ss), _p(sl), n_seg,
                                    _p(off), _p(cand), _p(stats), threads or self.threads)
        return off, cand[:int(off[-1])].copy(), stats

    def prepare_query(self, P, seq, qual, read_off, segs):
        sr, ss, sl = segs
        n_seg = len(sr)
        idx = np.empty((n_seg, P.locator_samples), np.uint16)
        h = np.empty((n_seg, P.locator_samples), np.uint32)
        self.lib.orc_prepare_query(C.byref(P), _p(seq), _p(qual), _p(read_off), _p(sr), _p(ss), _p(sl), n_seg,
                                   _p(idx), _p(h))
        return idx, h

    def locate(self, P, packed, b_start, b_len, seq, qual, read_off, segs, cand_off, cand, threads=None):
        sr, ss, sl = segs
        n = len(read_off) - 1
        loc_off = np.empty(n + 1, np.uint32)
        locs = np.empty(len(cand) + 1, LOC_DTYPE)
        self.lib.orc_locate(C.byref(P), _p(packed), _p(b_start), _p(b_len), len(b_start), _p(seq), _p(qual),
                            _p(read_off), n, _p(sr), _p(ss), _p(sl), len(sr), _p(cand_off), _p(cand), _p(loc_off),
                            _p(locs), threads or self.threads)
        return loc_off, locs[:int(loc_off[-1])].copy()

    def filter_best(self, P, read_off, loc_off, locs):
        n = len(read_off) - 1
        out_off = np.empty(n + 1, np.uint32)
        out = np.empty(len(locs) + 1, LOC_DTYPE)
        self.lib.orc_filter_best(C.byref(P), _p(read_off), n, _p(loc_off), _p(locs), _p(out_off), _p(out))
        return out_off, out[:int(out_off[-1])].copy()

    def align(self, P, packed, b_start, b_len, seq, read_off, loc_off, locs, threads=None):
        n = len(read_off) - 1
        aln = np.empty(len(locs), ALN_DTYPE)
        cap = max(1, len(locs) * 16)
        cig = np.empty(cap, np.uint32)
        need = self.lib.orc_align(C.byref(P), _p(packed), _p(b_start), _p(b_len), _p(seq), _p(read_off), n,
                                  _p(loc_off), _p(locs), _p(aln), _p(cig), cap, threads or self.threads)
        if need > cap:
            cig = np.empty(need, np.uint32)
            self.lib.orc_align(C.byref(P), _p(packed), _p(b_start), _p(b_len), _p(seq), _p(read_off), n, _p(loc_off),
                               _p(locs), _p(aln), _p(cig), need, threads or self.threads)
        return aln, cig[:need].copy()

    def cli_index(self, fasta, L, r, q, frac, x, y, index_dir, indicator):
        return self.lib.orc_cli_index(fasta.encode(), L, r, q, frac, x, y, index_dir.encode(), indicator.encode(),
                                      self.threads)

    def cli_map(self, fasta, P, index_dir, indicator, fastq, sam, align=False):
        return self.lib.orc_cli_map(fasta.encode(), C.byref(P), index_dir.encode(), indicator.encode(),
                                    fastq.encode(), sam.encode(), int(align), self.threads)
