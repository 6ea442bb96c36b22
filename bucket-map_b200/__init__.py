"""ctypes binding of libbucketmap_b200.so (the C-ABI in include/bucketmap_b200.h).

This is plumbing for tests/ and bench.py: the product is the shared library and the C++ host in host/.
There is no CPU fallback: if the library is missing or no CUDA device is usable, construction raises.
The directory name has a hyphen, so import it through the root-level ``bucketmap_b200`` module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbucketmap_b200.so")

(K_ROWPOP, K_INDEX, K_SAMPLE, K_MAP, K_LOCSAMPLE, K_LOCATE, K_MERGE, K_ALIGN, K_AUX, K_KTABLE, K_GIDX, K_LOCSCAN,
 K_MAPOVER, K_PACK, K_GATHER, K_COUNT) = range(16)
KERNEL_NAMES = ["rowpop", "index_build", "map_sample", "map_vote", "loc_sample", "locate", "merge", "align", "aux",
                "kmer_table", "locator_index", "locate_scan", "map_vote_handover", "pack_reads", "gather_pack"]
COMM_ID_BYTES = 128
PEER_HANDLE_BYTES = 128


class Params(C.Structure):
    """bm_params (main.cpp:12-39 plus the float-derived thresholds)."""
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

_u8p, _u32p, _u64p, _i32p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint32, C.c_uint64, C.c_int32))
_vp = C.c_void_p

# name -> (restype, argtypes); every symbol include/bucketmap_b200.h declares
SIGNATURES = {
    "bm_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "bm_ctx_destroy": (None, [_vp]),
    "bm_last_error": (C.c_char_p, [_vp]),
    "bm_version": (C.c_int, []),
    "bm_params_derive": (None, [C.POINTER(Params), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                C.c_float, C.c_uint32, C.c_float, C.c_float, C.c_uint32, C.c_uint32]),
    "bm_params_default": (None, [C.POINTER(Params), C.c_uint32]),
    "bm_ref_upload": (C.c_int, [_vp, _vp, C.c_uint64, _vp, _vp, C.c_uint32, C.c_uint32]),
    "bm_index_build": (C.c_int, [_vp, _vp, C.c_uint32, C.c_uint32, _vp]),
    "bm_index_load": (C.c_int, [_vp, _vp, C.c_uint32, _vp, C.c_uint32, C.c_uint32]),
    "bm_index_download": (C.c_int, [_vp, _vp]),
    "bm_index_zeros": (C.c_int, [_vp, C.c_float, _vp, C.POINTER(C.c_uint32)]),
    "bm_index_kmer_table": (C.c_int, [_vp, C.c_int, C.c_uint64]),
    "bm_index_kmer_table_info": (C.c_int, [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]),
    "bm_index_kmer_table_choice": (C.c_int, [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_double)]),
    "bm_map_handover": (C.c_int, [_vp, C.c_int]),
    "bm_index_fine_table": (C.c_int, [_vp, C.c_int]),
    "bm_index_fine_table_info": (C.c_int, [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]),
    "bm_pack_layout": (C.c_int, [_vp, C.c_uint32, _vp]),
    "bm_pack_reads": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _vp, _vp,
                                C.POINTER(C.c_uint32)]),
    "bm_pack_reads_v": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _vp, _vp,
                                  C.POINTER(C.c_uint32)]),
    "bm_prepare": (C.c_int, [_vp, C.POINTER(Params)]),
    "bm_reads_upload_packed": (C.c_int, [_vp, C.POINTER(Params), _vp, _vp, _vp, _vp, C.c_uint32]),
    "bm_map_locate_packed": (C.c_int, [_vp, C.POINTER(Params), _vp, _vp, _vp, _vp, C.c_uint32, C.c_int, _vp, _vp,
                                       C.c_uint64, _vp]),
    "bm_index_free": (C.c_int, [_vp]),
    "bm_locator_index": (C.c_int, [_vp, C.c_int]),
    "bm_index_alloc": (C.c_int, [_vp, _vp, C.c_uint32, C.c_uint32, C.c_uint32]),
    "bm_index_device": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(C.c_uint64)]),
    "bm_ref_device": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(C.c_uint64)]),
    "bm_reads_upload": (C.c_int, [_vp, C.POINTER(Params), _vp, _vp, _vp, C.c_uint32]),
    "bm_reads_segments": (C.c_int, [_vp, C.POINTER(C.c_uint32), _vp, _vp, _vp]),
    "bm_map_device": (C.c_int, [_vp, C.POINTER(Params)]),
    "bm_map_download": (C.c_int, [_vp, _vp, _vp, C.c_uint64, _vp]),
    "bm_map_reads": (C.c_int, [_vp, C.POINTER(Params), _vp, _vp, _vp, C.c_uint32, _vp, _vp, C.c_uint64, _vp]),
    "bm_locate_device": (C.c_int, [_vp, C.POINTER(Params), C.c_int]),
    "bm_locate_download": (C.c_int, [_vp, _vp, _vp, C.c_uint64]),
    "bm_locate_count": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "bm_locate_device_buffers": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_uint64)]),
    "bm_map_locate": (C.c_int, [_vp, C.POINTER(Params), _vp, _vp, _vp, C.c_uint32, C.c_int, _vp, _vp, C.c_uint64,
                                _vp]),
    "bm_pipeline_lanes": (C.c_int, [_vp, C.c_uint32]),
    "bm_align_noalloc": (C.c_int, [_vp, C.POINTER(Params)] + [C.c_void_p] * 6 + [C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                                                   C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bm_index_build_sharded": (C.c_int, [_vp, C.c_void_p, C.c_uint32, C.c_uint32]),
    "bm_ctx_create_lane": (C.c_int, [_vp, C.POINTER(Params), C.POINTER(_vp)]),
    "bm_align_device": (C.c_int, [_vp, C.POINTER(Params)]),
    "bm_align_download": (C.c_int, [_vp, _vp, _vp, C.c_uint64, C.POINTER(C.c_uint64)]),
    "bm_comm_unique_id": (C.c_int, [_vp]),
    "bm_comm_init": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "bm_comm_attach": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "bm_comm_destroy": (C.c_int, [_vp]),
    "bm_bcast_index": (C.c_int, [_vp, C.c_int]),
    "bm_bcast_ref": (C.c_int, [_vp, C.c_int]),
    "bm_gather_open": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint64, _vp]),
    "bm_gather_connect": (C.c_int, [_vp, _vp]),
    "bm_gather_results": (C.c_int, [_vp, _vp, _vp, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "bm_gather_sync": (C.c_int, [_vp]),
    "bm_gather_host_async": (C.c_int, [_vp, C.c_int]),
    "bm_gather_join": (C.c_int, [_vp]),
    "bm_gather_device_buffers": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_uint64),
                                           C.POINTER(C.c_uint64)]),
    "bm_gather_download": (C.c_int, [_vp, _vp, _vp, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "bm_gather_close": (C.c_int, [_vp]),
    "bm_timer_start": (C.c_int, [_vp]),
    "bm_timer_stop": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "bm_kernel_timing": (C.c_int, [_vp, C.c_int]),
    "bm_kernel_stats": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    "bm_kernel_stats_reset": (C.c_int, [_vp]),
    "bm_flush_l2": (C.c_int, [_vp, C.c_uint64]),
    "bm_sync": (C.c_int, [_vp]),
    "bm_host_alloc": (_vp, [C.c_uint64]),
    "bm_host_free": (None, [_vp]),
}


class BucketMapError(RuntimeError):
    pass


def load_library(path=LIB_PATH):
    """dlopen the product library and attach prototypes.  Raises if it has not been built (no fallback)."""
    if not os.path.exists(path):
        raise BucketMapError(f"{path} is missing: run `make lib` (python -c 'import __graft_entry__ as g; g.build()')")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here = stale library: rebuild with `make lib`
        fn.restype, fn.argtypes = res, args
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def pack_reads(lib, seq, qual, read_off, k, b, threads=1):
    """host-side packer (bm_pack_layout + bm_pack_reads): returns (planes uint64, qk uint32, pl_off uint32, n_low)"""
    n = len(read_off) - 1
    read_off = np.ascontiguousarray(read_off, np.uint64)
    pl_off = np.empty(n + 1, np.uint32)
    rc = lib.bm_pack_layout(_ptr(read_off), n, _ptr(pl_off))
    if rc:
        raise BucketMapError(f"bm_pack_layout failed ({rc})")
    planes = np.zeros(int(pl_off[n]) + 1, np.uint64)
    qk = np.zeros(int(pl_off[n]) + 1, np.uint32)
    low = C.c_uint32()
    rc = lib.bm_pack_reads(_ptr(seq), _ptr(qual), _ptr(read_off), _ptr(pl_off), 0, n, k, b, _ptr(planes), _ptr(qk),
                           C.byref(low))
    if rc:
        raise BucketMapError(f"bm_pack_reads failed ({rc})")
    return planes, qk, pl_off, low.value


def default_params(lib, bucket_len, **kw):
    """cmd_arguments defaults (main.cpp:12-39); keyword overrides use the CLI meaning (r,q,k,s,d,b,e,n,p,u)."""
    a = dict(r=300, q=9, k=12, s=15, d=0.5, b=25, e=0.4, n=0.02, p=10, u=40)
    a.update(kw)
    P = Params()
    lib.bm_params_derive(C.byref(P), bucket_len, a["r"], a["q"], a["k"], a["s"], a["d"], a["b"], a["e"], a["n"],
                         a["p"], a["u"])
    return P


class Context:
    """One bm_ctx (one GPU)."""

    def __init__(self, device=0, lib=None):
        self.lib = lib or load_library()
        h = _vp()
        rc = self.lib.bm_ctx_create(device, C.byref(h))
        if rc != 0:
            raise BucketMapError(f"bm_ctx_create(device={device}) failed with {rc}: no usable CUDA device "
                                 "(this library has no CPU fallback)")
        self.h = h
        self._keep = []

    def align_noalloc(self, P, d_packed, d_bstart, d_blen, d_planes, d_pl_off, d_read_off, n_reads, max_read_len,
                      d_loc_off, d_locs, n_aln, d_aln, d_cigar, d_cigar_cnt):
        """bm_align_noalloc: every d_* is a DEVICE address (int); enqueues on the context's stream, allocates nothing"""
        self._ck(self.lib.bm_align_noalloc(self.h, C.byref(P), d_packed, d_bstart, d_blen, d_planes, d_pl_off, d_read_off,
                                           n_reads, max_read_len, d_loc_off, d_locs, n_aln, d_aln, d_cigar, d_cigar_cnt),
                 "bm_align_noalloc")

    def index_build_sharded(self, k2i, q, n_rows):
        """collective: every rank indexes its share of the buckets, the matrices are summed over NCCL"""
        k2i = np.ascontiguousarray(k2i, np.int32)
        self._ck(self.lib.bm_index_build_sharded(self.h, _ptr(k2i), q, n_rows), "bm_index_build_sharded")
        self.n_rows = n_rows

    def create_lane(self, P):
        """a second context on this GPU viewing this one's reference, index and tables (bm_ctx_create_lane): drive it from
        another host thread to keep two batches in flight.  Close it before this context."""
        h = _vp()
        self._ck(self.lib.bm_ctx_create_lane(self.h, C.byref(P), C.byref(h)), "bm_ctx_create_lane")
        lane = Context.__new__(Context)
        lane.lib, lane.h, lane._keep = self.lib, h, []
        return lane

    def close(self):
        if getattr(self, "h", None):
            self.lib.bm_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            msg = self.lib.bm_last_error(self.h)
            raise BucketMapError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    # ---- reference + index ---------------------------------------------------------------------------------------
    def ref_upload(self, packed, n_bases, b_start, b_len, NB):
        packed = np.ascontiguousarray(packed, np.uint8)
        b_start = np.ascontiguousarray(b_start, np.uint64)
        b_len = np.ascontiguousarray(b_len, np.uint32)
        self._ck(self.lib.bm_ref_upload(self.h, _ptr(packed), n_bases, _ptr(b_start), _ptr(b_len), len(b_start), NB),
                 "bm_ref_upload")
        self.NB = NB

    def index_build(self, k2i, q, n_rows, download=True):
        k2i = np.ascontiguousarray(k2i, np.int32)
        W = (self.NB + 7) >> 3
        out = np.empty((n_rows, W), np.uint8) if download else None
        self._ck(self.lib.bm_index_build(self.h, _ptr(k2i), q, n_rows, _ptr(out)), "bm_index_build")
        self.n_rows = n_rows
        return out

    def index_load(self, k2i, q, rows, NB):
        k2i = np.ascontiguousarray(k2i, np.int32)
        rows = np.ascontiguousarray(rows, np.uint8)
        n_rows = rows.shape[0]
        self._ck(self.lib.bm_index_load(self.h, _ptr(k2i), q, _ptr(rows), n_rows, NB), "bm_index_load")
        self.NB, self.n_rows = NB, n_rows

    def index_download(self):
        out = np.empty((self.n_rows, (self.NB + 7) >> 3), np.uint8)
        self._ck(self.lib.bm_index_download(self.h, _ptr(out)), "bm_index_download")
        return out

    def index_zeros(self, d):
        z = np.empty(self.n_rows, np.uint32)
        v = C.c_uint32()
        self._ck(self.lib.bm_index_zeros(self.h, d, _ptr(z), C.byref(v)), "bm_index_zeros")
        return z, v.value

    def index_kmer_table(self, mode, max_bytes=0):
        """0: the mapper gathers the k-q+1 q-gram rows per sample; 1 (default): one k-mer row when 4^k rows fit."""
        self._ck(self.lib.bm_index_kmer_table(self.h, mode, max_bytes), "bm_index_kmer_table")

    def index_kmer_table_info(self):
        k, n = C.c_uint32(), C.c_uint64()
        self._ck(self.lib.bm_index_kmer_table_info(self.h, C.byref(k), C.byref(n)), "bm_index_kmer_table_info")
        return k.value, n.value

    def index_fine_table(self, mode):
        """0: passing groups are recounted from the .qgram words; 1 (default): from the full-resolution k-mer table when
        it fits"""
        self._ck(self.lib.bm_index_fine_table(self.h, mode), "bm_index_fine_table")

    def index_fine_table_info(self):
        k, n = C.c_uint32(), C.c_uint64()
        self._ck(self.lib.bm_index_fine_table_info(self.h, C.byref(k), C.byref(n)), "bm_index_fine_table_info")
        return k.value, n.value

    def index_kmer_table_choice(self):
        """(buckets per coarse bit of the resident table or 0, measured k-mer row density)"""
        g, d = C.c_uint32(), C.c_double()
        self._ck(self.lib.bm_index_kmer_table_choice(self.h, C.byref(g), C.byref(d)), "bm_index_kmer_table_choice")
        return g.value, d.value

    def map_handover(self, max_groups):
        """lists with more than max_groups passing coarse groups go to the full-width gather (-1: default)"""
        self._ck(self.lib.bm_map_handover(self.h, max_groups), "bm_map_handover")

    def pipeline_lanes(self, min_reads):
        """batches of >= min_reads reads go through bm_map_locate on two lanes (0 / 0xFFFFFFFF: never, the default)"""
        self._ck(self.lib.bm_pipeline_lanes(self.h, min_reads), "bm_pipeline_lanes")

    def locator_index(self, mode):
        """0: the locator scans each candidate bucket's text; 1 (default): genome k-mer position index when it fits."""
        self._ck(self.lib.bm_locator_index(self.h, mode), "bm_locator_index")

    def prepare(self, P):
        """build the derived tables now instead of at the first batch"""
        self._ck(self.lib.bm_prepare(self.h, C.byref(P)), "bm_prepare")

    def index_free(self):
        self._ck(self.lib.bm_index_free(self.h), "bm_index_free")

    def index_device(self):
        p, n = _vp(), C.c_uint64()
        self._ck(self.lib.bm_index_device(self.h, C.byref(p), C.byref(n)), "bm_index_device")
        return p.value, n.value

    def ref_device(self):
        p, n = _vp(), C.c_uint64()
        self._ck(self.lib.bm_ref_device(self.h, C.byref(p), C.byref(n)), "bm_ref_device")
        return p.value, n.value

    # ---- reads ---------------------------------------------------------------------------------------------------
    def reads_upload(self, P, seq, qual, read_off):
        self.n_reads = len(read_off) - 1
        self._ck(self.lib.bm_reads_upload(self.h, C.byref(P), _ptr(seq), _ptr(qual), _ptr(read_off), self.n_reads),
                 "bm_reads_upload")
        n = C.c_uint32()
        self._ck(self.lib.bm_reads_segments(self.h, C.byref(n), None, None, None), "bm_reads_segments")
        self.n_seg = n.value

    def reads_upload_packed(self, P, planes, qk, pl_off, read_off):
        self.n_reads = len(read_off) - 1
        self._ck(self.lib.bm_reads_upload_packed(self.h, C.byref(P), _ptr(planes), _ptr(qk), _ptr(pl_off), _ptr(read_off),
                                                 self.n_reads), "bm_reads_upload_packed")
        n = C.c_uint32()
        self._ck(self.lib.bm_reads_segments(self.h, C.byref(n), None, None, None), "bm_reads_segments")
        self.n_seg = n.value

    def segments(self):
        nq = C.c_uint32()
        self._ck(self.lib.bm_reads_segments(self.h, C.byref(nq), None, None, None), "bm_reads_segments")
        n = self.n_seg = nq.value               # the host-buffer entry points (bm_map_locate*) leave a batch resident too
        a, b, c = (np.empty(n, np.uint32) for _ in range(3))
        nn = C.c_uint32()
        self._ck(self.lib.bm_reads_segments(self.h, C.byref(nn), _ptr(a), _ptr(b), _ptr(c)), "bm_reads_segments")
        return a, b, c

    # ---- mapper --------------------------------------------------------------------------------------------------
    def map_device(self, P):
        self._ck(self.lib.bm_map_device(self.h, C.byref(P)), "bm_map_device")

    def map_download(self, max_candidates=30):
        off = np.empty(2 * self.n_seg + 1, np.uint32)
        stats = np.zeros(8, np.uint64)
        self._ck(self.lib.bm_map_download(self.h, _ptr(off), None, 0, _ptr(stats)), "bm_map_download")
        cand = np.empty(int(off[-1]), np.uint32)
        self._ck(self.lib.bm_map_download(self.h, None, _ptr(cand), len(cand), None), "bm_map_download")
        return off, cand, stats

    def map_reads(self, P, seq, qual, read_off):
        """host buffers in, host candidate lists out (q_gram_mapper::map for one batch)."""
        n = len(read_off) - 1
        nseg_max = n * max(1, P.num_segment_samples)
        off = np.empty(2 * nseg_max + 1, np.uint32)
        cand = np.empty(2 * nseg_max * P.max_candidates + 1, np.uint32)
        stats = np.zeros(8, np.uint64)
        self._ck(self.lib.bm_map_reads(self.h, C.byref(P), _ptr(seq), _ptr(qual), _ptr(read_off), n, _ptr(off),
                                       _ptr(cand), len(cand), _ptr(stats)), "bm_map_reads")
        self.n_reads = n
        ns = C.c_uint32()
        self.lib.bm_reads_segments(self.h, C.byref(ns), None, None, None)
        self.n_seg = ns.value
        off = off[:2 * self.n_seg + 1]
        return off, cand[:int(off[-1])].copy(), stats

    # ---- locator -------------------------------------------------------------------------------------------------
    def locate_device(self, P, filter_best):
        self._ck(self.lib.bm_locate_device(self.h, C.byref(P), int(filter_best)), "bm_locate_device")

    def locate_download(self):
        n = C.c_uint64()
        self._ck(self.lib.bm_locate_count(self.h, C.byref(n)), "bm_locate_count")
        off = np.empty(self.n_reads + 1, np.uint32)
        locs = np.empty(n.value, LOC_DTYPE)
        self._ck(self.lib.bm_locate_download(self.h, _ptr(off), _ptr(locs), n.value), "bm_locate_download")
        return off, locs

    def locate_download_into(self, off, locs):
        """bm_locate_download into the caller's arrays (pinned ones make the copy asynchronous-speed); returns n_locs"""
        n = C.c_uint64()
        self._ck(self.lib.bm_locate_count(self.h, C.byref(n)), "bm_locate_count")
        self._ck(self.lib.bm_locate_download(self.h, _ptr(off), _ptr(locs), len(locs)), "bm_locate_download")
        return n.value

    def align_download_into(self, aln, cig):
        """bm_align_download into the caller's arrays; returns the number of CIGAR ops"""
        used = C.c_uint64()
        self._ck(self.lib.bm_align_download(self.h, _ptr(aln), _ptr(cig), len(cig), C.byref(used)), "bm_align_download")
        return used.value

    def locate_device_buffers(self):
        """(loc_off device pointer, locs device pointer, n_locs) of the last locate_device"""
        a, b, n = _vp(), _vp(), C.c_uint64()
        self._ck(self.lib.bm_locate_device_buffers(self.h, C.byref(a), C.byref(b), C.byref(n)),
                 "bm_locate_device_buffers")
        return a.value, b.value, n.value

    def index_alloc(self, k2i, q, n_rows, NB):
        k2i = np.ascontiguousarray(k2i, np.int32)
        self._ck(self.lib.bm_index_alloc(self.h, _ptr(k2i), q, n_rows, NB), "bm_index_alloc")
        self.NB, self.n_rows = NB, n_rows

    def map_locate(self, P, seq, qual, read_off, filter_best=True):
        n = len(read_off) - 1
        nseg_max = n * max(1, P.num_segment_samples)
        cap = 2 * nseg_max * P.max_candidates + 1
        off = np.empty(n + 1, np.uint32)
        locs = np.empty(cap, LOC_DTYPE)
        stats = np.zeros(8, np.uint64)
        self._ck(self.lib.bm_map_locate(self.h, C.byref(P), _ptr(seq), _ptr(qual), _ptr(read_off), n,
                                        int(filter_best), _ptr(off), _ptr(locs), cap, _ptr(stats)), "bm_map_locate")
        self.n_reads = n
        return off, locs[:int(off[-1])].copy(), stats

    def map_locate_into(self, P, seq, qual, read_off, out_off, out_locs, filter_best=True):
        """bm_map_locate with caller-provided (ideally pinned) output arrays: out_off uint32[n+1], out_locs LOC_DTYPE[cap].
        Returns (n_locs, stats)."""
        n = len(read_off) - 1
        stats = np.zeros(8, np.uint64)
        self._ck(self.lib.bm_map_locate(self.h, C.byref(P), _ptr(seq), _ptr(qual), _ptr(read_off), n,
                                        int(filter_best), _ptr(out_off), _ptr(out_locs), len(out_locs), _ptr(stats)),
                 "bm_map_locate")
        self.n_reads = n
        return int(out_off[n]), stats

    def map_locate_packed_into(self, P, planes, qk, pl_off, read_off, out_off, out_locs, filter_best=True):
        """bm_map_locate_packed with caller-provided (ideally pinned) arrays.  Returns (n_locs, stats)."""
        n = len(read_off) - 1
        stats = np.zeros(8, np.uint64)
        self._ck(self.lib.bm_map_locate_packed(self.h, C.byref(P), _ptr(planes), _ptr(qk), _ptr(pl_off), _ptr(read_off), n,
                                               int(filter_best), _ptr(out_off), _ptr(out_locs),
                                               len(out_locs) if out_locs is not None else 0, _ptr(stats)),
                 "bm_map_locate_packed")
        self.n_reads = n
        return (int(out_off[n]) if out_off is not None else -1), stats

    # ---- several GPUs ----------------------------------------------------------------------------------------------
    def comm_unique_id(self):
        buf = (C.c_uint8 * COMM_ID_BYTES)()
        rc = self.lib.bm_comm_unique_id(buf)
        if rc:
            raise BucketMapError(f"bm_comm_unique_id failed ({rc}): libnccl.so.2 not loadable?")
        return bytes(buf)

    def comm_init(self, uid, nranks, rank):
        buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(uid)
        self._ck(self.lib.bm_comm_init(self.h, buf, nranks, rank), "bm_comm_init")

    def bcast_index(self, root=0):
        self._ck(self.lib.bm_bcast_index(self.h, root), "bm_bcast_index")

    def bcast_ref(self, root=0):
        self._ck(self.lib.bm_bcast_ref(self.h, root), "bm_bcast_ref")

    def gather_open(self, rank, nranks, root, max_reads, max_locs):
        buf = (C.c_uint8 * PEER_HANDLE_BYTES)()
        self._ck(self.lib.bm_gather_open(self.h, rank, nranks, root, max_reads, max_locs, buf), "bm_gather_open")
        return bytes(buf)

    def gather_connect(self, handles):
        blob = b"".join(handles)
        buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        self._ck(self.lib.bm_gather_connect(self.h, buf), "bm_gather_connect")

    def gather_results(self, out_off=None, out_locs=None):
        """every rank after locate_device; root with arrays: blocks and returns (n_reads_total, n_locs_total)"""
        nr, nl = C.c_uint64(), C.c_uint64()
        want = out_off is not None
        self._ck(self.lib.bm_gather_results(self.h, _ptr(out_off), _ptr(out_locs), len(out_locs) if out_locs is not None else 0,
                                            C.byref(nr) if want else None, C.byref(nl) if want else None),
                 "bm_gather_results")
        return nr.value, nl.value

    def gather_sync(self):
        self._ck(self.lib.bm_gather_sync(self.h), "bm_gather_sync")

    def gather_host_async(self, on):
        self._ck(self.lib.bm_gather_host_async(self.h, int(on)), "bm_gather_host_async")

    def gather_join(self):
        self._ck(self.lib.bm_gather_join(self.h), "bm_gather_join")

    def gather_download(self):
        """root: (loc_off, locs) of the last gathered batch, copied from the root's device buffers"""
        nr, nl = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.bm_gather_download(self.h, None, None, 0, C.byref(nr), C.byref(nl)), "bm_gather_download")
        off = np.empty(nr.value + 1, np.uint32)
        locs = np.empty(nl.value, LOC_DTYPE)
        self._ck(self.lib.bm_gather_download(self.h, _ptr(off), _ptr(locs), len(locs), C.byref(nr), C.byref(nl)),
                 "bm_gather_download")
        return off, locs

    def gather_close(self):
        self._ck(self.lib.bm_gather_close(self.h), "bm_gather_close")

    # ---- alignment -----------------------------------------------------------------------------------------------
    def align_device(self, P):
        self._ck(self.lib.bm_align_device(self.h, C.byref(P)), "bm_align_device")

    def align_download(self, n_locs):
        aln = np.empty(n_locs, ALN_DTYPE)
        used = C.c_uint64()
        self._ck(self.lib.bm_align_download(self.h, _ptr(aln), None, 0, C.byref(used)), "bm_align_download")
        cig = np.empty(max(1, used.value), np.uint32)
        self._ck(self.lib.bm_align_download(self.h, _ptr(aln), _ptr(cig), len(cig), C.byref(used)),
                 "bm_align_download")
        return aln, cig[:used.value]

    # ---- measurement ---------------------------------------------------------------------------------------------
    def timer_start(self):
        self._ck(self.lib.bm_timer_start(self.h), "bm_timer_start")

    def timer_stop(self):
        ms = C.c_float()
        self._ck(self.lib.bm_timer_stop(self.h, C.byref(ms)), "bm_timer_stop")
        return ms.value

    def kernel_timing(self, on):
        self._ck(self.lib.bm_kernel_timing(self.h, int(on)), "bm_kernel_timing")

    def kernel_stats(self):
        out = {}
        for kid, name in enumerate(KERNEL_NAMES):
            ms, n = C.c_double(), C.c_uint64()
            self._ck(self.lib.bm_kernel_stats(self.h, kid, C.byref(ms), C.byref(n)), "bm_kernel_stats")
            out[name] = (ms.value, n.value)
        return out

    def kernel_stats_reset(self):
        self._ck(self.lib.bm_kernel_stats_reset(self.h), "bm_kernel_stats_reset")

    def flush_l2(self, nbytes=256 << 20):
        self._ck(self.lib.bm_flush_l2(self.h, nbytes), "bm_flush_l2")

    def sync(self):
        self._ck(self.lib.bm_sync(self.h), "bm_sync")

    def pinned(self, shape, dtype):
        """numpy array over cudaMallocHost memory (freed with the context object's lifetime)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = self.lib.bm_host_alloc(max(n, 1))
        if not p:
            raise BucketMapError("bm_host_alloc failed")
        buf = (C.c_uint8 * max(n, 1)).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        self._keep.append(p)
        return arr
