"""GPU: bm_align_noalloc -- the verifier for callers that cannot allocate (static device scratch, caller's device pointers)
-- against bm_align_device + bm_align_download on the same locations, bit for bit, on two calls in a row.

The file sorts last on purpose: the entry point has no counterpart in the reference and a failure here must not hide the
parity tests.  torch only provides the caller's device buffers."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()


def _reference_results(oracle, bm, ctx, d, P):
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    ctx.reads_upload(P, d.seq, d.qual, d.read_off)
    ctx.map_device(P)
    ctx.locate_device(P, False)
    loff, locs = ctx.locate_download()
    ctx.align_device(P)
    aln, cig = ctx.align_download(len(locs))
    return loff, locs, aln, cig


def _call(torch, bm, ctx, d, P, loff, locs, max_len):
    """one bm_align_noalloc call on reads [0, n) whose locations are locs[:loff[n]]; returns host copies of the outputs"""
    n_reads = len(loff) - 1
    n_aln = int(loff[-1])
    read_off = np.ascontiguousarray(d.read_off[:n_reads + 1], np.uint64)
    planes, _, pl_off, _ = bm.pack_reads(ctx.lib, d.seq, d.qual, read_off, P.k, P.avg_base_quality)
    ins = [_dev(torch, x) for x in (d.packed, d.b_start.astype(np.uint64), d.b_len.astype(np.uint32), planes, pl_off,
                                    read_off, loff.astype(np.uint32), locs[:n_aln])]
    o_aln = torch.full((max(n_aln, 1) * bm.ALN_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device="cuda")
    o_cig = torch.full((max(n_aln, 1) * 16 * 4,), 0xCD, dtype=torch.uint8, device="cuda")
    o_cnt = torch.full((max(n_aln, 1) * 4,), 0xEF, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info()[0]
    ctx.align_noalloc(P, ins[0].data_ptr(), ins[1].data_ptr(), ins[2].data_ptr(), ins[3].data_ptr(), ins[4].data_ptr(),
                      ins[5].data_ptr(), n_reads, max_len, ins[6].data_ptr(), ins[7].data_ptr(), n_aln,
                      o_aln.data_ptr(), o_cig.data_ptr(), o_cnt.data_ptr())
    ctx.sync()
    assert torch.cuda.memory_allocated() == before
    aln = o_aln.cpu().numpy().view(bm.ALN_DTYPE)[:n_aln]
    cig = o_cig.cpu().numpy().view(np.uint32).reshape(-1, 16)[:n_aln]
    cnt = o_cnt.cpu().numpy().view(np.uint32)[:n_aln]
    return aln, cig, cnt, free_before - torch.cuda.mem_get_info()[0]


def _compare(aln, cig, cnt, want_aln, want_cig):
    n = len(aln)
    for f in ("score", "begin", "map_qual", "kept", "cigar_len"):
        assert np.array_equal(aln[f], want_aln[f][:n]), f
    assert np.array_equal(cnt, want_aln["cigar_len"][:n])
    assert np.array_equal(aln["cigar_off"], np.arange(n, dtype=np.uint32) * 16)
    short = 0
    for i in range(n):
        ln = int(want_aln["cigar_len"][i])
        if ln <= 16:
            o = int(want_aln["cigar_off"][i])
            assert np.array_equal(cig[i, :ln], want_cig[o:o + ln]), i
            short += 1
    return short


@pytest.mark.parametrize("kw,pk", [
    (dict(read_len=300, n_reads=900, sub=0.01, indel=0.001), dict()),          # C5-like, NW = 5
    (dict(read_len=150, n_reads=3000), dict()),                                # NW = 3
    (dict(read_len=500, n_reads=350, sub=0.01, indel=0.002), dict()),          # NW = 8
])
def test_noalloc_verifier_equals_bm_align_device_twice_in_a_row(oracle, bm, ctx, kw, pk):
    torch = pytest.importorskip("torch")
    d = dataset(**kw)
    P = bm.default_params(ctx.lib, d.L, **pk)
    loff, locs, want_aln, want_cig = _reference_results(oracle, bm, ctx, d, P)
    assert len(locs) > 0.9 * kw["n_reads"]
    max_len = int(np.diff(d.read_off).max())
    # call 1: all reads; call 2 straight after it: the first half only (the array still holds call 1's records)
    a1, c1, n1, grew1 = _call(torch, bm, ctx, d, P, loff, locs, max_len)
    short = _compare(a1, c1, n1, want_aln, want_cig)
    assert short > 0.9 * len(a1)
    half = (len(loff) - 1) // 2
    a2, c2, n2, grew2 = _call(torch, bm, ctx, d, P, loff[:half + 1], locs, max_len)
    _compare(a2, c2, n2, want_aln, want_cig)
    assert grew2 <= 0                      # no device memory left the free pool during the second call


def test_noalloc_verifier_refuses_more_than_the_array_holds(oracle, bm, ctx):
    torch = pytest.importorskip("torch")
    d = dataset(read_len=300, n_reads=1400)
    P = bm.default_params(ctx.lib, d.L)
    loff, locs, want_aln, want_cig = _reference_results(oracle, bm, ctx, d, P)
    assert int(loff[-1]) > 1024
    max_len = int(np.diff(d.read_off).max())
    with pytest.raises(bm.BucketMapError, match="at most 1024 alignments per call"):
        _call(torch, bm, ctx, d, P, loff, locs, max_len)
    # ... and the batch split in two goes through
    cut = int(np.searchsorted(loff, 1000, side="right")) - 1
    a1, c1, n1, _ = _call(torch, bm, ctx, d, P, loff[:cut + 1], locs, max_len)
    _compare(a1, c1, n1, want_aln, want_cig)
    with pytest.raises(bm.BucketMapError, match="null argument"):
        ctx.align_noalloc(P, 0, 0, 0, 0, 0, 0, 1, 300, 0, 0, 1, 0, 0, 0)
