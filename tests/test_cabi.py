"""CPU: the C-ABI library loads, exports every symbol include/bucketmap_b200.h declares, and fails loudly (no CPU
fallback) when there is no CUDA device.  No compute calls."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "bucketmap_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(bm_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_the_path():
    names = declared_functions()
    for must in ("bm_ctx_create", "bm_ref_upload", "bm_index_build", "bm_index_load", "bm_map_reads", "bm_map_locate",
                 "bm_locate_device", "bm_align_device", "bm_kernel_stats"):
        assert must in names


def test_library_exports_every_declared_symbol(bm):
    lib = C.CDLL(bm.LIB_PATH)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_python_binding_covers_every_declared_symbol(bm):
    assert sorted(bm.SIGNATURES) == declared_functions()
    bm.load_library()          # attaches prototypes; raises on a stale library


def test_struct_layouts_match_header(bm):
    # bm_params: 15 4-byte fields; bm_loc: 5; bm_aln: 6 (the numpy dtypes the tests read results through)
    assert C.sizeof(bm.Params) == 15 * 4
    assert bm.LOC_DTYPE.itemsize == 20 and bm.ALN_DTYPE.itemsize == 24


def test_params_derive_matches_oracle(bm, oracle):
    lib = bm.load_library()
    for kw in (dict(), dict(s=20, e=0.6, k=14, b=10), dict(r=150), dict(s=30, e=0.35, p=15, n=0.05, u=30),
               dict(r=100, n=0.03, e=0.45), dict(d=0.7, q=8, k=11)):
        a = bm.default_params(lib, 65536, **kw)
        b = oracle.params(65536, **kw)
        for name, _ in bm.Params._fields_:
            assert getattr(a, name) == getattr(b, name), (kw, name)
    d = bm.default_params(lib, 65536)
    # the float-derived thresholds of SURVEY section 0
    assert (d.mapper_fault, d.allowed_mismatch, d.allowed_indel) == (6, 4, 6)
    e = bm.default_params(lib, 65536, s=20, e=0.6, r=150)
    assert (e.mapper_fault, e.allowed_indel) == (12, 3)


def test_no_cpu_fallback(bm):
    """without a CUDA device the product refuses to construct a context (and never computes on the CPU)"""
    lib = bm.load_library()
    h = C.c_void_p()
    rc = lib.bm_ctx_create(0, C.byref(h))
    if rc == 0:
        lib.bm_ctx_destroy(h)
        pytest.skip("a CUDA device is present")
    assert rc == -2 and not h.value
    with pytest.raises(bm.BucketMapError):
        bm.Context(0)


def test_missing_library_is_an_error(bm, tmp_path):
    with pytest.raises(bm.BucketMapError):
        bm.load_library(str(tmp_path / "libbucketmap_b200.so"))


def test_product_does_not_reference_the_oracle():
    """nothing under bucket-map_b200/ may include, import or link oracle/"""
    pkg = os.path.join(ROOT, "bucket-map_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp", ".py")):
                text = open(os.path.join(dp, f), errors="ignore").read()
                for pat in (r"#\s*include[^\n]*oracle", r"^\s*(import|from)\s+[^\n]*oracle", r"libbm_oracle",
                            r"dlopen[^\n]*oracle", r"sys\.path[^\n]*oracle"):
                    assert not re.search(pat, text, re.M), (os.path.join(dp, f), pat)
