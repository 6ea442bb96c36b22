"""INTEGRATION.md path B, compiled: the reference's OWN main.cpp with the three-class swap applied, built against the
product's host classes (bucket-map_b200/host/bm_host.hpp) and libbucketmap_b200.so.  TEST INFRASTRUCTURE ONLY.

The patch is applied to a copy of /root/reference/bucket_map/main.cpp (read where it lies, never committed; the patched
copy and the binaries go to the git-ignored oracle/_ref/).  seqan3::debug_stream and sharg come from oracle/shim, as for
the unmodified reference build.  Needs /root/reference, i.e. this container; the binaries travel to the GPU box, where
tests/test_gpu_pathb.py runs them on the goldens."""
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_MAIN = "/root/reference/bucket_map/main.cpp"
OUT = os.path.join(HERE, "_ref")
CONFIGS = [(49, 8192)]          # (NB, L) of the goldens the GPU test runs


def binary(NB, L, align=False):
    return os.path.join(OUT, f"pathb_{'bucketmapalign' if align else 'bucketmap'}_{NB}_{L}")


def available(NB, L, align=False):
    return os.path.exists(binary(NB, L, align))


def can_build():
    return os.path.exists(REF_MAIN)


def patched_source():
    """the reference main.cpp after the substitutions INTEGRATION.md lists"""
    s = open(REF_MAIN).read()
    n = 0

    def sub(pat, rep, flags=0, count=1):
        nonlocal s, n
        s, k = re.subn(pat, rep, s, count=count, flags=flags)
        if k == 0:
            raise RuntimeError("INTEGRATION.md path B: pattern not found in the reference main.cpp: " + pat)
        n += k

    sub(r'#include "indexer/bucket_hash_indexer\.h"\n', '#include "bm_host.hpp"                       // bucket-map_b200/host\n')
    sub(r'#include "mapper/q_gram_mapper\.h"\n', "")
    sub(r'#include "locator/bucket_locator\.h"\n', "")
    sub(r"bucket_hash_indexer<BM_BUCKET_NUM>\s+ind\(", "bm::bucket_hash_indexer ind(")
    sub(r"q_gram_mapper<BM_BUCKET_NUM>\s+map\((.*?)\);", r"bm::q_gram_mapper map(\1);\n    map.set_num_buckets(BM_BUCKET_NUM);      // was the template argument",
        flags=re.S)
    sub(r"bucket_locator\s+loc\(&ind,\s*&map,(.*?)args\.average_base_quality\);",
        r"bm::bucket_locator loc(&ind, &map,\1args.average_base_quality, 5, BM_PATHB_ALIGN);", flags=re.S)
    return s


def build(NB, L):
    if not can_build():
        raise RuntimeError("the reference sources (/root/reference) are not present on this machine")
    os.makedirs(OUT, exist_ok=True)
    src = os.path.join(OUT, "pathb_main.cpp")
    text = patched_source()
    if not os.path.exists(src) or open(src).read() != text:
        open(src, "w").write(text)
    host = os.path.join(ROOT, "bucket-map_b200", "host")
    lib_dir = os.path.join(ROOT, "bucket-map_b200")
    deps = [src, os.path.join(host, "bm_host.hpp"), os.path.join(lib_dir, "libbucketmap_b200.so")] + \
           [os.path.join(host, f) for f in ("bm_classes.cpp", "bm_io.cpp", "bm_locator.cpp")]
    outs = []
    for align in (False, True):
        out = binary(NB, L, align)
        outs.append(out)
        if os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
            continue
        cmd = ["g++", "-std=c++20", "-O2", "-w", "-I", os.path.join(HERE, "shim"), "-I", "/root/reference/bucket_map",
               "-I", os.path.join(ROOT, "include"), "-I", host,
               f"-DBM_BUCKET_NUM={NB}", f"-DBM_BUCKET_LEN={L}", '-DBM_GENOME_PATH=std::getenv("BM_GENOME")',
               f"-DBM_PATHB_ALIGN={'true' if align else 'false'}"] + (["-DBM_ALIGN"] if align else []) + \
              ["-o", out, src] + deps[3:] + ["-x", "c", os.path.join(HERE, "ref_rand.c"), "-x", "none",
               "-L", lib_dir, "-lbucketmap_b200", "-Wl,-rpath," + lib_dir, "-Wl,-rpath,$ORIGIN/../../bucket-map_b200", "-pthread"]
        subprocess.check_call(cmd)
    return outs


def build_all():
    return [build(NB, L) for NB, L in CONFIGS]


if __name__ == "__main__":
    print(build_all())
