"""Host-side integer logic of the path that is not a kernel (product code; mirrors, not imports, the oracle).

  fracminhash   tools/hash_function_generator.h:95-117 + indexer/bucket_indexer.h:147-159
  num_buckets   CMakeLists.txt:13-50 (awk): NB = sum ceil(len/L)
  bucketize     utils.h:60-102
"""
import numpy as np

_PRIMES = [5, 11, 23, 47, 97, 199, 409, 823, 1741, 3469, 6949, 14033, 28411, 57557, 116731, 236897, 480881, 976369,
           1982627, 4026031, 8175383, 16601593, 33712729, 68460391, 139022417, 282312799, 573292817, 1164186217,
           2364114217, 4294967291]
HASH_TABLE_SIZE = 10000   # main.cpp:176


def fracminhash(q, x, y, frac):
    """kmer_to_index[i] = running row id if ((x*i + y) % p % 10000) <= (unsigned)(10000*frac) else -1"""
    p = next(c for c in _PRIMES if c > 10 * HASH_TABLE_SIZE)
    thr = int(np.float32(HASH_TABLE_SIZE) * np.float32(frac))
    i = np.arange(1 << (2 * q), dtype=np.uint64)
    h = (np.uint64(x) * i + np.uint64(y)) % np.uint64(p) % np.uint64(HASH_TABLE_SIZE)
    keep = h <= thr
    k2i = np.full(len(i), -1, np.int32)
    k2i[keep] = np.arange(int(keep.sum()), dtype=np.int32)
    return k2i, int(keep.sum())


def num_buckets(rec_len, L):
    return int(sum((int(n) + L - 1) // L for n in rec_len if int(n) > 0))
