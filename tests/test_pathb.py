"""CPU: INTEGRATION.md path B is more than prose -- the substitutions it lists apply to the reference's main.cpp as it is,
and the result compiles and links against the product's host classes and the C-ABI library (no GPU needed to build)."""
import os

import pytest

import pathb


@pytest.mark.skipif(not pathb.can_build(), reason="the reference sources are only present in the build container")
def test_patch_applies_and_builds():
    src = pathb.patched_source()
    assert '#include "bm_host.hpp"' in src and "bucket_hash_indexer<BM_BUCKET_NUM>" not in src
    assert "bm::q_gram_mapper map(" in src and "map.set_num_buckets(BM_BUCKET_NUM);" in src
    assert "bm::bucket_locator loc(&ind, &map," in src
    assert "hash_function_generator gen;" in src               # everything else is the reference's own text
    for outs in pathb.build_all():
        for o in outs:
            assert os.path.exists(o) and os.access(o, os.X_OK)
