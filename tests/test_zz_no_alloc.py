"""pfb_kat_trigger_static (include/pfbwt_b200.h): the trigger known-answer kernel with its scratch in a static device
array instead of per-call device allocations.  No counterpart in the reference.
The file is named so that pytest collects it last.

GPU part: through the C ABI, the static entry point returns bit for bit what pfb_kat_trigger returns (integer results:
the tolerance is zero, as in the other parity tests), on two calls in a row that share the array and differ in size, and
rejects n above PFB_KAT_STATIC_MAX.  CPU part: the symbol is exported and the limit in the header is the one the test uses."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pfbwt_b200_loader import binding  # noqa: E402


def header_limit():
    h = open(os.path.join(ROOT, "include", "pfbwt_b200.h")).read()
    return int(re.search(r"#define\s+PFB_KAT_STATIC_MAX\s+(\d+)", h).group(1))


def test_static_entry_point_is_declared_and_exported():
    import ctypes
    B = binding()
    assert header_limit() == 65536
    lib = ctypes.CDLL(B.LIB_PATH)
    assert hasattr(lib, "pfb_kat_trigger_static")
    assert "pfb_kat_trigger_static" in B.EXPORTS


@pytest.mark.gpu
def test_static_scratch_matches_allocating_entry_point_twice():
    B = binding()
    MAXN = header_limit()
    ctx = B.Context(device=0, bits=32, w=10, p=100)
    try:
        rng = np.random.default_rng(20261105)
        # call 1 fills the whole array; call 2 is shorter and has other k-mers and another modulus: stale bytes of call 1
        # beyond its n (and below it) must not show up in its results
        first = np.concatenate([np.array([0, 1, (1 << 20) - 1, (1 << 64) - 1], dtype=np.uint64),
                                rng.integers(0, 1 << 63, MAXN - 4, dtype=np.uint64)])
        second = rng.integers(0, 1 << 20, 12_345, dtype=np.uint64)
        for kmers, p in ((first, 100), (second, 7), (first[:1], 1), (second, 100)):
            h_ref, d_ref = ctx.kat_trigger(kmers, p)
            h, d = ctx.kat_trigger_static(kmers, p)
            assert np.array_equal(h, h_ref), f"hashes differ (n={kmers.shape[0]}, p={p})"
            assert np.array_equal(d, d_ref), f"divisibility flags differ (n={kmers.shape[0]}, p={p})"
            assert np.array_equal(d.astype(bool), (h % np.uint64(p)) == 0)
        assert int(ctx.kat_trigger_static(first[:1], 100)[0][0]) == 0x77cfa1eef01bca90     # wang_hash(0), include/hash.hpp:12-21
        # n = 0 is a no-op, n beyond the array is refused with a message that names the limit
        h, d = ctx.kat_trigger_static(np.empty(0, dtype=np.uint64), 100)
        assert h.shape[0] == 0 and d.shape[0] == 0
        with pytest.raises(B.PfbError) as ei:
            ctx.kat_trigger_static(np.zeros(MAXN + 1, dtype=np.uint64), 100)
        assert "PFB_KAT_STATIC_MAX" in str(ei.value) and str(MAXN) in str(ei.value)
        # the context still works after the refusal
        h, d = ctx.kat_trigger_static(second, 100)
        h_ref, d_ref = ctx.kat_trigger(second, 100)
        assert np.array_equal(h, h_ref) and np.array_equal(d, d_ref)
    finally:
        ctx.close()
