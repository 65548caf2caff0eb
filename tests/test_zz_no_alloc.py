"""psift_match_dev_noalloc (include/psift.h): the matcher with its device scratch in a static __device__ array instead
of cudaMalloc'd buffers.  No counterpart in the reference.  Checked through
the C ABI against the existing entry point psift_match_dev on two calls in a row (the array is shared by all calls and
keeps the last call's contents).  Collected last on purpose."""
import ctypes as C
import os
import numpy as np
import pytest
import refio
from popsift_b200 import capi

pytestmark = pytest.mark.gpu
NOALLOC_MAX = 65536          # PSIFT_NOALLOC_MATCH_MAX


def _desc(case):
    return np.ascontiguousarray(refio.RefFeatures(os.path.join(refio.GOLD, "ref", case + ".feat.bin")).desc, np.float32)


def _match(fn, l, r):
    m = np.full((l.n_desc, 3), -7, np.int32); d = np.full((l.n_desc, 2), -7.0, np.float32)
    rc = fn(0, l.info.desc, l.n_desc, r.info.desc, r.n_desc, m.ctypes.data, d.ctypes.data)
    assert rc == 0, capi.lib().psift_last_error().decode()
    return m, d


def test_noalloc_matcher_equals_allocating_matcher_twice():
    L = capi.lib()
    old, new = L.psift_match_dev, L.psift_match_dev_noalloc
    big, small = _desc("g640_default"), _desc("g256_default")
    with capi.DevFeatures.from_host(big) as a, capi.DevFeatures.from_host(small) as b:
        # call 1: many left rows; call 2: fewer left rows right after it (stale rows of call 1 stay behind in the array)
        for l, r in ((a, b), (b, a), (a, a)):
            want_m, want_d = _match(old, l, r)
            got_m, got_d = _match(new, l, r)
            assert np.array_equal(got_m, want_m)                       # best, second best, accept: integers, bit-exact
            assert np.array_equal(got_d.view(np.uint32), want_d.view(np.uint32))   # squared distances: same kernel, same bits
            assert (got_m[:, 2] >= 0).all() and (got_m[:, 0] >= 0).all()
        # an empty right-hand set and an empty left-hand set behave like the allocating entry point
        with capi.DevFeatures.from_host(np.zeros((0, 128), np.float32)) as e:
            want_m, _ = _match(old, b, e); got_m, _ = _match(new, b, e)
            assert np.array_equal(got_m, want_m)
            assert new(0, e.info.desc, 0, b.info.desc, b.n_desc, None, None) == 0


def test_noalloc_matcher_rejects_sizes_beyond_the_static_scratch():
    L = capi.lib()
    d = _desc("g256_default")
    with capi.DevFeatures.from_host(d) as b:
        m = np.zeros((4, 3), np.int32)
        rc = L.psift_match_dev_noalloc(0, b.info.desc, NOALLOC_MAX + 1, b.info.desc, b.n_desc, m.ctypes.data, None)
        assert rc != 0
        msg = L.psift_last_error().decode()
        assert "PSIFT_NOALLOC_MATCH_MAX" in msg and str(NOALLOC_MAX) in msg
        # and the entry point still works afterwards
        got_m, _ = _match(L.psift_match_dev_noalloc, b, b)
        want_m, _ = _match(L.psift_match_dev, b, b)
        assert np.array_equal(got_m, want_m)
