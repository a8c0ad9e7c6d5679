"""The allocation-free Merkle entry point (ahv_merkle_roots_noalloc_dev) against the regular one.

Some users of the library may not allocate device memory once they are running - their GPU memory
was divided up at start-up between this library and another framework, or they sit in a loop that
cannot tolerate the device-wide stall of cudaMalloc/cudaFree.  For them the library keeps a fixed
8 MiB scratch array in its own image.  This file checks, through the C ABI, that the variant gives
bit-identical roots to ahv_merkle_roots_dev on two calls in a row (the second call finds the
first call's leftovers in the shared array), and that an oversized batch is refused up front.

Named test_zz_* so that pytest collects it last: a failure here cannot hide any other test.
"""

import numpy as np
import pytest

from oracle import coracle as co

pytestmark = pytest.mark.gpu


def _roots(ctx, digs, rp, noalloc):
    S, n = len(rp) - 1, len(digs)
    maxl = int(np.diff(rp.astype(np.int64)).max())
    d_dig, d_rp = ctx.dev_alloc(max(digs.nbytes, 32)), ctx.dev_alloc(rp.nbytes)
    d_roots, d_empty = ctx.dev_alloc(32 * S), ctx.dev_alloc(S)
    try:
        ctx.h2d(d_dig, digs); ctx.h2d(d_rp, rp)
        # poison the outputs so that a call that wrote nothing cannot pass
        ctx.memset(d_roots, 0xA5, 32 * S); ctx.memset(d_empty, 0xA5, S)
        fn = ctx.merkle_roots_noalloc_dev if noalloc else ctx.merkle_roots_dev
        fn(d_dig, d_rp, S, n, maxl, d_roots, d_empty)
        roots, empty = np.empty((S, 32), np.uint8), np.empty(S, np.uint8)
        ctx.d2h(roots, d_roots); ctx.d2h(empty, d_empty)
        ctx.sync()
        return roots, empty
    finally:
        for p in (d_dig, d_rp, d_roots, d_empty):
            ctx.dev_free(p)


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    from agent_hypervisor_b200 import get_context
    ctx = get_context()
    rng = np.random.default_rng(2026)
    # call 1: ragged sessions, several of them deep enough to use the global level buffers
    sizes1 = [0, 1, 2, 3, 5, 17, 255, 256, 257, 700, 1000, 1025, 4099, 20000, 1, 0, 333]
    # call 2: a different, smaller shape: it runs on top of what call 1 left in the shared array
    sizes2 = [1000] * 40 + [7, 0, 513]
    for sizes in (sizes1, sizes2):
        rp = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
        digs = rng.integers(0, 256, (int(rp[-1]), 32), dtype=np.uint8)
        want_roots, want_empty = _roots(ctx, digs, rp, noalloc=False)
        got_roots, got_empty = _roots(ctx, digs, rp, noalloc=True)
        assert np.array_equal(got_empty, want_empty)
        assert np.array_equal(got_roots, want_roots)        # bit-exact, like every digest test here
        oracle_roots, oracle_empty = co.merkle_roots(digs, rp)             # ... and both against the oracle, not only each other
        assert np.array_equal(got_roots, oracle_roots) and np.array_equal(got_empty, oracle_empty)
        assert list(got_empty) == [1 if s == 0 else 0 for s in sizes]


def test_noalloc_refuses_oversized_batch_and_missing_max():
    from agent_hypervisor_b200 import NativeError, get_context
    ctx = get_context()
    n = 400_000                                   # 2 x 32 x n/2 = 12.8 MB of level buffers > 8 MiB
    rp = np.array([0, n], dtype=np.uint64)
    d_dig, d_rp, d_roots, d_empty = ctx.dev_alloc(32 * n), ctx.dev_alloc(rp.nbytes), ctx.dev_alloc(32), ctx.dev_alloc(1)
    try:
        ctx.h2d(d_rp, rp); ctx.sync()
        with pytest.raises(NativeError, match="fixed 8388608-byte array"):
            ctx.merkle_roots_noalloc_dev(d_dig, d_rp, 1, n, n, d_roots, d_empty)
        with pytest.raises(NativeError, match="max_session_leaves"):
            ctx.merkle_roots_noalloc_dev(d_dig, d_rp, 1, n, 0, d_roots, d_empty)
    finally:
        for p in (d_dig, d_rp, d_roots, d_empty):
            ctx.dev_free(p)
