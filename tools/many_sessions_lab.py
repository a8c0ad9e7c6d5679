#!/usr/bin/env python3
"""Ragged Merkle trees over MANY small sessions (10^6 x 10, 10^5 x 100, 4 x 10^6 x 3 leaves; digests resident on the
device): time per ahv_merkle_roots_dev call.  The regime where per-session work is tiny and anything serial over the
session list shows (round 2: 8.5 / 2.6 / 27.7 ms before the tiled row-pointer scan and the thread-per-session finish,
2.1 / 2.0 / 2.3 ms after)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from agent_hypervisor_b200 import Context
ctx = Context(0)
for S, L in ((1_000_000, 10), (100_000, 100), (4_000_000, 3)):
    n = S * L
    digs = np.random.default_rng(1).integers(0, 256, 32 * n, dtype=np.uint8)
    rp = (np.arange(S + 1, dtype=np.uint64) * L)
    d_digs = ctx.dev_alloc(32 * n); d_rp = ctx.dev_alloc(rp.nbytes); d_roots = ctx.dev_alloc(32 * S); d_empty = ctx.dev_alloc(S)
    ctx.h2d(d_digs, digs); ctx.h2d(d_rp, rp); ctx.sync()
    best = 1e9
    for rep in range(4):
        ctx.timer_start(1)
        ctx.merkle_roots_dev(d_digs, d_rp, S, n, L, d_roots, d_empty)
        ctx.timer_stop(1); ctx.sync(); best = min(best, ctx.timer_ms(1))
    print(f"S={S} L={L} ragged tree: {best:.3f} ms ({n} leaves)", flush=True)
    for p in (d_digs, d_rp, d_roots, d_empty): ctx.dev_free(p)
