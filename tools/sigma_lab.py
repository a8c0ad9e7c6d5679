#!/usr/bin/env python3
"""Latency of single sigma_eff tiles (development tool)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
def run(name, deg):
    deg = np.asarray(deg, dtype=np.int64); A = len(deg)
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint32); E = int(rowptr[-1])
    rng = np.random.default_rng(0)
    arrs = dict(rowptr=rowptr, bonded=rng.random(max(E, 1)), active=np.ones(max(E, 1), np.uint8),
                expiry=np.full(max(E, 1), np.iinfo(np.int64).max, np.int64), sigma=rng.random(A), omega=rng.random(A),
                consensus=np.zeros(A, np.uint8), ring=np.full(A, 3, np.uint8))
    d = {k: ctx.dev_alloc(v.nbytes) for k, v in arrs.items()}
    for k, v in arrs.items(): ctx.h2d(d[k], v)
    o = [ctx.dev_alloc(8 * A), ctx.dev_alloc(A), ctx.dev_alloc(A), ctx.dev_alloc(A)]
    ctx.sync()
    f = lambda: ctx.sigma_eff_rings_dev(d["rowptr"], d["bonded"], d["active"], d["expiry"], 0, d["sigma"], d["omega"],
                                        d["consensus"], d["ring"], A, *o)
    for _ in range(3): f()
    ctx.sync(); best = 1e9
    for _ in range(10):
        ctx.timer_start(1); f(); ctx.timer_stop(1); best = min(best, ctx.timer_ms(1))
    print(f"{name:45s} A={A} E={E}: {best*1e3:8.1f} us  ({best*1e6/max(E,1):.2f} ns/edge, {best*1e-3/max(deg.max(),1)*1.965e9:.1f} cyc per edge of the longest row)", flush=True)
    for p in list(d.values()) + o: ctx.dev_free(p)
run("empty tile", [0] * 256)
run("one row of 14017", [14017] + [0] * 255)
run("one row of 2048", [2048] + [0] * 255)
run("one row of 512", [512] + [0] * 255)
run("256 rows of 55", [55] * 256)
run("256 rows of 8", [8] * 256)
run("one row of 14017 + 1e6 rows of 10", [14017] + [10] * 999999)
run("1e6 rows of 10 + one row of 14017 last", [10] * 999999 + [14017])
run("1e6 rows of 10", [10] * 1000000)
