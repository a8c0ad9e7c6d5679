#!/usr/bin/env python3
"""Launches each kernel that tools/prof_kernels.sh profiles a few times at its BASELINE-config shape:
k_sigma_eff_tiled (config 4, Zipf then uniform), k_sha256_var (config-3 shape, 10^6 messages, length-sorted),
k_chain_capture (10^4 sessions x 200 x 256 B) and the tree kernels (10^4 sessions x 1000 leaves)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads
from agent_hypervisor_b200 import Context
ctx = Context(0)
which = set(sys.argv[1:]) or {"sigma", "var", "chain", "tree"}
REPS = 3

if "sigma" in which:
    A, E = 1_000_000, 10_000_000
    g = workloads.config4_graph(A, E)
    keys = ("rowptr", "bonded", "active", "expiry", "sigma", "omega", "consensus", "current_ring")
    d = {k: ctx.dev_alloc(g[k].nbytes) for k in keys}
    for k in keys:
        ctx.h2d(d[k], g[k])
    o = [ctx.dev_alloc(8 * A), ctx.dev_alloc(A), ctx.dev_alloc(A), ctx.dev_alloc(A)]
    for shape in ("zipf", "uniform"):
        if shape == "uniform":
            ctx.h2d(d["rowptr"], (np.arange(A + 1, dtype=np.uint64) * 10).astype(np.uint32))
        ctx.sync()
        for _ in range(REPS):
            ctx.sigma_eff_rings_dev(d["rowptr"], d["bonded"], d["active"], d["expiry"], g["now"], d["sigma"], d["omega"],
                                    d["consensus"], d["current_ring"], A, *o)
        ctx.sync()
    for p in list(d.values()) + o:
        ctx.dev_free(p)

if "var" in which:
    n = 1_000_000
    a, offs = workloads.config3_arena(n)
    d_a, d_o, d_d = ctx.dev_alloc(a.nbytes + 64), ctx.dev_alloc(offs.nbytes), ctx.dev_alloc(32 * n)
    ctx.h2d(d_a, a); ctx.h2d(d_o, offs); ctx.sync()
    for _ in range(REPS):
        ctx.sha256_batch_dev(d_a, a.nbytes, d_o, n, 0, d_d)
    ctx.sync()
    for p in (d_a, d_o, d_d):
        ctx.dev_free(p)

if "chain" in which or "tree" in which:
    S, L, FL = 10_000, (1000 if "tree" in which else 200), 256
    n = S * L
    arena = np.zeros(n * FL, np.uint8)
    offs, hexpos, rp = workloads.config2_fill(arena, 0, S, L)
    d = {k: ctx.dev_alloc(v) for k, v in dict(arena=arena.nbytes, offs=offs.nbytes, hexpos=hexpos.nbytes, rp=rp.nbytes,
                                               digs=32 * n, par=32 * n, kind=n, roots=32 * S, empty=S).items()}
    ctx.h2d(d["arena"], arena); ctx.h2d(d["offs"], offs); ctx.h2d(d["hexpos"], hexpos); ctx.h2d(d["rp"], rp); ctx.sync()
    if "chain" in which:
        Sc = S if L == 200 else S // 5                     # 2 x 10^6 deltas either way
        for _ in range(REPS):
            ctx.chain_capture_dev(d["arena"], arena.nbytes, d["offs"], d["hexpos"], d["rp"], Sc, Sc * L, d["digs"], d["par"], d["kind"])
        ctx.sync()
    if "tree" in which:
        ctx.sha256_batch_dev(d["arena"], arena.nbytes, 0, n, FL, d["digs"])
        for _ in range(REPS):
            ctx.commit_dev(d["arena"], arena.nbytes, 0, FL, 0, S, n, L, L, d["digs"], d["roots"], d["empty"])     # uniform tree
            ctx.merkle_roots_dev(d["digs"], d["rp"], S, n, L, d["roots"], d["empty"])                               # ragged tree
        ctx.sync()
print("prof_kernels done", sorted(which), ctx.launch_count(), "launches")
