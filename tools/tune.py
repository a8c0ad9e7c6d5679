#!/usr/bin/env python3
"""Kernel timing sweep on one GPU (development tool; not part of the product or the bench contract).

Times, with CUDA events on the context stream: the integer-pipe microbenchmarks, the fixed-size leaf
kernel variants x block sizes, the variable-length leaf kernel (sorted / unsorted), the tree, verify,
and the sigma_eff pass.  Writes a JSON summary to gpurun_out/tune.json."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads  # noqa: E402
from agent_hypervisor_b200 import Context  # noqa: E402


def timed(ctx, fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    ctx.sync()
    best = 1e30
    tot = 0.0
    for _ in range(reps):
        ctx.timer_start(1)
        fn()
        ctx.timer_stop(1)
        ms = ctx.timer_ms(1)
        best = min(best, ms)
        tot += ms
    return best, tot / reps


def main():
    S = int(os.environ.get("TUNE_SESSIONS", "10000"))
    L, FL = 1000, 256
    n = S * L
    ctx = Context(0)
    out = {"device": ctx.device_info(), "S": S}
    for mode, iters in ((0, 30000), (1, 3000), (2, 3000)):
        v, ms = ctx.int_peak(mode, iters)
        out[f"int_peak_mode{mode}_tinstr"] = v / 1e12
        print(f"int_peak mode {mode}: {v/1e12:.3f} Tinstr/s ({ms:.2f} ms)", flush=True)

    arena = np.empty(n * FL, np.uint8)
    offs, hexpos, rp = workloads.config2_fill(arena, 0, S, L, FL)
    d = {k: ctx.dev_alloc(v) for k, v in dict(arena=arena.nbytes, offs=offs.nbytes, hexpos=hexpos.nbytes, rp=rp.nbytes,
                                               digs=32 * n, par=32 * n, kind=n, roots=32 * S, empty=S, verdict=S,
                                               fb=8 * S, match=S, roots2=32 * S).items()}
    ctx.h2d(d["arena"], arena); ctx.h2d(d["offs"], offs); ctx.h2d(d["hexpos"], hexpos); ctx.h2d(d["rp"], rp)
    t0 = time.perf_counter()
    ctx.chain_capture_dev(d["arena"], arena.nbytes, d["offs"], d["hexpos"], d["rp"], S, n, d["digs"], d["par"], d["kind"])
    ctx.sync()
    out["chain_capture_ms"] = (time.perf_counter() - t0) * 1e3
    print("chain capture ms", out["chain_capture_ms"], flush=True)
    comp_leaf = n * 5

    res = {}
    for comp in (0, 1):
        ctx.set_option("compress", comp)
        best, avg = timed(ctx, lambda: ctx.sha256_batch_dev(d["arena"], arena.nbytes, 0, n, FL, d["digs"]))
        res[f"leaf_fixed_compress{comp}"] = {"best_ms": best, "Gcomp_s": comp_leaf / best / 1e6}
        b2, _ = timed(ctx, lambda: ctx.commit_dev(d["arena"], arena.nbytes, 0, FL, 0, S, n, L, L, d["digs"], d["roots"], d["empty"]), reps=3)
        b3, _ = timed(ctx, lambda: ctx.sha256_batch_dev(d["arena"], arena.nbytes, d["offs"], n, 0, d["digs"]), reps=3)
        print(f"compress={comp}: leaf fixed {best:.3f} ms ({comp_leaf/best/1e6:.2f} Gcomp/s), commit {b2:.3f} ms, leaf var {b3:.3f} ms", flush=True)
        res[f"commit_compress{comp}"] = {"best_ms": b2}
        res[f"leaf_var_compress{comp}"] = {"best_ms": b3}
    for variant in (0, 1, 2):
        for block in (64, 128, 256):
            ctx.set_option("fixed_variant", variant); ctx.set_option("block", block)
            best, avg = timed(ctx, lambda: ctx.sha256_batch_dev(d["arena"], arena.nbytes, 0, n, FL, d["digs"]))
            res[f"leaf_fixed_v{variant}_b{block}"] = {"best_ms": best, "avg_ms": avg, "Gcomp_s": comp_leaf / best / 1e6,
                                                      "Tinstr_s": comp_leaf * 1400 / best / 1e9}
            print(f"leaf fixed variant {variant} block {block}: best {best:.3f} ms avg {avg:.3f} ms "
                  f"-> {comp_leaf/best/1e6:.2f} Gcomp/s {comp_leaf*1400/best/1e9:.2f} Tinstr/s", flush=True)
    ctx.set_option("fixed_variant", 0); ctx.set_option("block", 128)
    for sort in (0, 1):
        ctx.set_option("sort", sort)
        best, avg = timed(ctx, lambda: ctx.sha256_batch_dev(d["arena"], arena.nbytes, d["offs"], n, 0, d["digs"]), reps=3)
        res[f"leaf_var_sort{sort}"] = {"best_ms": best, "avg_ms": avg, "Gcomp_s": comp_leaf / best / 1e6}
        print(f"leaf var (same data) sort={sort}: best {best:.3f} ms -> {comp_leaf/best/1e6:.2f} Gcomp/s", flush=True)
    ctx.set_option("sort", 1)
    _, node_c = workloads.count_compressions_fixed(n, FL, L)
    best, avg = timed(ctx, lambda: ctx.commit_dev(d["arena"], arena.nbytes, 0, FL, 0, S, n, L, L, d["digs"], d["roots"], d["empty"]))
    res["commit_uniform"] = {"best_ms": best, "avg_ms": avg}
    print(f"commit (leaf+tree uniform): best {best:.3f} ms", flush=True)
    ctx.sha256_batch_dev(d["arena"], arena.nbytes, 0, n, FL, d["digs"])
    for top in (256, 64, 32, 16, 8, 4, 2):
        ctx.set_option("top_max", top)
        # tree alone = commit minus leaves is not separable through the ABI; time the ragged tree and the uniform commit
        best, avg = timed(ctx, lambda: ctx.commit_dev(d["arena"], arena.nbytes, 0, FL, 0, S, n, L, L, d["digs"], d["roots"], d["empty"]), reps=3)
        res[f"commit_uniform_top{top}"] = {"best_ms": best}
        b2, _ = timed(ctx, lambda: ctx.merkle_roots_dev(d["digs"], d["rp"], S, n, L, d["roots2"], d["empty"]), reps=3)
        res[f"tree_ragged_top{top}"] = {"best_ms": b2, "Gcomp_s": node_c / b2 / 1e6}
        print(f"top_max {top}: commit {best:.3f} ms, ragged tree {b2:.3f} ms ({node_c/b2/1e6:.2f} Gcomp/s)", flush=True)
    ctx.set_option("top_max", 0)
    best, avg = timed(ctx, lambda: ctx.merkle_roots_dev(d["digs"], d["rp"], S, n, L, d["roots2"], d["empty"]))
    res["tree_ragged"] = {"best_ms": best, "avg_ms": avg, "Gcomp_s": node_c / best / 1e6}
    print(f"tree ragged path: best {best:.3f} ms -> {node_c/best/1e6:.2f} Gcomp/s", flush=True)
    best, avg = timed(ctx, lambda: ctx.verify_chains_dev(d["arena"], arena.nbytes, 0, FL, d["par"], d["kind"], 0, S, n, L, L,
                                                         d["roots"], d["verdict"], d["fb"], 0, d["roots2"], d["match"]))
    res["verify_uniform"] = {"best_ms": best, "avg_ms": avg}
    print(f"verify (leaf+links+tree+compare): best {best:.3f} ms", flush=True)
    out["audit"] = res
    for p in d.values():
        ctx.dev_free(p)

    # variable-length (config 3 shape, reduced n)
    n3 = int(os.environ.get("TUNE_N3", "1000000"))
    a3, o3 = workloads.config3_arena(n3)
    comp3 = int(((np.diff(o3.astype(np.int64)) + 9 + 63) // 64).sum())
    d3 = {k: ctx.dev_alloc(v) for k, v in dict(arena=a3.nbytes + 64, offs=o3.nbytes, digs=32 * n3).items()}
    ctx.h2d(d3["arena"], a3); ctx.h2d(d3["offs"], o3); ctx.sync()
    for sort in (0, 1):
        ctx.set_option("sort", sort)
        best, avg = timed(ctx, lambda: ctx.sha256_batch_dev(d3["arena"], a3.nbytes, d3["offs"], n3, 0, d3["digs"]), reps=3)
        out[f"config3_sort{sort}"] = {"best_ms": best, "Gcomp_s": comp3 / best / 1e6, "GBps": a3.nbytes / best / 1e6,
                                      "n": n3, "bytes": int(a3.nbytes)}
        print(f"config3 n={n3} sort={sort}: best {best:.3f} ms -> {comp3/best/1e6:.2f} Gcomp/s, {a3.nbytes/best/1e6:.1f} GB/s", flush=True)
    ctx.set_option("sort", 1)
    for p in d3.values():
        ctx.dev_free(p)

    # sigma_eff (config 4)
    A, E = 1_000_000, 10_000_000
    g = workloads.config4_graph(A, E)
    E = int(g["rowptr"][-1])
    names = dict(rowptr=g["rowptr"], bonded=g["bonded"], active=g["active"], expiry=g["expiry"], sigma=g["sigma"],
                 omega=g["omega"], consensus=g["consensus"], current_ring=g["current_ring"])
    dd = {k: ctx.dev_alloc(v.nbytes) for k, v in names.items()}
    for k, v in names.items():
        ctx.h2d(dd[k], v)
    o = {k: ctx.dev_alloc(sz) for k, sz in dict(se=8 * A, ring=A, dem=A, near=A).items()}
    ctx.sync()
    best, avg = timed(ctx, lambda: ctx.sigma_eff_rings_dev(dd["rowptr"], dd["bonded"], dd["active"], dd["expiry"], g["now"],
                                                          dd["sigma"], dd["omega"], dd["consensus"], dd["current_ring"], A,
                                                          o["se"], o["ring"], o["dem"], o["near"]), reps=10)
    by = workloads.config4_bytes(A, E)
    out["sigma_eff"] = {"best_ms": best, "avg_ms": avg, "GBps": by / best / 1e6, "bytes": by, "A": A, "E": E}
    print(f"sigma_eff A={A} E={E}: best {best:.4f} ms -> {by/best/1e6:.1f} GB/s", flush=True)
    # same sizes, every row exactly 10 edges (no long-row tail): the bulk streaming rate
    rp_u = (np.arange(A + 1, dtype=np.uint64) * (E // A)).astype(np.uint32)
    ctx.h2d(dd["rowptr"], rp_u); ctx.sync()
    E_u = (E // A) * A
    best, avg = timed(ctx, lambda: ctx.sigma_eff_rings_dev(dd["rowptr"], dd["bonded"], dd["active"], dd["expiry"], g["now"],
                                                          dd["sigma"], dd["omega"], dd["consensus"], dd["current_ring"], A,
                                                          o["se"], o["ring"], o["dem"], o["near"]), reps=10)
    by_u = workloads.config4_bytes(A, min(E_u, E))
    out["sigma_eff_uniform10"] = {"best_ms": best, "GBps": by_u / best / 1e6}
    print(f"sigma_eff uniform degree 10: best {best:.4f} ms -> {by_u/best/1e6:.1f} GB/s", flush=True)
    print("max degree in config-4 graph:", int(np.diff(g["rowptr"].astype(np.int64)).max()), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "tune.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
