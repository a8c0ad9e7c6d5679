#!/usr/bin/env python3
"""sigma_eff pass: single-tile latencies and the config-4 graphs at full size (development tool).

    python tools/sigma_lab.py            # tiles + config 4 (Zipf and uniform), every long-row cut
    python tools/sigma_lab.py --config4  # only the config-4 part (what tools/prof_kernels.sh profiles)
Writes gpurun_out/sigma_lab.json."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads
from agent_hypervisor_b200 import Context
ctx = Context(0)
out = {}
LAST_BATCH_MS = 0.0


def upload(arrs):
    d = {k: ctx.dev_alloc(max(v.nbytes, 8)) for k, v in arrs.items()}
    for k, v in arrs.items():
        ctx.h2d(d[k], v)
    ctx.sync()
    return d


def time_pass(d, A, now, reps=20):
    o = [ctx.dev_alloc(8 * A), ctx.dev_alloc(A), ctx.dev_alloc(A), ctx.dev_alloc(A)]
    f = lambda: ctx.sigma_eff_rings_dev(d["rowptr"], d["bonded"], d["active"], d["expiry"], now, d["sigma"], d["omega"],
                                        d["consensus"], d["current_ring"], A, *o)
    for _ in range(3):
        f()
    ctx.sync()
    ts = []
    for _ in range(reps):
        ctx.timer_start(1); f(); ctx.timer_stop(1); ts.append(ctx.timer_ms(1))
    # and 20 launches back to back between ONE pair of events (what bench.py reports): no per-launch event latency
    ctx.timer_start(1)
    for _ in range(reps):
        f()
    ctx.timer_stop(1)
    global LAST_BATCH_MS
    LAST_BATCH_MS = ctx.timer_ms(1) / reps
    for p in o:
        ctx.dev_free(p)
    return min(ts), float(np.median(ts))


def run_tile(name, deg):
    deg = np.asarray(deg, dtype=np.int64); A = len(deg)
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint32); E = int(rowptr[-1])
    rng = np.random.default_rng(0)
    arrs = dict(rowptr=rowptr, bonded=rng.random(max(E, 1)), active=np.ones(max(E, 1), np.uint8),
                expiry=np.full(max(E, 1), np.iinfo(np.int64).max, np.int64), sigma=rng.random(A), omega=rng.random(A),
                consensus=np.zeros(A, np.uint8), current_ring=np.full(A, 3, np.uint8))
    d = upload(arrs)
    res = {}
    for exact in (1, 0):
        ctx.set_option("sigma_exact", exact)
        best, med = time_pass(d, A, 0, reps=10)
        res["exact" if exact else "default"] = best * 1e3
    ctx.set_option("sigma_exact", 0)
    print(f"{name:45s} A={A} E={E}: exact {res['exact']:8.1f} us   default {res['default']:8.1f} us", flush=True)
    out.setdefault("tiles", {})[name] = res
    for p in d.values():
        ctx.dev_free(p)


def run_config4():
    A, E = 1_000_000, 10_000_000
    nbytes = workloads.config4_bytes(A, E)
    g = workloads.config4_graph(A, E)
    u = dict(g)
    rng = np.random.default_rng(1)
    u["rowptr"] = (np.arange(A + 1, dtype=np.uint64) * 10).astype(np.uint32)      # degree 10 everywhere, same edge arrays
    for name, gr in (("zipf", g), ("uniform", u)):
        arrs = {k: gr[k] for k in ("rowptr", "bonded", "active", "expiry", "sigma", "omega", "consensus", "current_ring")}
        d = upload(arrs)
        res = {}
        for label, opts in (("exact", dict(sigma_exact=1)), ("cut16", dict(sigma_long_cut=16)),
                            ("cut32", dict(sigma_long_cut=32)), ("cut64", dict(sigma_long_cut=64)), ("cut128", dict(sigma_long_cut=128)),
                            ("cut32_noshare", dict(sigma_long_cut=32, sigma_share=0)),
                            ("cut32_chunk1024", dict(sigma_long_cut=32, sigma_chunk=1024)),
                            ("cut32_chunk2048", dict(sigma_long_cut=32, sigma_chunk=2048)),
                            ("cut32_chunk1440", dict(sigma_long_cut=32, sigma_chunk=1440)),
                            ("cut64_chunk1440", dict(sigma_long_cut=64, sigma_chunk=1440)),
                            ("cut32_chunk1296", dict(sigma_long_cut=32, sigma_chunk=1296)),
                            ("cut64_chunk1296", dict(sigma_long_cut=64, sigma_chunk=1296)),
                            ("cut64_chunk1440_heavy2048", dict(sigma_long_cut=64, sigma_chunk=1440, sigma_heavy=2048)),
                            ("heavy2048", dict(sigma_heavy=2048)), ("heavy3072", dict(sigma_heavy=3072)),
                            ("heavy6144", dict(sigma_heavy=6144)), ("heavy8192", dict(sigma_heavy=8192)),
                            ("stream_cut32", dict(sigma_long_cut=32, sigma_variant=2)),
                            ("row_per_thread_exact", dict(sigma_exact=1, sigma_variant=0))):
            for k, v in opts.items():
                ctx.set_option(k, v)
            best, med = time_pass(d, A, gr["now"])
            ctx.set_option("sigma_exact", 0); ctx.set_option("sigma_long_cut", 64); ctx.set_option("sigma_variant", 1); ctx.set_option("sigma_chunk", 1440); ctx.set_option("sigma_share", 1); ctx.set_option("sigma_heavy", 3072)
            res[label] = {"best_us": best * 1e3, "median_us": med * 1e3, "GBps_best": nbytes / best / 1e6,
                          "GBps_median": nbytes / med / 1e6, "back_to_back_us": LAST_BATCH_MS * 1e3,
                          "GBps_back_to_back": nbytes / LAST_BATCH_MS / 1e6}
            print(f"config4 {name:8s} {label:22s} best {best*1e3:7.1f} us  median {med*1e3:7.1f} us  {nbytes/med/1e6:7.0f} GB/s   "
                  f"back-to-back {LAST_BATCH_MS*1e3:7.1f} us {nbytes/LAST_BATCH_MS/1e6:7.0f} GB/s", flush=True)
        out.setdefault("config4", {})[name] = res
        for p in d.values():
            ctx.dev_free(p)
    out["config4"]["algorithmic_bytes"] = nbytes
    out["config4"]["longest_row"] = int(np.diff(g["rowptr"].astype(np.int64)).max())


if "--config4" not in sys.argv:
    run_tile("empty tile", [0] * 256)
    run_tile("one row of 14017", [14017] + [0] * 255)
    run_tile("one row of 2048", [2048] + [0] * 255)
    run_tile("one row of 512", [512] + [0] * 255)
    run_tile("one row of 513", [513] + [0] * 255)
    run_tile("256 rows of 55", [55] * 256)
    run_tile("256 rows of 8", [8] * 256)
    run_tile("1e6 rows of 10 + one row of 14017 last", [10] * 999999 + [14017])
    run_tile("1e6 rows of 10", [10] * 1000000)
run_config4()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sigma_lab.json"), "w"), indent=1)
