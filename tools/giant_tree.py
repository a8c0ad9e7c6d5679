#!/usr/bin/env python3
"""SURVEY 8e / config 5 variant: ONE chain of N leaves (default 10^8) sharded over the ranks at
2^k-aligned cuts; each rank hashes its leaves and reduces its sub-trees, the sub-roots are all-gathered by the
library's own NCCL communicator (ahv_comm_all_gather_dev: one collective) and every rank finishes the top.
Checked against the whole tree rebuilt on rank 0 from the all-gathered leaf digests.  Run under torchrun (the
process launcher; torch itself is not imported); prints one JSON line on rank 0."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads
from agent_hypervisor_b200 import Context, sharding
from agent_hypervisor_b200 import comm as ahv_comm

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
N = int(os.environ.get("GIANT_LEAVES", "100000000")); FL = 256
ctx = Context(local)
ahv_comm.init_from_env(ctx, rank, world)
plan = sharding.plan_tree(N, world, k=20 if N >= (1 << 24) else None)
k, n_sub = plan["k"], plan["n_sub"]
lo, hi = plan["ranges"][rank]
first = lo << k
n_loc = max(0, min(hi << k, N) - first)
# leaves: distinct 256-byte messages (config-2 shaped rows with the global leaf index in the session-id digits)
L = 1000
S_loc = (n_loc + L - 1) // L + 1
h = ctx.host_alloc(S_loc * L * FL)
workloads.config2_fill(h, first // L, S_loc, L, FL)
skip = (first % L) * FL
d_arena = ctx.dev_alloc(max(n_loc, 1) * FL)
ctx.h2d(d_arena, h[skip: skip + n_loc * FL]); ctx.sync()
d_digs = ctx.dev_alloc(max(n_loc, 1) * 32)
n_sub_loc = hi - lo
rp = np.minimum(np.arange(n_sub_loc + 1, dtype=np.uint64) << np.uint64(k), np.uint64(n_loc))
d_rp = ctx.dev_alloc(rp.nbytes); ctx.h2d(d_rp, rp)
cap = max(h_ - l_ for l_, h_ in plan["ranges"])
d_sub = ctx.dev_alloc(cap * 32); ctx.memset(d_sub, 0, cap * 32)
d_all = ctx.dev_alloc(world * cap * 32)
d_top = ctx.dev_alloc(n_sub * 32)
d_empty = ctx.dev_alloc(max(n_sub, 1)); d_root = ctx.dev_alloc(32); d_rp_top = ctx.dev_alloc(16)
ctx.h2d(d_rp_top, np.array([0, n_sub], np.uint64))
ctx.sync()


def run():
    if n_loc:
        ctx.sha256_batch_dev(d_arena, n_loc * FL, 0, n_loc, FL, d_digs)
        ctx.merkle_roots_dev(d_digs, d_rp, n_sub_loc, n_loc, 1 << k, d_sub, d_empty)
        tail = int(rp[-1] - rp[-2])
        if tail < (1 << k) and n_sub > 1:                    # lift the ragged tail to level k (delta.py:129)
            depth = 0
            while (1 << depth) < tail: depth += 1
            x = np.empty((1, 32), np.uint8)
            ctx.d2h(x, d_sub + (n_sub_loc - 1) * 32); ctx.sync()
            for _ in range(k - depth):
                x, _ = ctx.merkle_roots(np.concatenate([x, x]), np.array([0, 2], np.uint64))
            ctx.h2d(d_sub + (n_sub_loc - 1) * 32, np.ascontiguousarray(x))
    # ONE collective: every rank's (padded) sub-roots
    ctx.comm_all_gather_dev(d_sub, d_all, cap * 32)
    # compact the padded blocks into the ordered list of sub-roots (device-to-device copies through the host API)
    pos = 0
    allsub = np.empty((world, cap, 32), np.uint8)
    ctx.d2h(allsub, d_all); ctx.sync()
    subs = np.concatenate([allsub[r, : plan["ranges"][r][1] - plan["ranges"][r][0]] for r in range(world)])
    ctx.h2d(d_top, subs)
    ctx.merkle_roots_dev(d_top, d_rp_top, 1, n_sub, n_sub, d_root, d_empty)
    root = np.empty(32, np.uint8); ctx.d2h(root, d_root); ctx.sync()
    return root


root = run()                                  # warm-up (sizes the scratch slab)
ctx.barrier()
t0 = time.perf_counter()
root = run()
dt = float(ctx.comm_all_gather(np.array([time.perf_counter() - t0])).max())
# checker: whole tree on rank 0 from the all-gathered leaf digests
counts = [max(0, min(h_ << k, N) - (l_ << k)) for l_, h_ in plan["ranges"]]
capd = max(counts)
d_pad = ctx.dev_alloc(capd * 32); ctx.memset(d_pad, 0, capd * 32)
loc = np.empty(max(n_loc, 1) * 32, np.uint8); ctx.d2h(loc, d_digs); ctx.sync()
ctx.h2d(d_pad, loc[: n_loc * 32]); ctx.sync()
d_alld = ctx.dev_alloc(world * capd * 32)
ctx.comm_all_gather_dev(d_pad, d_alld, capd * 32); ctx.sync()
if rank == 0:
    alld = np.empty((world, capd * 32), np.uint8); ctx.d2h(alld, d_alld); ctx.sync()
    whole = np.concatenate([alld[r, : counts[r] * 32] for r in range(world)])
    d_whole = ctx.dev_alloc(N * 32); ctx.h2d(d_whole, whole)
    d_rp_w = ctx.dev_alloc(16); ctx.h2d(d_rp_w, np.array([0, N], np.uint64))
    d_r2 = ctx.dev_alloc(32)
    ctx.merkle_roots_dev(d_whole, d_rp_w, 1, N, N, d_r2, d_empty)
    r2 = np.empty(32, np.uint8); ctx.d2h(r2, d_r2); ctx.sync()
    ok = bool(np.array_equal(r2, root))
    print(json.dumps({"leaves": N, "n_gpus": world, "k": k, "sub_roots": n_sub, "seconds": dt,
                      "leaves_per_s": N / dt, "root": root.tobytes().hex(), "matches_whole_tree": ok,
                      "collective": "ahv_comm_all_gather_dev (ncclAllGather inside libahv_b200.so)", "comm": ctx.comm_info()}), flush=True)
ctx.barrier()
ctx.comm_destroy()
