#!/usr/bin/env python3
"""SURVEY 8e / config 5 variant: ONE chain of N leaves (default 10^8) sharded over the ranks at
2^k-aligned cuts; each rank hashes its leaves and reduces its sub-trees, the sub-roots are all-gathered
(NCCL) and every rank finishes the top.  Checked against the whole tree rebuilt on rank 0 from the
all-gathered leaf digests.  Run under torchrun; prints one JSON line on rank 0."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import workloads
from agent_hypervisor_b200 import Context, sharding

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
N = int(os.environ.get("GIANT_LEAVES", "100000000")); FL = 256
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = Context(local)
dev = torch.device("cuda", local)
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
t_digs = torch.empty(max(n_loc, 1) * 32, dtype=torch.uint8, device=dev)
n_sub_loc = hi - lo
rp = np.minimum(np.arange(n_sub_loc + 1, dtype=np.uint64) << np.uint64(k), np.uint64(n_loc))
d_rp = ctx.dev_alloc(rp.nbytes); ctx.h2d(d_rp, rp)
cap = max(h_ - l_ for l_, h_ in plan["ranges"])
t_sub = torch.zeros(cap * 32, dtype=torch.uint8, device=dev)
t_all = torch.empty(world * cap * 32, dtype=torch.uint8, device=dev)
d_empty = ctx.dev_alloc(max(n_sub, 1)); d_root = ctx.dev_alloc(32); d_rp_top = ctx.dev_alloc(16)
ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
ctx.sync()

def run():
    if n_loc:
        ctx.sha256_batch_dev(d_arena, n_loc * FL, 0, n_loc, FL, t_digs.data_ptr())
        ctx.merkle_roots_dev(t_digs.data_ptr(), d_rp, n_sub_loc, n_loc, 1 << k, t_sub.data_ptr(), d_empty)
        tail = int(rp[-1] - rp[-2])
        if tail < (1 << k) and n_sub > 1:                    # lift the ragged tail to level k (delta.py:129)
            depth = 0
            while (1 << depth) < tail: depth += 1
            ctx.sync()
            x = t_sub[(n_sub_loc - 1) * 32: n_sub_loc * 32].cpu().numpy().reshape(1, 32)
            for _ in range(k - depth):
                x, _ = ctx.merkle_roots(np.concatenate([x, x]), np.array([0, 2], np.uint64))
            t_sub[(n_sub_loc - 1) * 32: n_sub_loc * 32] = torch.from_numpy(np.ascontiguousarray(x)).to(dev).reshape(-1)
    if world > 1:
        with torch.cuda.stream(ext):
            dist.all_gather_into_tensor(t_all, t_sub)
        allsub = torch.cat([t_all[r * cap * 32: r * cap * 32 + (plan["ranges"][r][1] - plan["ranges"][r][0]) * 32] for r in range(world)])
    else:
        allsub = t_sub[: n_sub * 32]
    allsub = allsub.contiguous()
    rp_top = np.array([0, n_sub], np.uint64); ctx.h2d(d_rp_top, rp_top)
    torch.cuda.synchronize()
    ctx.merkle_roots_dev(allsub.data_ptr(), d_rp_top, 1, n_sub, n_sub, d_root, d_empty)
    root = np.empty(32, np.uint8); ctx.d2h(root, d_root); ctx.sync()
    return root

root = run()                                  # warm-up (sizes the scratch slab)
if world > 1: dist.barrier()
torch.cuda.synchronize(); t0 = time.perf_counter()
root = run()
torch.cuda.synchronize(); dt = time.perf_counter() - t0
if world > 1:
    t = torch.tensor([dt], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
# checker: whole tree on rank 0 from the all-gathered leaf digests
ok = None
counts = [max(0, min(h_ << k, N) - (l_ << k)) for l_, h_ in plan["ranges"]]
capd = max(counts)
pad = torch.zeros(capd * 32, dtype=torch.uint8, device=dev); pad[: n_loc * 32] = t_digs[: n_loc * 32]
if world > 1:
    alld = torch.empty(world * capd * 32, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(alld, pad)
    whole = torch.cat([alld[r * capd * 32: r * capd * 32 + counts[r] * 32] for r in range(world)]).contiguous()
else:
    whole = pad
if rank == 0:
    d_rp_w = ctx.dev_alloc(16); ctx.h2d(d_rp_w, np.array([0, N], np.uint64)); torch.cuda.synchronize()
    d_r2 = ctx.dev_alloc(32)
    ctx.merkle_roots_dev(whole.data_ptr(), d_rp_w, 1, N, N, d_r2, d_empty)
    r2 = np.empty(32, np.uint8); ctx.d2h(r2, d_r2); ctx.sync()
    ok = bool(np.array_equal(r2, root))
    print(json.dumps({"leaves": N, "n_gpus": world, "k": k, "sub_roots": n_sub, "seconds": dt,
                      "leaves_per_s": N / dt, "root": root.tobytes().hex(), "matches_whole_tree": ok}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
