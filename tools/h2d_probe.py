#!/usr/bin/env python3
"""Host -> device copy ceiling of the box, beside the end-to-end audit call.

Run under torchrun with N ranks (one per GPU; torch is not imported).  With K = 1, 2, 4, ... N of the ranks active
at a time (the others wait at the barrier), each active rank copies its 2.6 GB pinned arena to its GPU:
  * one cudaMemcpyAsync for the whole arena, and 20 chunked copies (what ahv_audit_sessions issues);
  * pinned (cudaHostAllocDefault) and write-combined pinned memory;
  * with and without the process pinned to the cores NVML names for its GPU (set BEFORE the allocation);
  * device -> host of 320 MB of digests alone;
  * the end-to-end call (ahv_audit_sessions: upload, commit + verify, results back) on the same arena.
Rank 0 prints one JSON line: per-rank and aggregate GB/s for every K."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
import workloads

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
AFFINITY = os.environ.get("PROBE_AFFINITY", "1") == "1"
host = {"applied": False}
if AFFINITY:
    sys.path.insert(0, ROOT)
    from bench import pin_near_gpu
    host = pin_near_gpu(local)
from agent_hypervisor_b200 import Context
from agent_hypervisor_b200 import comm as ahv_comm
ctx = Context(local)
ahv_comm.init_from_env(ctx, rank, world)

S, L, FL = int(os.environ.get("PROBE_SESSIONS", "10000")), 1000, 256
n = S * L
nbytes = n * FL
h_def = ctx.host_alloc(nbytes)
offs, hexpos, rp = workloads.config2_fill(h_def, rank * S, S, L, FL)
h_wc = ctx.host_alloc(nbytes, write_combined=True)
h_wc[:] = h_def
d_arena = ctx.dev_alloc(nbytes)
d_digs = ctx.dev_alloc(32 * n)
h_digs = ctx.host_alloc(32 * n)
h_pos = ctx.host_alloc(4 * n, np.uint32); h_pos[:] = hexpos
h_out = {k: ctx.host_alloc(v) for k, v in dict(roots=32 * S, empty=S, verdict=S, fb=8 * S, match=S).items()}
ctx.sync()
vp = C.c_void_p


def copy_whole(h):
    ctx.h2d(d_arena, h); ctx.sync()


def copy_chunked(h, chunks=20):
    per = (nbytes // chunks + 255) & ~255
    for k in range(chunks):
        lo, hi = k * per, min(nbytes, (k + 1) * per)
        if lo < hi:
            ctx.lib.ahv_memcpy_h2d(ctx.h, vp(d_arena + lo), vp(h.ctypes.data + lo), hi - lo)
    ctx.sync()


def d2h_digs():
    ctx.d2h(h_digs, d_digs); ctx.sync()


def e2e(h, with_digests=True):
    rc = ctx.lib.ahv_audit_sessions(ctx.h, vp(h.ctypes.data), None, FL, None, None, vp(h_pos.ctypes.data), None, S, L, None,
                                    vp(h_digs.ctypes.data) if with_digests else None, vp(h_out["roots"].ctypes.data),
                                    vp(h_out["empty"].ctypes.data), vp(h_out["verdict"].ctypes.data), vp(h_out["fb"].ctypes.data),
                                    vp(h_out["match"].ctypes.data))
    assert rc == 0, ctx.lib.ahv_last_error()


def measure(fn, active, reps=3):
    """Seconds per repetition on the active ranks (max over them); inactive ranks only take part in the barriers."""
    if active:
        fn()
    ctx.barrier()
    t0 = time.perf_counter()
    if active:
        for _ in range(reps):
            fn()
    dt = (time.perf_counter() - t0) / reps if active else 0.0
    ctx.barrier()
    return float(ctx.comm_all_gather(np.array([dt])).max())


# the chained arena (so that the e2e call verifies): chain on the device once, copy back into both host buffers
d_tmp = {k: ctx.dev_alloc(v.nbytes) for k, v in dict(offs=offs, hexpos=hexpos, rp=rp).items()}
ctx.h2d(d_arena, h_def); ctx.h2d(d_tmp["offs"], offs); ctx.h2d(d_tmp["hexpos"], hexpos); ctx.h2d(d_tmp["rp"], rp)
d_par, d_kind = ctx.dev_alloc(32 * n), ctx.dev_alloc(n)
ctx.chain_capture_dev(d_arena, nbytes, d_tmp["offs"], d_tmp["hexpos"], d_tmp["rp"], S, n, d_digs, d_par, d_kind)
ctx.d2h(h_def, d_arena); ctx.sync()
h_wc[:] = h_def
ctx.dev_free(d_par); ctx.dev_free(d_kind)

out = {"world": world, "sessions_per_rank": S, "arena_bytes_per_rank": nbytes, "affinity": host, "by_active_ranks": {}}
K = 1
while K <= world:
    act = rank < K
    row = {}
    for name, fn, bytes_ in (("h2d_whole_pinned", lambda: copy_whole(h_def), nbytes),
                             ("h2d_20chunks_pinned", lambda: copy_chunked(h_def), nbytes),
                             ("h2d_whole_write_combined", lambda: copy_whole(h_wc), nbytes),
                             ("d2h_digests_pinned", d2h_digs, 32 * n),
                             ("e2e_audit_pinned", lambda: e2e(h_def), nbytes + 4 * n),
                             ("e2e_audit_write_combined", lambda: e2e(h_wc), nbytes + 4 * n),
                             ("e2e_audit_pinned_no_digests", lambda: e2e(h_def, False), nbytes + 4 * n)):
        dt = measure(fn, act)
        row[name] = {"seconds": dt, "GBps_per_rank": bytes_ / dt / 1e9, "GBps_aggregate": K * bytes_ / dt / 1e9}
        if name.startswith("e2e"):
            row[name]["deltas_per_s_aggregate"] = K * n / dt
    out["by_active_ranks"][str(K)] = row
    K *= 2
ok = bool(h_out["verdict"].all() and h_out["match"].all())
if rank == 0:
    out["verified"] = ok
    print(json.dumps(out), flush=True)
ctx.barrier()
ctx.comm_destroy()
