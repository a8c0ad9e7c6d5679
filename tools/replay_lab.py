#!/usr/bin/env python3
"""Replay / import of stored sessions from COLUMNS: host C++ packer -> one upload ->
commit + verify on the GPU, with the reference's json.dumps expression timed beside the packer.

    python tools/replay_lab.py [--sessions 2000] [--deltas 1000] [--out profiles/r01/replay_lab.json]

Benchmark-shaped deltas (benchmarks/bench_hypervisor.py:147-184): one VFSChange each, chained parents.
"""
import argparse, json, os, sys, time
from datetime import datetime, timedelta, timezone
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
from agent_hypervisor_b200.audit.packing import pack_columns, _str_col

ap = argparse.ArgumentParser()
ap.add_argument("--sessions", type=int, default=2000)
ap.add_argument("--deltas", type=int, default=1000)
ap.add_argument("--out", default="")
ap.add_argument("--no-gpu", action="store_true")
args = ap.parse_args()
S, L = args.sessions, args.deltas
n = S * L
T0 = datetime(2026, 1, 1, tzinfo=timezone.utc)
EPOCH = datetime(1970, 1, 1, tzinfo=timezone.utc)
t0_us = int((T0 - EPOCH).total_seconds()) * 1_000_000

t = time.perf_counter()
idx = np.arange(n)
turn = (idx % L) + 1
cols = {
    "n": n,
    "delta_id": _str_col([f"delta:{i}" for i in turn]),
    "session_id": _str_col([f"session:{s:08d}" for s in range(S) for _ in range(L)]),
    "agent_did": _str_col(["did:mesh:a"] * n),
    "turn_id": turn.astype(np.int64),
    "timestamp_us": (t0_us + (idx % L) * 1_000_000 + 123456).astype(np.int64),
    "timestamp_utcoffset_s": np.zeros(n, np.int32),
    "parent_kind": np.where(idx % L == 0, 0, 1).astype(np.uint8),            # None for the first of a session
    "parent_digest": np.zeros(32 * n, np.uint8),                             # placeholders, chained below
    "parent_other": _str_col([""] * n),
    "change_rowptr": np.arange(n + 1, dtype=np.uint64),
    "change_path": _str_col([f"/f{i}" for i in turn]),
    "change_op": _str_col(["add"] * n),
    "change_content": _str_col([f"h{i}" for i in turn]),
    "change_prev": _str_col([None] * n),
}
t_cols = time.perf_counter() - t
res = {"sessions": S, "deltas_per_session": L, "deltas": n, "columns_build_s": t_cols, "host_cores": os.cpu_count()}

# -- the C++ packer (all host threads), twice: the second run has warm pages
for k in range(2):
    t = time.perf_counter()
    arena, offs, ppos = pack_columns(cols, want_parent_pos=True)
    dt = time.perf_counter() - t
res.update(packer_s=dt, packer_deltas_per_s=n / dt, packer_GBps=arena.nbytes / dt / 1e9, arena_bytes=int(arena.nbytes),
           mean_delta_bytes=arena.nbytes / n)

# -- the reference's expression (audit/delta.py:43-62) on a sample, one core
m = min(n, 50_000)
t = time.perf_counter()
for i in range(m):
    tid = int(turn[i])
    json.dumps({"delta_id": f"delta:{tid}", "turn_id": tid, "session_id": f"session:{i // L:08d}", "agent_did": "did:mesh:a",
                "timestamp": (T0 + timedelta(seconds=tid - 1, microseconds=123456)).isoformat(),
                "changes": [{"path": f"/f{tid}", "operation": "add", "content_hash": f"h{tid}", "previous_hash": None}],
                "parent_hash": None if tid == 1 else "0" * 64}, sort_keys=True).encode()
dt = time.perf_counter() - t
res.update(json_dumps_deltas_per_s_one_core=m / dt, json_dumps_sample=m)

if not args.no_gpu:
    ctx = Context(0)
    rp = (np.arange(S + 1, dtype=np.uint64) * np.uint64(L))
    arena = np.array(arena)                                  # writable: the chain is written into it
    t = time.perf_counter()
    digs, par, kind = ctx.chain_capture(arena, offs, ppos, rp)               # chains the parents on the device
    res["chain_capture_s"] = time.perf_counter() - t
    for k in range(3):
        t = time.perf_counter()
        o = ctx.audit_sessions(arena, offs, None, None, rp, parent_pos=ppos)
        dt = time.perf_counter() - t
    assert o["verdict"].all() and o["root_match"].all() and np.array_equal(o["digests"], digs)
    res.update(audit_sessions_pageable_s=dt, audit_pageable_deltas_per_s=n / dt)
    # the same with every buffer pinned (what a replay service would hold on to): the packer writes straight
    # into the pinned arena
    h_arena = ctx.host_alloc(arena.nbytes); h_arena[:] = arena
    h_offs = ctx.host_alloc(offs.nbytes, np.uint64); h_offs[:] = offs
    h_pos = ctx.host_alloc(ppos.nbytes, np.uint32); h_pos[:] = ppos
    h_rp = ctx.host_alloc(rp.nbytes, np.uint64); h_rp[:] = rp
    out = dict(digests=ctx.host_alloc(32 * n).reshape(n, 32), roots=ctx.host_alloc(32 * S).reshape(S, 32),
               empty=ctx.host_alloc(S), verdict=ctx.host_alloc(S), first_bad=ctx.host_alloc(8 * S, np.int64),
               root_match=ctx.host_alloc(S))
    for k in range(3):
        t = time.perf_counter()
        o = ctx.audit_sessions(h_arena, h_offs, None, None, h_rp, parent_pos=h_pos, out=out)
        dt = time.perf_counter() - t
    assert o["verdict"].all() and o["root_match"].all() and np.array_equal(o["digests"], digs)
    scratch = ctx.host_alloc(arena.nbytes)
    t = time.perf_counter()
    pack_columns(cols, arena=scratch, want_parent_pos=True)
    res["packer_into_pinned_s"] = time.perf_counter() - t
    res.update(audit_sessions_s=dt, audit_deltas_per_s=n / dt, h2d_bytes=int(arena.nbytes + offs.nbytes + ppos.nbytes + rp.nbytes),
               d2h_bytes=int(32 * n + 43 * S),
               columns_to_verdicts_deltas_per_s=n / (dt + res["packer_into_pinned_s"]))
    # -- pipelined: the same sessions in 8 chunks, packing chunk k+1 while chunk k is audited
    from agent_hypervisor_b200 import ReplayPipeline
    K = 8
    Sc = S // K
    def chunk_cols(k):
        lo, hi = k * Sc * L, (k + 1) * Sc * L
        c = {"n": hi - lo}
        for nm, v in cols.items():
            if nm == "n":
                continue
            if isinstance(v, tuple):                          # (blob, offsets, nulls): string columns
                blob, off, null = v
                c[nm] = (blob, (off[lo:hi + 1] - off[lo]).astype(np.uint64) if len(off) == n + 1 else off, null[lo:hi] if len(null) == n else null)
                c[nm] = (blob[int(off[lo]):int(off[hi])] if len(off) == n + 1 and off[hi] > off[lo] else np.zeros(1, np.uint8), c[nm][1], c[nm][2])
            elif nm == "parent_digest":
                c[nm] = par.reshape(-1)[32 * lo:32 * hi]      # the chained parents computed above
            elif nm == "change_rowptr":
                c[nm] = (v[lo:hi + 1] - v[lo]).astype(np.uint64)
            else:
                c[nm] = v[lo:hi]
        return c, np.arange(Sc + 1, dtype=np.uint64) * np.uint64(L)
    chunks = [chunk_cols(k) for k in range(K)]
    with ReplayPipeline(int(arena.nbytes // K * 1.05) + 4096, Sc * L, Sc, ctx) as pipe:
        ok = all(bool(r["verdict"].all()) and np.array_equal(r["digests"], digs[k * Sc * L:(k + 1) * Sc * L])
                 for k, r in enumerate(pipe.run(chunks, copy=False)))                  # checked outside the timing
        for rep in range(3):
            t = time.perf_counter()
            nd = 0
            for r in pipe.run(chunks, copy=False):
                ok = ok and bool(r["verdict"].all())
                nd += len(r["digests"])
            dt = time.perf_counter() - t
    assert ok and nd == K * Sc * L
    res.update(pipelined_chunks=K, pipelined_s=dt, pipelined_columns_to_verdicts_deltas_per_s=nd / dt)
    # -- the same chunks with the packer ON THE DEVICE: columns staged in pinned memory, one upload per chunk on the second
    #    stream, ahv_pack_deltas_dev -> extract parents -> commit -> verify; no host core serialises anything
    from agent_hypervisor_b200 import DeviceReplay
    from agent_hypervisor_b200.audit.packing import column_bytes
    colb = max(column_bytes(c[0]) for c in chunks)
    with DeviceReplay(int(arena.nbytes // K * 1.05) + 4096, Sc * L, Sc, colb, ctx) as dev:
        ok = all(bool(r["verdict"].all()) and np.array_equal(r["digests"], digs[k * Sc * L:(k + 1) * Sc * L])
                 for k, r in enumerate(dev.run(chunks, copy=False)))
        for rep in range(3):
            t = time.perf_counter()
            nd = 0
            for r in dev.run(chunks, copy=False):
                ok = ok and bool(r["verdict"].all())
                nd += len(r["digests"])
            dt = time.perf_counter() - t
        # columns already resident in pinned host memory (an archive reader filling pinned buffers): no host copy
        staged = [dev.prestage(*ch) for ch in chunks]
        for rep in range(3):
            t = time.perf_counter()
            nd2 = 0
            for r in dev.run(staged, copy=False):
                ok = ok and bool(r["verdict"].all())
                nd2 += len(r["digests"])
            dt_staged = time.perf_counter() - t
        ok = ok and all(np.array_equal(r["digests"], digs[k * Sc * L:(k + 1) * Sc * L]) for k, r in enumerate(dev.run(staged, copy=False)))
        for sc in staged:
            sc.close()
        res.update(device_packer_prestaged_s=dt_staged, device_packer_prestaged_columns_to_verdicts_deltas_per_s=nd2 / dt_staged)
        # where the time goes for one chunk: host staging copy, upload, device packer, the audit kernels
        t = time.perf_counter(); ch = dev._stage_chunk(0, *chunks[0]); ctx.sync(); t_stage = time.perf_counter() - t
        ctx.upload_fence()
        ctx.timer_start(2)
        ctx.pack_deltas_dev(ch["st"], dev._d["offs"], dev._d["arena"], dev.limits[0], dev._d["ppos"])
        ctx.timer_stop(2)
        t_pack_ms = ctx.timer_ms(2)
        ctx.timer_start(2); dev._enqueue(ch); ctx.timer_stop(2)
        t_all_ms = ctx.timer_ms(2); ctx.sync()
    assert ok and nd == K * Sc * L
    res.update(device_packer_pipelined_s=dt, device_packer_columns_to_verdicts_deltas_per_s=nd / dt, column_bytes_per_delta=colb / (Sc * L),
               device_packer_chunk={"deltas": Sc * L, "stage_and_upload_s": t_stage, "pack_kernels_ms": t_pack_ms,
                                    "pack_deltas_per_s": Sc * L / t_pack_ms * 1e3, "pack_parents_commit_verify_download_ms": t_all_ms})
print(json.dumps(res, indent=1))
if args.out:
    with open(os.path.join(ROOT, args.out), "w") as f:
        json.dump(res, f, indent=1)
