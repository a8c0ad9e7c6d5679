#!/usr/bin/env python3
"""Slashing pass (ahv_slash_cascade_dev) on device-resident batches: time per launch, sessions/s,
clips/s, and the plain-C restatement on the host cores beside it.  Development / measurement tool:

    python tools/slash_lab.py [--sessions 1000000] [--out profiles/r01/slash_lab.json]
"""
import argparse, ctypes as C, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads
from agent_hypervisor_b200 import Context
from agent_hypervisor_b200 import _native as nat

ap = argparse.ArgumentParser()
ap.add_argument("--sessions", type=int, default=1_000_000)
ap.add_argument("--out", default="")
ap.add_argument("--no-cpu", action="store_true")
ap.add_argument("--slice", type=int, default=0, help="lane kernel: bytes of shared memory per warp (0: what the host entry picks)")
ap.add_argument("--variant", type=int, default=0, help="slash_variant for --only runs")
ap.add_argument("--only", default="", help="agents,records[,omega]: run just this shape (for ncu)")
args = ap.parse_args()
ctx = Context(0)
results = []


def host_entry_slice(S, A, E):
    """The slice ahv_slash_cascade (host entry, which knows E) picks for a batch - auto_lane_slice in
    csrc/ahv_slashing.cu; this tool times ahv_slash_cascade_dev on resident arrays and sets the same value."""
    avg = 11 * -(-A // S) + 2 * -(-E // S) + 16
    if 32 * avg <= 7040:
        return 7040
    if 32 * avg <= 9216:
        return 9216
    if 32 * avg <= 10752:
        return 10752
    return 7040 if 4 * avg <= 7040 else 10752


def run(name, b, cpu=True):
    S = len(b["session_agent_rowptr"]) - 1
    A = len(b["scores"]); E = len(b["voucher"]); NE = len(b["event_agent"])
    host = dict(rowptr=b["rowptr"], voucher=b["voucher"], active=b["active"], expiry=b["expiry"],
                session_agent_rowptr=b["session_agent_rowptr"], event_rowptr=b["event_rowptr"], event_agent=b["event_agent"],
                event_sigma=b["event_sigma"], event_omega=b["event_omega"], scores=b["scores"])
    outs = dict(touched=A, slash_log=32 * (NE + E), clip_log=24 * E, slash_count=4 * S, clip_count=4 * S, status=S)
    d = {k: ctx.dev_alloc(max(v.nbytes, 16)) for k, v in host.items()}
    d.update({k: ctx.dev_alloc(max(n, 16)) for k, n in outs.items()})
    for k, v in host.items():
        ctx.h2d(d[k], v)
    sb = nat.SlashBatch()
    sb.n_agents, sb.now, sb.n_sessions = A, int(b["now"]), S
    for k in list(host) + list(outs):
        setattr(sb, k, d[k])
    sb.event_depth = None
    lib, H = ctx.lib, ctx.h
    slice_bytes = args.slice or host_entry_slice(S, A, E)
    ctx.set_option("slash_slice", slice_bytes)

    def reset():
        ctx.h2d(d["active"], host["active"]); ctx.h2d(d["scores"], host["scores"])
        ctx.memset(d["touched"], 0, A)

    times = []
    for it in range(6):
        reset(); ctx.sync()
        ctx.timer_start(1); ctx.slash_cascade_dev(sb); ctx.timer_stop(1)
        ctx.sync()
        if it:
            times.append(ctx.timer_ms(1))
    ns = np.empty(S, np.uint32); nc = np.empty(S, np.uint32); stt = np.empty(S, np.uint8)
    ctx.d2h(ns, d["slash_count"]); ctx.d2h(nc, d["clip_count"]); ctx.d2h(stt, d["status"]); ctx.sync()
    assert not stt.any()
    ms = float(np.median(times))
    n_slash, n_clip = int(ns.sum()), int(nc.sum())
    algo = workloads.slash_bytes(b, n_slash, n_clip)
    r = dict(case=name, lane_slice=slice_bytes, sessions=S, agents=A, records=E, slash_calls=n_slash, clips=n_clip, ms=ms, ms_min=float(min(times)),
             sessions_per_s=S / (ms * 1e-3), clips_per_s=n_clip / (ms * 1e-3), algorithmic_bytes=algo,
             algorithmic_GBps=algo / (ms * 1e-3) / 1e9)
    if cpu and not args.no_cpu:
        from oracle import coracle as co
        t0 = time.perf_counter()
        want = co.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                                b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"])
        r["cpu_c_restatement_ms"] = (time.perf_counter() - t0) * 1e3
        r["cpu_threads"] = co.num_threads()
        sc = np.empty(A); ctx.d2h(sc, d["scores"]); ctx.sync()
        r["matches_cpu"] = bool(np.array_equal(want[5], ns) and np.array_equal(want[6], nc)
                                and np.array_equal(np.nan_to_num(want[0], nan=-1.0), np.nan_to_num(sc, nan=-1.0)))
    print(json.dumps(r), flush=True)
    results.append(r)
    for p in d.values():
        ctx.dev_free(p)


S = args.sessions
if args.only:
    ctx.set_option("slash_variant", args.variant)
    p = args.only.split(",")
    run(f"{S} sessions x {p[0]} agents x {p[1]} records", workloads.slash_sessions(S, int(p[0]), int(p[1]), omega=float(p[2]) if len(p) > 2 else 0.95), cpu=False)
    sys.exit(0)
for variant in (1, 2, 0):
    ctx.set_option("slash_variant", variant)
    tag = {0: "lanes+staged", 1: "global-only", 2: "staged"}[variant]
    run(f"[{tag}] {S} sessions x 16 agents x 48 records, omega 0.95", workloads.slash_sessions(S, 16, 48), cpu=False)
    run(f"[{tag}] {S // 4} sessions x 64 agents x 512 records, omega 0.95", workloads.slash_sessions(S // 4, 64, 512), cpu=False)
    run(f"[{tag}] {S // 2} sessions x 32 agents x 256 records, omega 0.95", workloads.slash_sessions(S // 2, 32, 256), cpu=False)
    run(f"[{tag}] {4 * S} sessions x 4 agents x 6 records, omega 0.95", workloads.slash_sessions(4 * S, 4, 6), cpu=False)
run(f"{S} sessions x 16 agents x 48 records, omega 0.95", workloads.slash_sessions(S, 16, 48))
run(f"{S} sessions x 16 agents x 48 records, omega 0.5 (no cascades)", workloads.slash_sessions(S, 16, 48, omega=0.5))
run(f"{S // 4} sessions x 64 agents x 512 records, omega 0.95", workloads.slash_sessions(S // 4, 64, 512))
run(f"{S // 2} sessions x 32 agents x 256 records, omega 0.95", workloads.slash_sessions(S // 2, 32, 256))
run(f"{S // 8} sessions x 128 agents x 1024 records, omega 0.95 (3 sessions per warp slice)", workloads.slash_sessions(S // 8, 128, 1024))
run(f"{S // 16} sessions x 256 agents x 2048 records, omega 0.95 (warp per session)", workloads.slash_sessions(S // 16, 256, 2048))
run("one session x 1e5 agents x 1e6 records, omega 0.95", workloads.slash_sessions(1, 100_000, 1_000_000))
if args.out:
    with open(os.path.join(ROOT, args.out), "w") as f:
        json.dump(results, f, indent=1)
