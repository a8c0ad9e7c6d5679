#!/usr/bin/env python3
"""bench.py - headline benchmark: audit deltas Merkle-committed + verified per second
(BASELINE.json configs[1]: 10k sessions x 1k fixed-size 256 B deltas per GPU).

    python bench.py --gpus N --steps K --warmup W            # CUDA arm (this repo)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU pipeline (hashlib port)

One step = one commit pass (leaf SHA-256 + Merkle roots, audit/delta.py:63,117-134) plus one
verify pass (re-hash, parent-link compare, root rebuild + compare, audit/delta.py:136-152 and
audit/commitment.py:59-64) over every session resident on the GPU.  Prints ONE JSON line.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
REF_COPY = os.path.join(ROOT, "baseline", "_ref")          # copy of the reference tree made by __graft_entry__.build()

SESSIONS_PER_GPU = 10_000
LEAVES = 1000
MSG_LEN = 256
METRIC = "audit deltas Merkle-committed+verified/sec"
UNIT = "deltas/s"
ALGO_INSTR_PER_COMPRESSION = 1400


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock / throttle-reason samples DURING the timed region: NVML in-process every 10 ms (a few
    microseconds per query, measured not to perturb the step); `nvidia-smi -lms 200` is the fallback."""

    def __init__(self, device: int):
        self.device, self.rows, self.proc, self._stop, self._thr, self.max_mhz = device, [], None, False, None, None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.device)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}

            def loop():
                while not self._stop:
                    try:
                        mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                        r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                        self.rows.append((time.perf_counter(), mhz, [k for k, b in bits.items() if r & b], pw))
                    except Exception:
                        pass
                    time.sleep(0.01)

            self._thr = threading.Thread(target=loop, daemon=True)
            self._thr.start()
            self.source = "nvml"
        except Exception:
            self.source = "nvidia-smi"
            try:
                q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                     "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                     "clocks_event_reasons.sw_power_cap")
                self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                              "-lms", "200", "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
                threading.Thread(target=self._read, daemon=True).start()
            except Exception:
                self.proc = None

    def _read(self):
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            try:
                self.max_mhz = float(f[1])
                self.rows.append((time.perf_counter(), float(f[0]),
                                  [n for n, v in zip(names, f[3:7]) if v.lower().startswith("active")], float(f[2])))
            except Exception:
                pass

    def stop(self):
        self._stop = True
        if self.proc:
            self.proc.terminate()

    def summary(self, t0, t1):
        rows = [r for r in self.rows if t0 <= r[0] <= t1] or self.rows[-3:]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        sm = sorted(r[1] for r in rows)
        reasons = sorted({x for r in rows for x in r[2]})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "power_w_max": max(r[3] for r in rows), "samples": len(rows), "source": self.source}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
_SHARED = {}     # inherited by forked workers (no pickling of the arena)


def _cpu_commit_verify(args):
    """One worker: the reference's per-session pipeline at arena level (hashlib), sessions [s0, s1)."""
    import hashlib
    from oracle import audit_oracle as ao
    s0, s1 = args
    arena, L, flen, parents_hex = _SHARED["arena"], _SHARED["L"], _SHARED["flen"], _SHARED["parents"]
    mv = memoryview(arena)
    roots, ok = [], True
    for s in range(s0, s1):
        base = s * L * flen
        pre = [mv[base + i * flen: base + (i + 1) * flen] for i in range(L)]
        hexes = [hashlib.sha256(p).hexdigest() for p in pre]            # delta.py:63 (what capture stored)
        roots.append(ao.merkle_root_hex(hexes))                         # delta.py:117-134
        v, _, _ = ao.verify_chain_session(pre, parents_hex[s])          # delta.py:136-152 (re-hashes)
        ok = ok and v
    return roots, ok


def cpu_pool(arena: np.ndarray, L: int, flen: int, parents_hex, workers: int):
    import multiprocessing as mp
    _SHARED.update(arena=arena, L=L, flen=flen, parents=parents_hex)
    return mp.get_context("fork").Pool(workers)


def cpu_pipeline(S: int, workers: int, pool):
    """Commit + verify sessions [0, S) of the shared arena.  Returns (seconds, roots, ok)."""
    per = max(1, (S + 4 * workers - 1) // (4 * workers))
    jobs = [(w, min(S, w + per)) for w in range(0, S, per)]
    t0 = time.perf_counter()
    res = pool.map(_cpu_commit_verify, jobs)
    dt = time.perf_counter() - t0
    roots = [r for rs, _ in res for r in rs]
    return dt, roots, all(ok for _, ok in res)


def cpu_chain_sample(S: int, L: int, flen: int, first_session: int = 0):
    """Config-2 sessions chained on the CPU with hashlib (reference arm input synthesis)."""
    import hashlib
    import workloads
    arena = np.zeros(S * L * flen, np.uint8)
    offs, hexpos, rp = workloads.config2_fill(arena, first_session, S, L, flen)
    a = arena.reshape(S, L, flen)
    parents = []
    for s in range(S):
        ph = [None]
        parent = None
        for i in range(L):
            if i:
                a[s, i, hexpos[i]:hexpos[i] + 64] = np.frombuffer(parent.encode(), np.uint8)
                ph.append(parent)
            parent = hashlib.sha256(a[s, i].tobytes()).hexdigest()
        parents.append(ph)
    return arena, parents


def reference_own(cores: int, budget_s: float = 60.0):
    """The UNMODIFIED reference, from the copy under baseline/_ref (never from the original tree): its own
    benchmarks/bench_hypervisor.py run as a subprocess, config 1 (one session, 1000 benchmark-shaped deltas:
    capture, compute_merkle_root, verify_chain timed separately) and the commit+verify metric of this bench
    with the real DeltaEngine objects on all host cores.  None if the copy is not there."""
    src = os.path.join(REF_COPY, "src")
    if not os.path.isdir(os.path.join(src, "hypervisor")):
        return None
    out = {"kind": "reference", "tree": "baseline/_ref (copy of the reference, unmodified)"}
    # (1) the reference's own benchmark file, as it is
    script = os.path.join(REF_COPY, "benchmarks", "bench_hypervisor.py")
    try:
        t0 = time.perf_counter()
        r = subprocess.run([sys.executable, script], cwd=os.path.join(REF_COPY, "benchmarks"), capture_output=True,
                           text=True, timeout=600)
        own = {"rc": r.returncode, "seconds": time.perf_counter() - t0}
        res = os.path.join(REF_COPY, "benchmarks", "results", "benchmarks.json")
        if r.returncode == 0 and os.path.exists(res):
            data = json.load(open(res))
            rows = data.get("benchmarks", data if isinstance(data, list) else [])
            keep = {}
            for row in rows:
                if isinstance(row, dict) and "name" in row:
                    keep[row["name"]] = {k: row[k] for k in ("mean_us", "median_us", "p99_us", "ops_per_sec") if k in row}
            own["results"] = keep
        else:
            own["stdout_tail"] = (r.stdout or "")[-600:]
            own["stderr_tail"] = (r.stderr or "")[-300:]
        out["bench_hypervisor_py"] = own
    except Exception as e:                                   # a reported baseline must not break the bench line
        out["bench_hypervisor_py"] = {"error": repr(e)}
    # (2) + (3) with the real classes, in worker processes that import the copy
    code = r"""
import json, statistics, sys, time
sys.path.insert(0, sys.argv[1])
from hypervisor.audit.delta import DeltaEngine, VFSChange
mode, n_sessions = sys.argv[2], int(sys.argv[3])
def build(sid):
    de = DeltaEngine(sid)
    for i in range(1000):
        de.capture("did:mesh:a", [VFSChange(path=f"/f{i}", operation="add", content_hash=f"h{i}")])
    return de
if mode == "config1":
    def med(fn, reps, warm):
        ts = []
        for i in range(reps + warm):
            t0 = time.perf_counter(); r = fn(); dt = time.perf_counter() - t0
            if i >= warm: ts.append(dt)
        return statistics.median(ts) * 1e3, r
    t_cap, de = med(lambda: build("bench-session"), 5, 1)
    t_root, root = med(de.compute_merkle_root, 20, 2)
    t_ver, ok = med(de.verify_chain, 20, 2)
    print(json.dumps({"capture_1000_ms": t_cap, "merkle_root_ms": t_root, "verify_chain_ms": t_ver, "verify": bool(ok), "root": root}))
else:
    engines = [build(f"session:{k}") for k in range(n_sessions)]
    t0 = time.perf_counter()
    ok = True
    for de in engines:
        r = de.compute_merkle_root()          # core.py:204
        ok = de.verify_chain() and ok         # audit/delta.py:136-152 (re-serialises and re-hashes every delta)
    print(json.dumps({"seconds": time.perf_counter() - t0, "deltas": 1000 * n_sessions, "verify": bool(ok)}))
"""
    try:
        r = subprocess.run([sys.executable, "-c", code, src, "config1", "1"], capture_output=True, text=True, timeout=300)
        out["config1"] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"rc": r.returncode, "stderr_tail": r.stderr[-300:]}
        per = 4                                                        # sessions per worker: ~0.1 s of timed work each
        procs = [subprocess.Popen([sys.executable, "-c", code, src, "pipeline", str(per)], stdout=subprocess.PIPE, text=True)
                 for _ in range(cores)]
        outs = [json.loads(p.communicate(timeout=budget_s * 5)[0].strip().splitlines()[-1]) for p in procs]
        worst = max(o["seconds"] for o in outs)
        n = sum(o["deltas"] for o in outs)
        out["commit_verify"] = {"value": n / worst, "unit": UNIT, "cores": cores, "verify": all(o["verify"] for o in outs),
                                "sample": f"{cores} processes x {per} sessions x 1000 benchmark-shaped deltas: compute_merkle_root + "
                                          f"verify_chain of the real DeltaEngine (json.dumps + hashlib per delta); throughput = all "
                                          f"deltas / slowest process"}
    except Exception as e:
        out["commit_verify"] = {"error": repr(e)}
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on all host cores.  The line's value is
    its hashlib pipeline restated at arena level (oracle port: pre-packed pre-images of the configs[1] recipe, so
    WITHOUT the reference's json.dumps cost - this flatters the reference); the unmodified reference itself,
    run from the copy under baseline/_ref, is reported beside it under cpu_baseline.reference_own."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    S = args.ref_sessions or max(cores, min(64 * cores, 4096))
    if not args.ref_sessions and cores > 64:
        S = min(S, 16 * cores)
    arena, parents = cpu_chain_sample(S, LEAVES, MSG_LEN)
    with cpu_pool(arena, LEAVES, MSG_LEN, parents, cores) as pool:
        for _ in range(max(args.warmup, 1)):
            cpu_pipeline(min(S, cores), cores, pool)
        total = 0.0
        ok = True
        for _ in range(args.steps):
            dt, roots, v = cpu_pipeline(S, cores, pool)
            total += dt
            ok = ok and v
    n = S * LEAVES
    value = n * args.steps / total
    sample = f"{S} sessions x {LEAVES} x {MSG_LEN} B per step (config-2 recipe), hashlib arena-level port, {cores} processes"
    own = None if args.no_reference_own else reference_own(cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"configs[1] recipe (1k fixed-size 256 B deltas per session, commit + verify), bounded sample of "
                               f"{S} sessions per step (the CUDA arm runs 10k sessions per GPU per step; throughput is per delta)",
                   "sample_sessions": S, "leaves_per_session": LEAVES, "delta_bytes": MSG_LEN,
                   "note": "pre-packed arenas: the reference's json.dumps serialisation cost is NOT included in `value`; "
                           "cpu_baseline.reference_own has the unmodified reference with it"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "reference_own": own if own is not None else "baseline/_ref not present (built only where /root/reference exists)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "verified": bool(ok),
    }
    emit_line(line)
    return 0


# ----------------------------------------------------------------------------- CUDA arm
def pin_near_gpu(device: int):
    """Restrict this process to the host cores next to its GPU (NVML's ideal CPU affinity) BEFORE any pinned
    buffer is allocated, so that first touch places the upload buffers on the GPU's NUMA node."""
    info = {"applied": False}
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(device)
        words = (os.cpu_count() + 63) // 64
        mask = nv.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        info["nvml_cpus"] = len(cpus)
        try:
            info["numa_node"] = int(open(f"/sys/bus/pci/devices/{nv.nvmlDeviceGetPciInfo(h).busId.decode().lower()[4:]}/numa_node").read())
        except Exception:
            pass
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            info["applied"] = True
        info["cpus_used"] = len(use or allowed)
    except Exception as e:
        info["error"] = repr(e)[:120]
    return info


def timed_ms(ctx, fn, reps=3, warm=1):
    """CUDA-event time of fn on the context's stream: (best, mean) over reps after warm-up."""
    for _ in range(warm):
        fn()
    ctx.sync()
    ts = []
    for _ in range(reps):
        ctx.timer_start(1); fn(); ctx.timer_stop(1)
        ts.append(ctx.timer_ms(1))
    return min(ts), float(np.mean(ts))


def timed_batch_ms(ctx, fn, reps=20, warm=3):
    """Per-launch time of fn when `reps` launches are enqueued back to back between ONE pair of CUDA events: for a
    kernel of tens of microseconds a pair of events around every single launch would add its own launch / event
    latency (an empty kernel measures ~9 us that way) to each sample."""
    for _ in range(warm):
        fn()
    ctx.sync()
    ctx.timer_start(1)
    for _ in range(reps):
        fn()
    ctx.timer_stop(1)
    return ctx.timer_ms(1) / reps


def extra_configs(ctx, peaks, peak_dual, check: bool):
    """The other BASELINE.json configs at their stated sizes on one GPU, timed with CUDA events after the headline
    region (device-resident inputs); with check=True a sample of each is compared with the oracle (checker only)."""
    import workloads
    out = {}
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    co = None
    if check:
        from oracle import coracle as co
    # ---- configs[2]: variable-length deltas 64 B..16 KB, 10^7 leaves: hash, hash + 10^4 trees, verify
    n1, tiles = 1_000_000, 10
    a1, o1 = workloads.config3_arena(n1)
    B1 = int(o1[-1])
    pad = (-B1) % 16
    if pad:                                                   # keep every tile 16-byte aligned: the last message of a tile grows by the pad
        a1 = np.concatenate([a1, np.zeros(pad, np.uint8)])
        o1 = o1.copy(); o1[-1] += np.uint64(pad)
    B1p = B1 + pad
    n = n1 * tiles
    offs = np.concatenate([o1[:-1] + np.uint64(k * B1p) for k in range(tiles)] + [np.array([tiles * B1p], np.uint64)])
    comp = int(((np.diff(o1.astype(np.int64)) + 9 + 63) // 64).sum()) * tiles
    S3, L3 = n // 1000, 1000
    d = {k: ctx.dev_alloc(v) for k, v in dict(arena=B1p * tiles + 64, offs=offs.nbytes, digs=32 * n, par=32 * n, kind=n,
                                               roots=32 * S3, roots2=32 * S3, empty=S3, verdict=S3, fb=8 * S3, match=S3).items()}
    for k in range(tiles):
        ctx.h2d(d["arena"] + k * B1p, a1)
    ctx.h2d(d["offs"], offs)
    ctx.sync()
    hash_ms, _ = timed_ms(ctx, lambda: ctx.sha256_batch_dev(d["arena"], B1p * tiles, d["offs"], n, 0, d["digs"]), reps=2)
    commit_ms, _ = timed_ms(ctx, lambda: ctx.commit_dev(d["arena"], B1p * tiles, d["offs"], 0, 0, S3, n, L3, L3, d["digs"],
                                                        d["roots"], d["empty"]), reps=2)
    # parents = digest of the message before (first of a session: None): built on the host from the digests
    h_digs = np.empty((n, 32), np.uint8)
    ctx.d2h(h_digs, d["digs"]); ctx.sync()
    par = np.zeros((n, 32), np.uint8); par[1:] = h_digs[:-1]
    kind = np.ones(n, np.uint8); kind[::L3] = 0
    ctx.h2d(d["par"], par); ctx.h2d(d["kind"], kind); ctx.sync()
    verify_ms, _ = timed_ms(ctx, lambda: ctx.verify_chains_dev(d["arena"], B1p * tiles, d["offs"], 0, d["par"], d["kind"], 0, S3, n,
                                                               L3, L3, d["roots"], d["verdict"], d["fb"], 0, d["roots2"], d["match"]),
                            reps=2)
    h_verdict = np.empty(S3, np.uint8); h_match = np.empty(S3, np.uint8); h_roots = np.empty((S3, 32), np.uint8)
    ctx.d2h(h_verdict, d["verdict"]); ctx.d2h(h_match, d["match"]); ctx.d2h(h_roots, d["roots"]); ctx.sync()
    ok3 = bool(h_verdict.all() and h_match.all()) and all(np.array_equal(h_digs[:n1], h_digs[k * n1:(k + 1) * n1]) for k in (1, tiles - 1))
    if check:
        ns = 20_000                                           # first 20 sessions of tile 0 against the C oracle
        want = co.sha256_batch(a1, o1[: ns + 1])
        wr, _ = co.merkle_roots(want, np.arange(ns // L3 + 1, dtype=np.uint64) * np.uint64(L3))
        ok3 = ok3 and np.array_equal(want, h_digs[:ns]) and np.array_equal(wr, h_roots[: ns // L3])
    for p_ in d.values():
        ctx.dev_free(p_)
    msg_bytes = B1 * tiles
    out["var_len"] = {
        "workload": "configs[2]: 10^7 messages, lengths log-uniform 64 B..16 KB (a 10^6-message arena tiled 10x on the "
                    "device), tightly packed (every byte alignment), length-sorted launch order; 10^4 sessions x 1000",
        "messages": n, "message_bytes": msg_bytes, "compressions": comp,
        "hash_ms": hash_ms, "compressions_per_s": comp / hash_ms * 1e3, "message_GBps": msg_bytes / hash_ms / 1e6,
        "frac_of_int_peak": comp * ALGO_INSTR_PER_COMPRESSION / (hash_ms * 1e-3) / peak_dual,
        "commit_ms": commit_ms, "verify_ms": verify_ms,
        "deltas_committed_verified_per_s": n / ((commit_ms + verify_ms) * 1e-3), "verified": bool(ok3),
        "checked_against_oracle": "first 20 000 digests + 20 roots" if check else None}
    del h_digs, par, kind, a1, offs

    # ---- configs[3]: sigma_eff + ring gating, 10^6 agents, 10^7 vouch records (Zipf-like in-degree, and degree 10)
    A, E = 1_000_000, 10_000_000
    g = workloads.config4_graph(A, E)
    nbytes = workloads.config4_bytes(A, E)
    keys = ("rowptr", "bonded", "active", "expiry", "sigma", "omega", "consensus", "current_ring")
    dd = {k: ctx.dev_alloc(g[k].nbytes) for k in keys}
    for k in keys:
        ctx.h2d(dd[k], g[k])
    o = [ctx.dev_alloc(8 * A), ctx.dev_alloc(A), ctx.dev_alloc(A), ctx.dev_alloc(A)]
    ctx.sync()
    run = lambda: ctx.sigma_eff_rings_dev(dd["rowptr"], dd["bonded"], dd["active"], dd["expiry"], g["now"], dd["sigma"],
                                          dd["omega"], dd["consensus"], dd["current_ring"], A, *o)
    sig = {"workload": "configs[3]: 10^6 agents, 10^7 vouch records, one session; sigma_eff, ring, demotion flag, near-threshold list",
           "algorithmic_bytes": nbytes, "hbm_peak_gbs": hbm_peak, "longest_row": int(np.diff(g["rowptr"].astype(np.int64)).max()),
           "sum_order": "rows of up to 64 records in the reference's order (bit-identical), longer rows by one warp in a fixed shape"}
    for shape in ("zipf", "uniform"):
        if shape == "uniform":
            ctx.h2d(dd["rowptr"], (np.arange(A + 1, dtype=np.uint64) * 10).astype(np.uint32)); ctx.sync()
        best, single = timed_ms(ctx, run, reps=20, warm=3)
        mean = timed_batch_ms(ctx, run, reps=20, warm=3)
        sig[shape] = {"ms": mean, "GBps": nbytes / mean / 1e6, "frac": nbytes / mean / 1e6 / hbm_peak,
                      "timing": "20 launches back to back between one pair of CUDA events (inputs of 202 MB exceed the L2)",
                      "single_launch_ms": single, "single_launch_best_ms": best,
                      "single_launch_note": "one pair of events per launch: includes ~5-9 us of launch / event latency"}
        if shape == "zipf":
            ctx.set_option("sigma_exact", 1)
            m2 = timed_batch_ms(ctx, run, reps=5, warm=1)
            ctx.set_option("sigma_exact", 0)
            sig["zipf_exact_order"] = {"ms": m2, "GBps": nbytes / m2 / 1e6, "frac": nbytes / m2 / 1e6 / hbm_peak}
            if check:
                run(); got = [np.empty(A, np.float64), np.empty(A, np.uint8), np.empty(A, np.uint8), np.empty(A, np.uint8)]
                for arr, ptr in zip(got, o):
                    ctx.d2h(arr, ptr)
                ctx.sync()
                want = co.sigma_eff_rings(*[g[k] for k in ("rowptr", "bonded", "active", "expiry", "now", "sigma", "omega",
                                                           "consensus", "current_ring")])
                rel = np.abs(got[0] - want[0]) / np.maximum(np.abs(want[0]), 1e-300)
                listed = (got[3] != 0) | (want[3] != 0)
                short = np.diff(g["rowptr"].astype(np.int64)) <= 64
                sig["max_rel_err_vs_oracle"] = float(rel.max())
                sig["bit_identical_rows"] = int((got[0] == want[0]).sum())
                sig["near_threshold_agents"] = int(listed.sum())
                sig["verified"] = bool(rel.max() <= 1e-6 and np.array_equal(got[0][short], want[0][short]) and
                                       np.array_equal(got[1][~listed], want[1][~listed]) and
                                       np.array_equal(got[2][~listed], want[2][~listed]))
    for p_ in list(dd.values()) + o:
        ctx.dev_free(p_)
    out["sigma_eff"] = sig
    return out


def run_cuda(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun (used only as the process launcher)
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), __file__] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"WORLD_SIZE={world} but --gpus {args.gpus}")

    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    host = pin_near_gpu(local) if not args.no_affinity else {"applied": False}

    import workloads
    from agent_hypervisor_b200 import Context
    from agent_hypervisor_b200 import comm as ahv_comm
    ctx = Context(local)
    # the library's own NCCL communicator (one process per GPU; no torch in this process)
    ahv_comm.init_from_env(ctx, rank, world)
    ctx.barrier()
    info = ctx.device_info()
    S, L, FL = args.sessions, LEAVES, MSG_LEN
    n = S * L
    arena_bytes = n * FL

    def max_over_ranks(x: float) -> float:
        return float(ctx.comm_all_gather(np.array([x], np.float64)).max()) if world > 1 else x

    # ---- inputs: template arena in pinned host memory -> device -> chained on the device
    h_arena = ctx.host_alloc(arena_bytes)
    offs, hexpos, rp = workloads.config2_fill(h_arena, rank * S, S, L, FL)
    stride = ctx.gather_stride(S)
    names = dict(arena=arena_bytes, offs=offs.nbytes, hexpos=hexpos.nbytes, rp=rp.nbytes, digs=32 * n, par=32 * n,
                 kind=n, empty=S, verdict=S, fb=8 * S, match=S, roots_v=32 * S, roots=32 * S, gathered=world * stride)
    d = {k: ctx.dev_alloc(v) for k, v in names.items()}
    d_roots, d_verdict = d["roots"], d["verdict"]
    ctx.h2d(d["arena"], h_arena); ctx.h2d(d["offs"], offs); ctx.h2d(d["hexpos"], hexpos); ctx.h2d(d["rp"], rp)
    chain_ms, _ = timed_ms(ctx, lambda: ctx.chain_capture_dev(d["arena"], arena_bytes, d["offs"], d["hexpos"], d["rp"], S, n,
                                                              d["digs"], d["par"], d["kind"]), reps=2, warm=1)
    ctx.d2h(h_arena, d["arena"])          # chained arena back in pinned host memory for the e2e leg
    h_par = ctx.host_alloc(32 * n); h_kind = ctx.host_alloc(n)
    ctx.d2h(h_par, d["par"]); ctx.d2h(h_kind, d["kind"])
    ctx.sync()

    def step():
        # commit: leaf digests + roots ; verify: re-hash, link compare, root rebuild + compare
        ctx.commit_dev(d["arena"], arena_bytes, 0, FL, 0, S, n, L, L, d["digs"], d_roots, d["empty"])
        ctx.verify_chains_dev(d["arena"], arena_bytes, 0, FL, d["par"], d["kind"], 0, S, n, L, L, d_roots,
                              d_verdict, d["fb"], 0, d["roots_v"], d["match"])
        if world > 1 and not args.no_gather:
            # ONE all-gather of every rank's roots + verdicts (33 B per session), issued by the library on its own
            # stream right behind the kernels: no host synchronisation inside a step
            ctx.gather_results_dev(d_roots, d_verdict, S, d["gathered"])

    for _ in range(max(args.warmup, 3)):
        step()
    ctx.sync()
    sampler = ClockSampler(local)
    if not args.no_clocks:                     # every rank watches ITS GPU: one throttled GPU slows every step of all ranks
        sampler.start()
        time.sleep(0.1)
    ctx.sync(); ctx.barrier(); ctx.sync()      # every rank enters the timed region together
    ctx.set_option("leaf_timer", 32)
    launches0 = ctx.launch_count()
    t_wall0 = time.perf_counter()
    ctx.timer_start(0)
    for _ in range(args.steps):
        step()
    ctx.timer_stop(0)
    ms_total = ctx.timer_ms(0)
    ctx.sync()
    t_wall1 = time.perf_counter()
    launches = ctx.launch_count() - launches0
    ctx.set_option("leaf_timer", 0)
    ctx.barrier()
    ms_own = ms_total
    ms_total = max_over_ranks(ms_total)
    if world > 1:
        launches = int(ctx.comm_all_gather(np.array([launches], np.int64)).sum())
    time.sleep(0.15)
    sampler.stop()
    clocks = sampler.summary(t_wall0, t_wall1)
    if world > 1:
        rnames = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        mask = sum(1 << i for i, nm in enumerate(rnames) if nm in clocks["reasons"])
        other = (1 << 8) if any(r_ not in rnames for r_ in clocks["reasons"]) else 0
        mine = np.array([clocks["sm_mhz"] or 0.0, float(mask | other), clocks.get("power_w_max") or 0.0, ms_own], np.float64)
        allr = ctx.comm_all_gather(mine).reshape(world, 4)
        if rank == 0:
            clocks["per_rank_sm_mhz"] = [float(x) for x in allr[:, 0]]
            clocks["sm_mhz"] = float(allr[:, 0].min())                        # the slowest GPU is the one that counts
            union = int(np.bitwise_or.reduce(allr[:, 1].astype(np.int64)))
            clocks["reasons"] = [nm for i, nm in enumerate(rnames) if union >> i & 1] + (["clock sampling unavailable on a rank"] if union >> 8 & 1 else [])
            clocks["per_rank_power_w_max"] = [float(x) for x in allr[:, 2]]
            clocks["per_rank_ms_total"] = [float(x) for x in allr[:, 3]]
    if rank != 0:
        clocks = None

    # leaf-kernel launch durations recorded live inside the timed region (CUDA events on the ctx stream)
    n_leaf = min(16, 2 * args.steps)
    leaf_ms = [ctx.timer_ms(32 + k) for k in range(n_leaf)]
    leaf_avg_ms = float(np.mean(leaf_ms))

    # results of the last step
    h_verdict = np.empty(S, np.uint8); h_match = np.empty(S, np.uint8); h_roots = np.empty(32 * S, np.uint8)
    ctx.d2h(h_verdict, d_verdict); ctx.d2h(h_match, d["match"]); ctx.d2h(h_roots, d_roots)
    h_all = np.empty(world * stride, np.uint8)
    if world > 1:
        ctx.d2h(h_all, d["gathered"])
    ctx.sync()
    verified = bool(h_verdict.all() and h_match.all())
    if world > 1 and not args.no_gather:
        from agent_hypervisor_b200 import sharding
        all_roots, all_verdicts = sharding.unpack_gathered(h_all, S, world)
        verified = verified and np.array_equal(all_roots[rank * S:(rank + 1) * S].reshape(-1), h_roots) and bool(all_verdicts.all())
        # every rank holds every rank's roots: all sessions distinct, none left unwritten
        verified = verified and len({r.tobytes() for r in all_roots[::max(1, S // 64)]}) == len(all_roots[::max(1, S // 64)])

    # ---- end-to-end through the host-pointer C ABI (pinned host buffers, copies inside the timed region)
    e2e_steps = max(1, min(args.steps, 3))
    h_digs = ctx.host_alloc(32 * n)
    lib, H = ctx.lib, ctx.h
    import ctypes as C
    vp = C.c_void_p
    h_roots_e = ctx.host_alloc(32 * S); h_empty_e = ctx.host_alloc(S)
    h_verdict_e = ctx.host_alloc(S); h_fb_e = ctx.host_alloc(8 * S, np.int64); h_match_e = ctx.host_alloc(S)

    h_pos = ctx.host_alloc(4 * n, np.uint32)
    h_pos[:] = hexpos                            # where each pre-image holds its parent_hash value

    def e2e_step(with_digests=True):
        # one upload of the batch (pre-images + 4-byte parent positions), then the commit pass and the
        # verify pass on the device, results back
        rc = lib.ahv_audit_sessions(H, vp(h_arena.ctypes.data), None, FL, None, None, vp(h_pos.ctypes.data),
                                    None, S, L, None, vp(h_digs.ctypes.data) if with_digests else None,
                                    vp(h_roots_e.ctypes.data), vp(h_empty_e.ctypes.data), vp(h_verdict_e.ctypes.data),
                                    vp(h_fb_e.ctypes.data), vp(h_match_e.ctypes.data))
        assert rc == 0, lib.ahv_last_error()

    def wall(fn, reps):
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return max_over_ranks(time.perf_counter() - t0)

    e2e_step()                                   # warm-up (sizes the staging slab)
    e2e_s = wall(e2e_step, e2e_steps)
    verified = verified and bool(h_verdict_e.all() and h_match_e.all()) and np.array_equal(h_roots_e, h_roots)
    e2e_nodig_s = wall(lambda: e2e_step(False), e2e_steps)
    # the plain-copy ceiling beside it: the same pinned arena, host -> device, nothing else, all ranks at once
    def copy_only():
        ctx.h2d(d["arena"], h_arena); ctx.sync()
    copy_only()
    copy_s = wall(copy_only, e2e_steps)
    h2d = arena_bytes + 4 * n                            # arena + parent positions, once per step
    d2h = 32 * n + 32 * S + S + S + 8 * S + S            # digests, roots, empty, verdict, first_bad, match

    if rank != 0:
        ctx.barrier()
        ctx.comm_destroy()
        return 0

    # ---- roofline of the dominant kernel (leaf SHA-256; 2 launches per step)
    # integer rates measured live by register-only kernels on all SMs (csrc/ahv_peak.cu)
    peak_alu, _ = ctx.int_peak(0, 30000)          # SHF+LOP3 only: the ALU pipe alone (64 lanes/clk/SM)
    peak_dual, _ = ctx.int_peak(10, 30000)        # SHF+LOP3 co-issued with IMAD: both integer-capable pipes
    peak_native, _ = ctx.int_peak(2, 3000)        # whole compressions, plain additions (compiler's placement)
    peak_scaled, _ = ctx.int_peak(15, 3000)       # whole compressions, additions as IMAD chains (what the kernels run)
    comp_per_launch = n * ((FL + 9 + 63) // 64)
    algo_instr = comp_per_launch * ALGO_INSTR_PER_COMPRESSION
    achieved = algo_instr / (leaf_avg_ms * 1e-3)
    leaf_c, node_c = workloads.count_compressions_fixed(n, FL, L)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_bytes = n * FL + 32 * n
    traffic, traffic_src = None, None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "leaf_kernel_traffic.json")))
        traffic, traffic_src = tr["dram_bytes_per_launch"], tr.get("source", "ncu --set full, round 1 (profiles/r01/prof_audit_raw.csv)")
    except Exception:
        pass
    # instructions SHA-256 can only issue on the ALU pipe (funnel shifts, 3-input logic, byte swaps):
    # 1040 per data block, 656 for the constant-schedule padding block
    alu_only = n * (4 * 1040 + 656)
    roofline = {
        "kernel": "k_sha256_fixed (leaf SHA-256, one 256 B delta per thread)",
        "bound": "int32-pipe", "achieved": achieved / 1e12, "peak": peak_dual / 1e12, "unit": "Tinstr/s",
        "frac": achieved / peak_dual, "traffic": traffic,
        "traffic_source": f"not observed by this run: {traffic_src}" if traffic is not None else None,
        "algorithmic": f"{ALGO_INSTR_PER_COMPRESSION} int instr/compression x {comp_per_launch} compressions/launch",
        "launch_ms": leaf_avg_ms, "launches_timed": n_leaf,
        "peak_source": "measured live: ahv_int_peak mode 10 (SHF+LOP3 on the ALU pipe co-issued with IMAD on the FMA "
                       "pipe = every pipe that executes 32-bit integer work)",
        "binding_pipe": {"pipe": "alu", "peak_tinstr": peak_alu / 1e12,
                         "alu_only_instr_per_launch": alu_only,
                         "frac": alu_only / (leaf_avg_ms * 1e-3) / peak_alu,
                         "note": "shift/logic instructions that cannot leave the ALU pipe, over its measured peak; "
                                 "additions run as IMAD on the FMA pipe"},
        "alu_pipe_peak_tinstr": peak_alu / 1e12, "frac_of_alu_pipe": achieved / peak_alu,
        # whole-compression microbenchmarks (register-only, no loads): comparison points, NOT upper bounds - the leaf
        # kernel's fifth block is a constant-schedule padding block that costs fewer instructions than these loops' blocks
        "compress_microbench_native_tinstr": peak_native / 1e12, "compress_microbench_scaled_tinstr": peak_scaled / 1e12,
        "vs_compress_microbench": achieved / peak_scaled,
        "hbm": {"achieved": hbm_bytes / (leaf_avg_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": hbm_bytes / (leaf_avg_ms * 1e-3) / 1e9 / hbm_peak,
                "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback"},
    }
    if roofline["frac"] > 1.0 or roofline["binding_pipe"]["frac"] > 1.0:
        roofline["inconsistent"] = "a fraction above 1: the peak microbenchmark is not an upper bound for this kernel"

    # ---- CPU baseline beside it (oracle port on the host cores, bounded sample) + parity check
    cpu = None
    configs = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        Sc = min(S, max(cores, min(16 * cores, 1024)))
        sample = np.array(h_arena[: Sc * L * FL])
        par_hex = []
        hp = h_par[: Sc * L * 32].reshape(Sc, L, 32)
        for s in range(Sc):
            par_hex.append([None] + [hp[s, i].tobytes().hex() for i in range(1, L)])
        with cpu_pool(sample, L, FL, par_hex, cores) as pool:
            cpu_pipeline(min(Sc, cores), cores, pool)
            dt, roots_cpu, ok_cpu = cpu_pipeline(Sc, cores, pool)
        gpu_roots = [h_roots[32 * s: 32 * s + 32].tobytes().hex() for s in range(Sc)]
        verified = verified and ok_cpu and roots_cpu == gpu_roots
        cpu = {"value": Sc * L / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {Sc} of the GPU run's sessions ({Sc * L} deltas), hashlib arena-level port of "
                         f"delta.py:63,117-152, {cores} processes; roots compared with the GPU's"}
    if world == 1 and not args.no_configs:
        for k_ in ("arena", "digs", "par", "kind"):           # make room: config 3 needs 30 GB + digests
            ctx.dev_free(d.pop(k_))
        configs = extra_configs(ctx, peaks, peak_dual, check=not args.no_cpu_baseline)
        configs["chain_capture"] = {
            "workload": "configs[1] arena: 10^4 sessions x 1000 x 256 B chained on the device (audit/delta.py:101-114), one lane per session",
            "ms": chain_ms, "deltas_per_s": n / chain_ms * 1e3,
            "verified": "the chained arena is the input of the headline commit + verify step (all links verify)"}
        verified = verified and bool(configs["var_len"]["verified"]) and bool(configs["sigma_eff"].get("verified", True))

    seconds = ms_total * 1e-3
    value = world * n * args.steps / seconds
    e2e_value = world * n * e2e_steps / e2e_s
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": "configs[1]: 10k sessions x 1k fixed-size 256 B deltas per GPU, commit + verify",
                   "sessions_per_gpu": S, "leaves_per_session": L, "delta_bytes": FL,
                   "l2_note": "inputs (2.56 GB arena per GPU) exceed the 126 MB L2; no flush needed",
                   "sharding": ("sessions sharded by rank; roots + verdicts collected by ONE ncclAllGather per step, issued "
                                "inside libahv_b200.so (ahv_gather_results_dev)") if world > 1 else "single GPU"},
        "sha256_message_GBps": world * 2 * n * FL * args.steps / seconds / 1e9,
        "compressions_per_s": world * 2 * (leaf_c + node_c) * args.steps / seconds,
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "api": "ahv_audit_sessions (host-pointer C ABI, pinned host buffers; commit pass + verify pass on one upload, pipelined in session chunks)",
                "upload_GBps_all_ranks": world * h2d * e2e_steps / e2e_s / 1e9,
                "h2d_ceiling_gbs": world * arena_bytes * e2e_steps / copy_s / 1e9,
                "h2d_ceiling_note": "the same pinned arena copied host -> device with nothing else running, all ranks at once",
                "frac_of_h2d_ceiling": (h2d / e2e_s) / (arena_bytes / copy_s),
                # on boxes whose host side is one shared resource for both directions (8 ranks: uploads and the digests coming
                # back do not overlap for free) the bytes of BOTH directions are what the ceiling has to carry
                "both_directions_over_h2d_ceiling": ((h2d + d2h) / e2e_s) / (arena_bytes / copy_s),
                "both_directions_note": "a ratio, not a fraction: above 1 where uploads and downloads overlap on a full-duplex link (one rank)",
                "without_digests": {"value": world * n * e2e_steps / e2e_nodig_s, "d2h_bytes_per_step": d2h - 32 * n,
                                    "note": "out_digests = NULL: roots, verdicts, first-bad indices and root matches only"}},
        "comm": ctx.comm_info(), "host": host, "configs": configs,
        "verified": verified, "device": info,
    }
    emit_line(line)
    if world > 1:
        ctx.barrier()
        ctx.comm_destroy()
    return 0 if verified else 1


_RESULT_OUT = None


def claim_stdout() -> None:
    """Keep the real stdout for the one JSON line; anything a library prints to fd 1 from here on (NCCL's
    version banner at NCCL_DEBUG=VERSION, for one) lands on stderr instead."""
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit_line(line: dict) -> None:
    out = _RESULT_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--sessions", type=int, default=SESSIONS_PER_GPU, help="sessions per GPU (default: configs[1])")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs[2]/configs[3] legs (variable-length deltas, sigma_eff)")
    ap.add_argument("--no-affinity", action="store_true", help="do not pin the process to the cores next to its GPU")
    ap.add_argument("--ref-sessions", type=int, default=0, help="--impl reference: sessions per step (default: bounded by the core count)")
    ap.add_argument("--no-reference-own", action="store_true", help="--impl reference: skip the unmodified reference's own runs")
    ap.add_argument("--no-clocks", action="store_true", help="diagnostic: do not poll nvidia-smi during the timed region")
    ap.add_argument("--no-gather", action="store_true", help="diagnostic: skip the per-step all-gather")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
