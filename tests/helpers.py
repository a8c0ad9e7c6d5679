"""Test helpers: turn recipe field dicts into packed arenas the way the
reference would serialise them (via the ORACLE's restatement of
delta.py:43-62), chaining parents with hashlib."""

from __future__ import annotations

import hashlib

import numpy as np

from oracle import audit_oracle as ao


def chain_preimages(fields):
    """[(preimage bytes, parent_hash str|None)] with parent = sha256(prev preimage)."""
    out = []
    parent = None
    for f in fields:
        pre = ao.canonical_preimage(
            f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
            f["timestamp"].isoformat(), f["changes"], parent)
        out.append((pre, parent))
        parent = hashlib.sha256(pre).hexdigest()
    return out


def pack_sessions(sessions):
    """sessions: list of list[(preimage, parent_hash)] -> arena, offsets, parents, kind, rowptr."""
    blobs, offs, par, kind, rp = [], [0], [], [], [0]
    for sess in sessions:
        for pre, parent in sess:
            blobs.append(pre)
            offs.append(offs[-1] + len(pre))
            k, raw = ao.classify_parent(parent)
            kind.append(k)
            par.append(raw)
        rp.append(len(kind))
    arena = np.frombuffer(b"".join(blobs), dtype=np.uint8).copy() if blobs else np.zeros(0, np.uint8)
    parents = np.frombuffer(b"".join(par), dtype=np.uint8).reshape(-1, 32).copy() if par else np.zeros((0, 32), np.uint8)
    return (arena, np.array(offs, dtype=np.uint64), parents,
            np.array(kind, dtype=np.uint8), np.array(rp, dtype=np.uint64))


def config2_session(s: int, L: int = 1000):
    """chained, every pre-image exactly 256 B."""
    import recipes
    out = []
    parent = None
    for i in range(L):
        f0 = recipes.config2_delta(s, i, 0)
        base = ao.canonical_preimage(f0["delta_id"], f0["turn_id"], f0["session_id"], f0["agent_did"],
                                     f0["timestamp"].isoformat(), [], parent)
        pad = recipes.CONFIG2_LEN - len(base)
        f = recipes.config2_delta(s, i, pad)
        pre = ao.canonical_preimage(f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
                                    f["timestamp"].isoformat(), [], parent)
        assert len(pre) == recipes.CONFIG2_LEN
        out.append((pre, parent))
        parent = hashlib.sha256(pre).hexdigest()
    return out


def hexes(digests: np.ndarray):
    return [bytes(d).hex() for d in np.asarray(digests, dtype=np.uint8).reshape(-1, 32)]


# ---- slashing cases (recipes.random_slash_case) as the CSR batch the C ABI takes ------------------
NOW_US = 1_800_000_000_000_000                   # any instant: expiries are +-10 years around it
EXPIRY_NONE = np.iinfo(np.int64).max


def fromhex(s: str) -> float:
    return float("nan") if s == "nan" else float.fromhex(s)


def same_f64(a, b) -> bool:
    """Bit-equal except that any NaN equals any NaN."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return a.shape == b.shape and bool(np.all((a.view(np.uint64) == b.view(np.uint64)) | (np.isnan(a) & np.isnan(b))))


def slash_batch(cases):
    """Concatenate slash cases, one session each.  Returns dict of arrays + ``edge_of`` (CSR position ->
    (case, original edge index)) + ``present`` (agents that had a key in the initial agent_scores)."""
    rowptr = [0]; voucher = []; active = []; expiry = []; sess_rp = [0]; ev_rp = [0]
    ev_agent = []; ev_sigma = []; ev_omega = []; ev_depth = []; scores = []; present = []; edge_of = []
    for ci, c in enumerate(cases):
        base = sess_rp[-1]
        A = c["n_agents"]
        rows = [[] for _ in range(A)]
        for k, (a, b, act, ek) in enumerate(c["edges"]):
            rows[b].append((k, a, act, ek))
        for r in rows:
            for (k, a, act, ek) in r:
                voucher.append(base + a); active.append(1 if act else 0)
                expiry.append(EXPIRY_NONE if ek == 0 else (NOW_US + 315_360_000_000_000 if ek == 1 else NOW_US - 315_360_000_000_000))
                edge_of.append((ci, k))
            rowptr.append(len(voucher))
        sc = np.zeros(A); pr = np.zeros(A, np.uint8)
        for a, v in c["scores"]:
            sc[a] = v; pr[a] = 1
        scores.append(sc); present.append(pr)
        for (v, sig, om, dep) in c["events"]:
            ev_agent.append(base + v); ev_sigma.append(sig); ev_omega.append(om); ev_depth.append(dep)
        ev_rp.append(len(ev_agent))
        sess_rp.append(base + A)
    return dict(
        rowptr=np.array(rowptr, np.uint32), voucher=np.array(voucher, np.uint32), active=np.array(active, np.uint8),
        expiry=np.array(expiry, np.int64), now=NOW_US, session_agent_rowptr=np.array(sess_rp, np.uint64),
        event_rowptr=np.array(ev_rp, np.uint64), event_agent=np.array(ev_agent, np.uint32),
        event_sigma=np.array(ev_sigma, np.float64), event_omega=np.array(ev_omega, np.float64),
        event_depth=np.array(ev_depth, np.uint8),
        scores=np.concatenate(scores) if scores else np.zeros(0), present=np.concatenate(present) if present else np.zeros(0, np.uint8),
        edge_of=edge_of)


def check_slash_against_golden(cases, golden_cases, b, out):
    """out = (scores, touched, active, slash_log, clip_log, n_slash, n_clip) for batch b; golden_cases the
    reference's records for the same cases (tests/golden/slashing_golden.json)."""
    scores, touched, active, slash_log, clip_log, n_slash, n_clip = out
    for s, (c, g) in enumerate(zip(cases, golden_cases)):
        a_lo = int(b["session_agent_rowptr"][s])
        e_base = int(b["rowptr"][a_lo])
        sl = slash_log[int(b["event_rowptr"][s]) + e_base:][: int(n_slash[s])]
        cl = clip_log[e_base:][: int(n_clip[s])]
        hist = g["history"]
        assert len(sl) == len(hist), (s, len(sl), len(hist))
        tops = [j for j in range(len(sl)) if sl[j]["parent"] < 0]
        assert tops == g["top_level"]
        assert [int(sl[j]["depth"]) for j in tops] == [e[3] for e in c["events"]]
        for j, h in enumerate(hist):
            r = sl[j]
            assert int(r["agent"]) - a_lo == h["vouchee"] and int(r["depth"]) == h["depth"]
            assert same_f64(r["sigma_before"], fromhex(h["sigma_before"]))
            if h["reason"] != "golden":      # reason: f"Cascade from {parent vouchee}: {parent reason}" (slashing.py:135)
                ph = hist[int(r["parent"])]
                assert h["reason"] == f"Cascade from did:{ph['vouchee']}: {ph['reason']}"
                assert ph["depth"] == h["depth"] - 1
            else:
                assert int(r["parent"]) == -1
            mine = cl[int(r["clip_begin"]): int(r["clip_begin"]) + int(r["clip_count"])]
            assert len(mine) == len(h["clips"])
            ev = int(r["event"]) - int(b["event_rowptr"][s])
            for q, hc in zip(mine, h["clips"]):
                assert int(q["slash"]) == j
                ci, k = b["edge_of"][int(q["edge"])]
                assert ci == s and k == hc[4] and int(b["voucher"][int(q["edge"])]) - a_lo == hc[0]
                assert same_f64(q["sigma_before"], fromhex(hc[1])) and same_f64(q["sigma_after"], fromhex(hc[2]))
                assert same_f64(c["events"][ev][2], fromhex(hc[3]))
        assert sum(int(r["clip_count"]) for r in sl) == len(cl)
        # final agent_scores: the keys the reference's dict ends with, and their values
        A = c["n_agents"]
        keys = {a for a, _ in g["scores"]}
        for a, v in g["scores"]:
            assert same_f64(scores[a_lo + a], fromhex(v)), (s, a)
        for a in range(A):
            assert (a in keys) == bool(b["present"][a_lo + a] or touched[a_lo + a]), (s, a)
            if a not in keys:
                assert scores[a_lo + a] == 0.0
        # released bonds
        for pos in range(e_base, int(b["rowptr"][a_lo + A])):
            ci, k = b["edge_of"][pos]
            assert int(active[pos]) == g["active"][k], (s, k)


# ---- sigma_eff: the fixed-shape sum the CUDA path uses for rows longer than "sigma_long_cut" -------------
SIGMA_LONG_CUT = 64


def fixed_shape_row_sum(vals: np.ndarray) -> float:
    """csrc/ahv_liability.cu (k_sigma_eff_tiled, PAR): 32 partials over row positions p, p+32, ... added in
    that order, then an xor butterfly over the 32 lanes (strides 16, 8, 4, 2, 1); lane 0's value.
    ``vals`` = the row's contributions (+0.0 for a record that is not live)."""
    d = len(vals)
    pad = (-d) % 32
    m = np.concatenate([np.asarray(vals, np.float64), np.zeros(pad)]).reshape(-1, 32)
    x = np.zeros(32)
    with np.errstate(all="ignore"):
        for r in m:
            x = x + r
        idx = np.arange(32)
        for o in (16, 8, 4, 2, 1):
            x = x + x[idx ^ o]
    return float(x[0])


def sigma_expected_default_path(exact, rowptr, bonded, active, expiry, now, sigma, omega, consensus, current_ring,
                                cut: int = SIGMA_LONG_CUT):
    """What the default (non-exact) CUDA path must return, bit for bit: the oracle's result ``exact``
    (sigma_eff, ring, demote, near) with the rows longer than ``cut`` recomputed in the fixed shape."""
    se, ring, dem, near = (np.array(x, copy=True) for x in exact)
    rp = np.asarray(rowptr, np.int64)
    long_rows = np.nonzero(np.diff(rp) > cut)[0]
    with np.errstate(all="ignore"):
        for a in long_rows:
            sl = slice(rp[a], rp[a + 1])
            live = (np.asarray(active[sl]) != 0) & ~((np.asarray(expiry[sl]) != EXPIRY_NONE) & (now > np.asarray(expiry[sl])))
            vals = np.where(live, np.asarray(bonded[sl], np.float64), 0.0)
            s = np.float64(sigma[a]) + np.float64(omega[a]) * np.float64(fixed_shape_row_sum(vals))
            s = np.float64(1.0) if 1.0 < s else s
            se[a] = s
            ring[a] = 1 if (s > 0.95 and consensus[a]) else (2 if s > 0.60 else 3)
            dem[a] = 1 if (2 if s > 0.60 else 3) > int(current_ring[a]) else 0
            near[a] = 1 if (abs(s - 0.60) <= 1e-6 * 0.60 or abs(s - 0.95) <= 1e-6 * 0.95) else 0
    return se, ring, dem, near, long_rows
