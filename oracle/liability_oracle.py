"""Plain-Python / numpy restatement of the sigma_eff + ring-gating pass
(TEST INFRASTRUCTURE; the product package never imports this module).

Reference lines followed (paths under ``/root/reference/src/hypervisor``):
``liability/vouching.py:34-38,128-151,186-197``, ``models.py:34-42,122-132``,
``rings/enforcer.py:44-137``, ``liability/slashing.py:53-141``.

Parity: pinned by ``tests/test_oracle_golden.py`` against the real
``VouchingEngine`` / ``RingEnforcer`` / ``SlashingEngine`` (``tests/golden/make_golden.py``).
"""

from __future__ import annotations

import numpy as np

EXPIRY_NONE = np.iinfo(np.int64).max  # VouchRecord.expiry is None

RING_1_THRESHOLD = 0.95  # rings/enforcer.py:38
RING_2_THRESHOLD = 0.60  # rings/enforcer.py:39
NEAR_REL_TOL = 1e-6      # BASELINE.json north_star: agents within this of a threshold are listed

# action flag bits (host encoding of ActionDescriptor, models.py:122-132)
ACT_ADMIN = 1
ACT_READ_ONLY = 2
ACT_REV_NONE = 4

# RingEnforcer.check outcome codes (rings/enforcer.py:58-128, in order)
REASON_GRANTED = 0
REASON_NEEDS_WITNESS = 1
REASON_RING1_SIGMA = 2
REASON_NEEDS_CONSENSUS = 3
REASON_RING2_SIGMA = 4
REASON_RING_INSUFFICIENT = 5


def ring_from_sigma(sigma_eff: float, has_consensus: bool) -> int:
    """models.py:34-42 (strict '>'; never returns ring 0)."""
    if sigma_eff > 0.95 and has_consensus:
        return 1
    elif sigma_eff > 0.60:
        return 2
    return 3


def required_ring(flags: int) -> int:
    """models.py:122-132."""
    if flags & ACT_ADMIN:
        return 0
    elif (flags & ACT_REV_NONE) and not (flags & ACT_READ_ONLY):
        return 1
    elif flags & ACT_READ_ONLY:
        return 3
    return 2


def sigma_eff_rings(rowptr, bonded, active, expiry, now, sigma, omega, consensus, current_ring):
    """CSR form of vouching.py:128-151 + models.py:34-42 + enforcer.py:134-137.

    Row r = the vouch records whose vouchee is agent r (one session), in
    insertion order (``_vouches`` dict order, vouching.py:186-197).  A record
    contributes iff ``is_active and not is_expired`` where
    ``is_expired = expiry is not None and now > expiry`` (vouching.py:34-38).
    Sum is sequential fp64 from 0.0 (:146-148); sigma_eff = sigma + (omega*sum)
    (:150); Python ``min(x, 1.0)`` (:151) keeps NaN.
    """
    rowptr = np.asarray(rowptr)
    A = len(rowptr) - 1
    out_sigma = np.empty(A, dtype=np.float64)
    out_ring = np.empty(A, dtype=np.uint8)
    out_demote = np.empty(A, dtype=np.uint8)
    out_near = np.empty(A, dtype=np.uint8)
    for a in range(A):
        contrib = 0.0
        for e in range(int(rowptr[a]), int(rowptr[a + 1])):
            if active[e] and not (expiry[e] != EXPIRY_NONE and now > expiry[e]):
                contrib += float(bonded[e])
        se = float(sigma[a]) + (float(omega[a]) * contrib)
        se = min(se, 1.0)
        out_sigma[a] = se
        out_ring[a] = ring_from_sigma(se, bool(consensus[a]))
        # should_demote computes the ring WITHOUT consensus (enforcer.py:136)
        out_demote[a] = 1 if ring_from_sigma(se, False) > int(current_ring[a]) else 0
        out_near[a] = (
            1
            if (
                abs(se - RING_2_THRESHOLD) <= NEAR_REL_TOL * RING_2_THRESHOLD
                or abs(se - RING_1_THRESHOLD) <= NEAR_REL_TOL * RING_1_THRESHOLD
            )
            else 0
        )
    return out_sigma, out_ring, out_demote, out_near


def ring_check(agent_ring, action_flags, sigma, consensus, witness):
    """rings/enforcer.py:44-128 as (allowed, reason_code, required_ring) arrays."""
    n = len(agent_ring)
    allowed = np.zeros(n, dtype=np.uint8)
    reason = np.zeros(n, dtype=np.uint8)
    req_out = np.zeros(n, dtype=np.uint8)
    for i in range(n):
        req = required_ring(int(action_flags[i]))
        req_out[i] = req
        s = float(sigma[i])
        if req == 0 and not witness[i]:
            reason[i] = REASON_NEEDS_WITNESS
        elif req == 1 and s < RING_1_THRESHOLD:
            reason[i] = REASON_RING1_SIGMA
        elif req == 1 and not consensus[i]:
            reason[i] = REASON_NEEDS_CONSENSUS
        elif req == 2 and s < RING_2_THRESHOLD:
            reason[i] = REASON_RING2_SIGMA
        elif int(agent_ring[i]) > req:
            reason[i] = REASON_RING_INSUFFICIENT
        else:
            reason[i] = REASON_GRANTED
            allowed[i] = 1
    return allowed, reason, req_out


def exposure_sums(rowptr, bonded, active, expiry, now, compensated):
    """Per-row sum of live bonded amounts: VouchingEngine.get_total_exposure (vouching.py:157-166) and
    LiabilityMatrix.total_exposure (liability/__init__.py:64-66), both written with Python's built-in
    sum().  compensated=True restates CPython >= 3.12's float sum (Neumaier, Python/bltinmodule.c);
    False the plain left-to-right additions of CPython <= 3.11."""
    rowptr = np.asarray(rowptr)
    R = len(rowptr) - 1
    out = np.zeros(R, dtype=np.float64)
    for r in range(R):
        f, c = 0.0, 0.0
        for e in range(int(rowptr[r]), int(rowptr[r + 1])):
            if not (active[e] and not (expiry[e] != EXPIRY_NONE and now > expiry[e])):
                continue
            x = float(bonded[e])
            if compensated:
                t = f + x
                if abs(f) >= abs(x):
                    c += (f - t) + x
                else:
                    c += (x - t) + f
                f = t
            else:
                f = f + x
        if compensated and c and np.isfinite(c):
            f += c
        out[r] = f
    return out


# ---- collateral slashing (liability/slashing.py) -----------------------------------------------
SIGMA_FLOOR = 0.05            # slashing.py:54
MAX_CASCADE_DEPTH = 2         # slashing.py:53

SLASH_DT = np.dtype([("agent", "<u4"), ("parent", "<i4"), ("event", "<u4"), ("depth", "<u4"),
                     ("clip_begin", "<u4"), ("clip_count", "<u4"), ("sigma_before", "<f8")])
CLIP_DT = np.dtype([("edge", "<u4"), ("slash", "<u4"), ("sigma_before", "<f8"), ("sigma_after", "<f8")])


def slash_cascade(rowptr, voucher, active, expiry, now, session_agent_rowptr, event_rowptr, event_agent,
                  event_sigma, event_omega, scores, event_depth=None):
    """CSR restatement of SlashingEngine.slash (slashing.py:62-141) for many sessions.

    Rows = vouch records by vouchee in insertion order (get_vouchers_for -> vouching.py:153-155,186-197);
    ``voucher[e]`` = row index of record e's voucher; session s owns agent rows
    ``session_agent_rowptr[s]:[s+1]`` and the events ``event_rowptr[s]:[s+1]``, applied in order.
    ``scores`` is the dense ``agent_scores`` (missing key = 0.0, slashing.py:94).  Returns copies:
    (scores, touched, active, slash_log, clip_log, slash_count[S], clip_count[S]); session s's slash
    records start at ``event_rowptr[s] + rowptr[a_lo]`` and its clip records at ``rowptr[a_lo]``.
    """
    rowptr = np.asarray(rowptr); voucher = np.asarray(voucher)
    active = np.array(active, dtype=np.uint8, copy=True)
    scores = np.array(scores, dtype=np.float64, copy=True)
    touched = np.zeros(len(scores), dtype=np.uint8)
    S = len(session_agent_rowptr) - 1
    E = int(rowptr[-1])
    n_ev = int(event_rowptr[-1])
    slash_log = np.zeros(n_ev + E, dtype=SLASH_DT)
    clip_log = np.zeros(E, dtype=CLIP_DT)
    n_slash = np.zeros(S, dtype=np.uint32)
    n_clip = np.zeros(S, dtype=np.uint32)

    def live(e):      # vouching.py:34-38,191-196
        return bool(active[e]) and not (expiry[e] != EXPIRY_NONE and now > expiry[e])

    for s in range(S):
        a_lo = int(session_agent_rowptr[s])
        e_base = int(rowptr[a_lo])
        sl = slash_log[int(event_rowptr[s]) + e_base:]
        cl = clip_log[e_base:]
        cnt = [0, 0]

        def slash(v, sigma_before, omega, depth, parent, ev):
            scores[v] = 0.0                                          # :88
            touched[v] = 1
            me = cnt[0]; cnt[0] += 1
            c0 = cnt[1]
            for e in range(int(rowptr[v]), int(rowptr[v + 1])):     # :92
                if not live(e):
                    continue
                u = int(voucher[e])
                before = float(scores[u])                            # :94 (.get(..., 0.0): dense zeros)
                new = before * (1.0 - omega)                         # :95
                new = max(new, SIGMA_FLOOR)                          # :96 (Python max keeps a NaN first arg)
                scores[u] = new; touched[u] = 1                      # :97
                cl[cnt[1]] = (e, me, before, new); cnt[1] += 1       # :99-105
                active[e] = 0                                        # :108 release_bond
            c1 = cnt[1]
            sl[me] = (v, parent, ev, depth, c0, c1 - c0, sigma_before)   # :110-119
            if depth < MAX_CASCADE_DEPTH:                            # :122
                for c in range(c0, c1):
                    after = float(cl[c]["sigma_after"])
                    if after < SIGMA_FLOOR + 0.01:                   # :124
                        u = int(voucher[int(cl[c]["edge"])])
                        if any(live(e) for e in range(int(rowptr[u]), int(rowptr[u + 1]))):   # :126-129
                            slash(u, after, omega, depth + 1, me, ev)

        for ev in range(int(event_rowptr[s]), int(event_rowptr[s + 1])):
            slash(int(event_agent[ev]), float(event_sigma[ev]), float(event_omega[ev]),
                  0 if event_depth is None else int(event_depth[ev]), -1, ev)      # cascade_depth argument, :70
        n_slash[s], n_clip[s] = cnt
    return scores, touched, active, slash_log, clip_log, n_slash, n_clip
