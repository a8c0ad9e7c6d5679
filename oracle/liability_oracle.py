"""Plain-Python / numpy restatement of the sigma_eff + ring-gating pass
(TEST INFRASTRUCTURE; the product package never imports this module).

Reference lines followed (paths under ``/root/reference/src/hypervisor``):
``liability/vouching.py:34-38,128-151,186-197``, ``models.py:34-42,122-132``,
``rings/enforcer.py:44-137``.

Parity: pinned by ``tests/test_oracle_golden.py`` against the real
``VouchingEngine`` / ``RingEnforcer`` (``tests/golden/make_golden.py``).
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
