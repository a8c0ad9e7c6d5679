"""Deterministic input recipes shared by the golden-vector generator
(``tests/golden/make_golden.py``, which feeds them to the REAL reference) and
by the tests (which feed them to the oracle and to the CUDA path).

Pure Python + hashlib; imports neither the reference, the oracle nor the
product.  Recipes follow SURVEY.md section 8c/8d.
"""

from __future__ import annotations

import hashlib
import random
from datetime import datetime, timedelta, timezone

T0 = datetime(2026, 1, 1, 0, 0, 0, tzinfo=timezone.utc)

SHA_KAT_LENGTHS = [0, 1, 3, 54, 55, 56, 57, 63, 64, 65, 118, 119, 120, 121, 127, 128, 129,
                   183, 184, 191, 192, 255, 256, 257, 300, 511, 512, 1000, 4095, 4096, 16383, 16384]


def kat_message(seed: int, length: int) -> bytes:
    """Deterministic pseudo-random bytes (SHAKE-256 of the seed)."""
    return hashlib.shake_256(f"ahv-kat:{seed}".encode()).digest(length) if length else b""


def change(path, operation, content_hash=None, previous_hash=None, agent_did=None) -> dict:
    return {"path": path, "operation": operation, "content_hash": content_hash,
            "previous_hash": previous_hash, "agent_did": agent_did}


# ---- SURVEY 8c golden session -------------------------------------------------
def golden_session(n: int) -> list[dict]:
    """Field dicts (without parent_hash: chained by whoever consumes them)."""
    return [
        dict(delta_id=f"delta:{i+1}", turn_id=i + 1, session_id="session:golden",
             agent_did="did:mesh:a", timestamp=T0 + timedelta(seconds=i),
             changes=[change(f"/f{i}", "add", f"h{i}")])
        for i in range(n)
    ]


# ---- benchmark-shaped deltas (benchmarks/bench_hypervisor.py:157-162), fixed clock ----
def bench_shape_session(n: int, session_id: str = "bench-session") -> list[dict]:
    return [
        dict(delta_id=f"delta:{i+1}", turn_id=i + 1, session_id=session_id, agent_did="did:mesh:a",
             timestamp=T0 + timedelta(seconds=i, microseconds=(i * 7919) % 1000000 or 1),
             changes=[change(f"/f{i}", "add", f"h{i}")])
        for i in range(n)
    ]


# ---- every pre-image exactly 256 B ---------------------------
def config2_delta(s: int, i: int, pad: int) -> dict:
    return dict(delta_id=f"delta:{i+1}", turn_id=i + 1, session_id=f"session:{s:08d}",
                agent_did="did:mesh:a" + "x" * pad, timestamp=T0 + timedelta(seconds=i), changes=[])


CONFIG2_LEN = 256


# ---- randomly shaped sessions ----------------------------------------------------
_ALPHABETS = [
    "abcdefghijklmnopqrstuvwxyz0123456789:/-_.",
    "ABCXYZ \"\\/\b\f\n\r\t<>&'",          # JSON escapes
    "éüñ中文Ж☃",  # BMP non-ASCII -> \uXXXX
    "\U0001F600\U0001F680\U00010348",        # astral -> surrogate pairs
    "\x00\x01\x1f\x7f",                      # control chars
]


def _rand_str(rng: random.Random, lo=0, hi=24) -> str:
    k = rng.randint(lo, hi)
    alpha = rng.choice(_ALPHABETS) if rng.random() < 0.35 else _ALPHABETS[0]
    return "".join(rng.choice(alpha) for _ in range(k))


def random_session(seed: int) -> list[dict]:
    """0..40 deltas with random field shapes (unicode, escapes, None, long lists)."""
    rng = random.Random(seed)
    n = rng.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 32, 33, 40]) if seed % 7 else rng.randint(0, 40)
    sid = "session:" + _rand_str(rng, 1, 20)
    t = T0 + timedelta(seconds=rng.randint(0, 10**8), microseconds=rng.choice([0, 0, 1, 999999, rng.randint(0, 999999)]))
    out = []
    for i in range(n):
        nch = rng.choice([0, 1, 1, 1, 2, 3, 8]) if rng.random() < 0.95 else rng.randint(20, 60)
        chs = [
            change(
                "/" + _rand_str(rng, 0, 30),
                rng.choice(["add", "modify", "delete", "permission"]),
                rng.choice([None, _rand_str(rng, 0, 16), hashlib.sha256(str(rng.random()).encode()).hexdigest()]),
                rng.choice([None, None, hashlib.sha256(str(rng.random()).encode()).hexdigest()]),
                rng.choice([None, "did:x"]),
            )
            for _ in range(nch)
        ]
        t = t + timedelta(microseconds=rng.randint(0, 5_000_000))
        tz = timezone.utc if rng.random() < 0.9 else timezone(timedelta(hours=rng.choice([-7, 5]), minutes=rng.choice([0, 30])))
        out.append(dict(
            delta_id=(f"delta:{i+1}" if rng.random() < 0.7 else _rand_str(rng, 1, 40)),
            turn_id=i + 1, session_id=sid,
            agent_did="did:mesh:" + _rand_str(rng, 0, 20),
            timestamp=t.astimezone(tz), changes=chs))
    return out


# ---- vouch graphs ----------------------------------------------------------------
def random_vouch_graph(seed: int, n_agents: int, n_edges: int):
    """Insertion-ordered edge list for ONE session, plus per-agent inputs.

    Returns dict with: edges = [(voucher_idx, vouchee_idx, voucher_sigma, pct, active, expiry_kind)]
    where expiry_kind in {0: none, 1: far future, 2: far past}; sigma/omega/consensus/current_ring lists.
    """
    rng = random.Random(seed)
    edges = []
    for _ in range(n_edges):
        a = rng.randrange(n_agents)
        b = rng.randrange(n_agents)
        vs = rng.uniform(0.5, 1.0)
        pct = rng.uniform(0.0, 0.8)
        active = rng.random() < 0.9
        ek = rng.choices([0, 1, 2], weights=[0.7, 0.2, 0.1])[0]
        edges.append((a, b, vs, pct, active, ek))
    sigma = [rng.choice([rng.random(), rng.random(), 0.6, 0.95, 0.0, 1.0, 0.5999999, 0.9500001]) for _ in range(n_agents)]
    omega = [rng.choice([0.2, 0.65, 0.95]) for _ in range(n_agents)]
    consensus = [rng.random() < 0.3 for _ in range(n_agents)]
    current_ring = [rng.choice([1, 2, 3]) for _ in range(n_agents)]
    return dict(edges=edges, sigma=sigma, omega=omega, consensus=consensus, current_ring=current_ring)


# ---- slashing cases --------------------------------------------------------------
SLASH_OMEGAS = [0.95, 0.95, 0.95, 0.5, 0.3, 0.65, 0.2, 0.99, 1.0, 0.0]


def random_slash_case(seed: int) -> dict:
    """One session's vouch records, a score table and a few slash events (all deterministic per seed).

    edges  = [(voucher_idx, vouchee_idx, active, expiry_kind)] in insertion order; duplicates, self-edges
             and cycles are allowed (the records are placed into ``_vouches`` directly, as the reference's
             own tests do); expiry_kind in {0: none, 1: far future, 2: far past}
    scores = [(agent_idx, value)] initial ``agent_scores`` dict in insertion order (some agents missing)
    events = [(vouchee_idx, vouchee_sigma, risk_weight, cascade_depth)] applied in order to the same state
             (cascade_depth is slash()'s last argument: 0 for a normal call)
    """
    rng = random.Random(0x51A5 + seed)
    n_agents = rng.choice([1, 2, 3, 5, 8, 13, 40])
    dense = rng.choice([0.5, 1.5, 4.0])
    n_edges = 0 if seed % 11 == 0 else int(n_agents * dense) + rng.randrange(3)
    hubs = [rng.randrange(n_agents) for _ in range(2)]
    edges = []
    for _ in range(n_edges):
        b = rng.choice(hubs) if rng.random() < 0.4 else rng.randrange(n_agents)
        a = rng.randrange(n_agents)
        if a == b and rng.random() < 0.8:
            a = (a + 1) % n_agents
        edges.append((a, b, rng.random() < 0.9, rng.choices([0, 1, 2], weights=[0.7, 0.2, 0.1])[0]))
    special = [0.05, 0.06, 0.060000000000000005, 0.0600000001, 0.0, 1.0, 1.2, 1.1999999, 0.059999, -0.3,
               float("nan"), 1e-300, 5e-324]
    scores = []
    order = list(range(n_agents))
    rng.shuffle(order)
    for a in order:
        if rng.random() < 0.8:
            scores.append((a, rng.choice(special) if rng.random() < 0.25 else rng.random()))
    events = []
    for _ in range(rng.choice([1, 1, 2, 3])):
        v = rng.choice(hubs) if rng.random() < 0.7 else rng.randrange(n_agents)
        events.append((v, rng.random(), rng.choice(SLASH_OMEGAS), rng.choice([0, 0, 0, 0, 0, 0, 1, 2, 3])))
    return dict(n_agents=n_agents, edges=edges, scores=scores, events=events)
