"""Synthetic workloads of the BASELINE.json configs (bench + full-size tests).

Builds inputs only; never imports ``oracle``.  Pre-images come from the product's own host
packer (``agent_hypervisor_b200.audit.packing``), so they are byte-identical to what the
reference would serialise for the same field values (checked by tests against golden vectors).
"""

from __future__ import annotations

from datetime import datetime, timedelta, timezone

import numpy as np

from agent_hypervisor_b200.audit.delta import SemanticDelta
from agent_hypervisor_b200.audit.packing import canonical_preimage

T0 = datetime(2026, 1, 1, 0, 0, 0, tzinfo=timezone.utc)
NO_PARENT = 0xFFFFFFFF


def config2_templates(L: int = 1000, target_len: int = 256):
    """SURVEY 8d config 2 rows for one session with session id and parent hash left as
    placeholders.  Returns (rows u8[L,target_len], pos_sid i64[L], pos_par u32[L])."""
    rows = np.zeros((L, target_len), dtype=np.uint8)
    pos_sid = np.zeros(L, dtype=np.int64)
    pos_par = np.full(L, NO_PARENT, dtype=np.uint32)
    sid = "session:00000000"
    for i in range(L):
        parent = None if i == 0 else "0" * 64

        def mk(pad):
            return SemanticDelta(delta_id=f"delta:{i+1}", turn_id=i + 1, session_id=sid,
                                 agent_did="did:mesh:a" + "x" * pad, timestamp=T0 + timedelta(seconds=i),
                                 changes=[], parent_hash=parent)

        pad = target_len - len(canonical_preimage(mk(0)))
        if pad < 0:
            raise ValueError(f"delta {i} does not fit in {target_len} bytes")
        pre = canonical_preimage(mk(pad))
        assert len(pre) == target_len
        rows[i] = np.frombuffer(pre, dtype=np.uint8)
        pos_sid[i] = pre.index(b'"session_id": "session:') + len(b'"session_id": "session:')
        if i:
            pos_par[i] = pre.index(b'"parent_hash": "') + len(b'"parent_hash": "')
    return rows, pos_sid, pos_par


def config2_fill(out: np.ndarray, first_session: int, S: int, L: int = 1000, target_len: int = 256,
                 templates=None):
    """Fill ``out`` (u8[S*L*target_len], e.g. a pinned buffer) with the config-2 arena for sessions
    first_session .. first_session+S-1, parent hashes still placeholders.  Returns
    (offsets u64[n+1], hexpos u32[n], session_rowptr u64[S+1])."""
    rows, pos_sid, pos_par = templates or config2_templates(L, target_len)
    a = out[: S * L * target_len].reshape(S, L, target_len)
    a[:] = rows[None, :, :]
    sids = np.arange(first_session, first_session + S, dtype=np.int64)
    li = np.arange(L)
    for k in range(8):
        digit = (sids // 10 ** (7 - k)) % 10 + ord("0")
        a[:, li, pos_sid + k] = digit[:, None].astype(np.uint8)
    n = S * L
    offsets = np.arange(n + 1, dtype=np.uint64) * np.uint64(target_len)
    hexpos = np.tile(pos_par, S)
    rowptr = np.arange(S + 1, dtype=np.uint64) * np.uint64(L)
    return offsets, hexpos, rowptr


def philox_like_lengths(n: int, seed: int, lo: int = 64, hi: int = 16384) -> np.ndarray:
    """Config 3 lengths: log-uniform integers in [lo, hi] (deterministic per seed)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    u = rng.random(n)
    return np.minimum(np.floor(lo * (hi / lo) ** u), hi).astype(np.uint64)


def config3_arena(n: int, seed: int = 0xA6E27, lo: int = 64, hi: int = 16384):
    """Variable-length messages of pseudo-random bytes.  Returns (arena, offsets)."""
    lens = philox_like_lengths(n, seed, lo, hi)
    offsets = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=offsets[1:])
    rng = np.random.Generator(np.random.Philox(key=seed + 1))
    total = int(offsets[-1])
    arena = rng.integers(0, 256, size=total, dtype=np.uint8)
    return arena, offsets


def config4_graph(A: int, E: int, seed: int = 0x51C3A, now_us: int = 1_800_000_000_000_000):
    """Zipf-like in-degree vouch graph as CSR + per-agent inputs."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    u = rng.random(A)
    deg = np.minimum(np.floor(1.0 / (1.0 - u) ** 0.7), 4096.0)
    deg = np.floor(deg * (E / deg.sum())).astype(np.int64)
    short = E - int(deg.sum())
    if short > 0:
        np.add.at(deg, rng.integers(0, A, short), 1)
    rowptr = np.zeros(A + 1, dtype=np.uint32)
    np.cumsum(deg, out=rowptr[1:])
    E = int(rowptr[-1])
    bonded = rng.uniform(0.5, 1.0, E) * rng.uniform(0.0, 0.8, E)       # voucher_sigma * pct (vouching.py:105)
    active = (rng.random(E) < 0.9).astype(np.uint8)
    expiry = np.full(E, np.iinfo(np.int64).max, dtype=np.int64)
    r = rng.random(E)
    expiry[r < 0.05] = now_us - 1_000_000
    expiry[(r >= 0.05) & (r < 0.15)] = now_us + 3_600_000_000
    sigma = rng.random(A)
    omega = rng.choice(np.array([0.2, 0.65, 0.95]), A)
    consensus = (rng.random(A) < 0.1).astype(np.uint8)
    current_ring = rng.integers(1, 4, A).astype(np.uint8)
    return dict(rowptr=rowptr, bonded=bonded, active=active, expiry=expiry, now=now_us, sigma=sigma,
                omega=omega, consensus=consensus, current_ring=current_ring)


def config4_bytes(A: int, E: int) -> int:
    """Algorithmic bytes of one sigma_eff pass."""
    return E * (8 + 1 + 8) + A * (4 + 8 + 8 + 1 + 1 + 8 + 1 + 1)


def count_compressions_fixed(n: int, msg_len: int, leaves_per_session: int) -> tuple[int, int]:
    """(leaf compressions, node compressions) for n fixed-size leaves in uniform sessions."""
    leaf = n * ((msg_len + 9 + 63) // 64)
    nodes, c = 0, leaves_per_session
    while c > 1:
        c = (c + 1) // 2
        nodes += c
    return leaf, 3 * nodes * (n // leaves_per_session)


def slash_sessions(S: int, agents: int = 16, records: int = 48, seed: int = 0x51A5, now_us: int = 1_800_000_000_000_000,
                   omega: float = 0.95):
    """S independent sessions for the slashing pass: ``agents`` participants and ``records``
    vouch records each (uniform voucher / vouchee, 90 % active, 5 % expired), one violation per session at the
    risk weight core.py:271 uses.  Returns the arrays of ``ahv_slash_batch`` (host side)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    A, E = S * agents, S * records
    base = np.repeat(np.arange(S, dtype=np.int64) * agents, records)
    vouchee = base + rng.integers(0, agents, E)
    voucher = base + rng.integers(0, agents, E)
    order = np.argsort(vouchee, kind="stable")
    vouchee, voucher = vouchee[order], voucher[order]
    rowptr = np.zeros(A + 1, dtype=np.int64)
    np.add.at(rowptr, vouchee + 1, 1)
    rowptr = np.cumsum(rowptr).astype(np.uint32)
    expiry = np.full(E, np.iinfo(np.int64).max, dtype=np.int64)
    expiry[rng.random(E) < 0.05] = now_us - 1_000_000
    return dict(
        rowptr=rowptr, voucher=voucher.astype(np.uint32), active=(rng.random(E) < 0.9).astype(np.uint8), expiry=expiry,
        now=now_us, session_agent_rowptr=np.arange(S + 1, dtype=np.uint64) * np.uint64(agents),
        event_rowptr=np.arange(S + 1, dtype=np.uint64),
        event_agent=(np.arange(S, dtype=np.int64) * agents + rng.integers(0, agents, S)).astype(np.uint32),
        event_sigma=rng.random(S), event_omega=np.full(S, omega), scores=rng.random(A))


def slash_bytes(b: dict, n_slash: int, n_clip: int) -> int:
    """Algorithmic bytes of one slashing pass: per slash call its row bounds, the vouchee's score and its
    record (8 + 8 + 32); per record looked at (every record of a slashed row) voucher + active + expiry
    (13); per clip the score read and written, the released flag and the clip record (16 + 1 + 24);
    per session its bounds, event and counters (8 + 8 + 4 + 8 + 8 + 9)."""
    S = len(b["session_agent_rowptr"]) - 1
    return n_slash * 48 + n_clip * (13 + 41) + S * 45
