"""VouchingEngine with the reference's API (``liability/vouching.py:41-230``).

The mutating, per-edge protocol (``vouch``, cycle check, release) stays host Python as in
the reference; ``compute_sigma_eff`` - the arithmetic on the path - runs on the GPU over a
CSR of the vouch records (``ahv_sigma_eff_rings``).  ``compute_sigma_eff_batch`` evaluates
every agent of a session in one launch instead of one O(records) scan per agent.
"""

from __future__ import annotations

import uuid
from dataclasses import dataclass, field
from datetime import datetime, timezone
from typing import Optional, Sequence

import numpy as np

from .. import _native as nat

_EPOCH = datetime(1970, 1, 1, tzinfo=timezone.utc)


def _to_us(t: datetime) -> int:
    """Exact microseconds since the epoch (datetime resolution)."""
    d = t - _EPOCH
    return (d.days * 86400 + d.seconds) * 1_000_000 + d.microseconds


@dataclass
class VouchRecord:
    """vouching.py:19-38."""

    vouch_id: str
    voucher_did: str
    vouchee_did: str
    session_id: str
    bonded_sigma_pct: float
    bonded_amount: float
    created_at: datetime = field(default_factory=lambda: datetime.now(timezone.utc))
    expiry: Optional[datetime] = None
    is_active: bool = True
    released_at: Optional[datetime] = None

    @property
    def is_expired(self) -> bool:
        if self.expiry is None:
            return False
        return datetime.now(timezone.utc) > self.expiry


class VouchingError(Exception):
    """Raised for vouching protocol violations (vouching.py:233)."""


@dataclass
class VouchGraphCSR:
    """Vouch records of ONE session grouped by vouchee, rows in insertion order."""

    agents: list            # row index -> agent DID
    rowptr: np.ndarray      # uint32[A+1]
    bonded: np.ndarray      # float64[E]
    active: np.ndarray      # uint8[E]
    expiry: np.ndarray      # int64[E] microseconds since epoch, EXPIRY_NONE when no expiry

    @staticmethod
    def from_records(records, session_id: str, agents: Sequence[str]) -> "VouchGraphCSR":
        index: dict = {}
        for i, a in enumerate(agents):                  # the same DID may be asked for more than once
            index.setdefault(a, []).append(i)
        rows: list[list] = [[] for _ in agents]
        for v in records:                               # dict order == insertion order (vouching.py:190)
            if v.session_id == session_id:
                for r in index.get(v.vouchee_did, ()):
                    rows[r].append(v)
        rowptr = np.zeros(len(agents) + 1, dtype=np.uint32)
        np.cumsum([len(r) for r in rows], out=rowptr[1:])
        E = int(rowptr[-1])
        bonded = np.empty(E, dtype=np.float64)
        active = np.empty(E, dtype=np.uint8)
        expiry = np.empty(E, dtype=np.int64)
        k = 0
        for r in rows:
            for v in r:
                bonded[k] = v.bonded_amount
                active[k] = 1 if v.is_active else 0
                expiry[k] = nat.EXPIRY_NONE if v.expiry is None else _to_us(v.expiry)
                k += 1
        return VouchGraphCSR(list(agents), rowptr, bonded, active, expiry)


class VouchingEngine:
    """Joint-liability vouching/bonding (vouching.py:41-230)."""

    SCORE_SCALE = 1000.0
    MIN_VOUCHER_SCORE = 0.50
    DEFAULT_BOND_PCT = 0.20
    DEFAULT_MAX_EXPOSURE = 0.80

    # The drop-in class returns the reference's own bits: every row is summed left to right
    # (vouching.py:146-148).  Set False to let rows of more than 64 records take the faster fixed-shape
    # warp sum of ahv_sigma_eff_rings (a few ulp away, inside the 1e-6 tolerance; same value every run).
    exact_order = True

    def __init__(self, max_exposure: Optional[float] = None) -> None:
        self._vouches: dict[str, VouchRecord] = {}
        self.max_exposure = max_exposure or self.DEFAULT_MAX_EXPOSURE

    # ---- host-side protocol (per edge, mutating: same rules as the reference) ----
    def vouch(self, voucher_did: str, vouchee_did: str, session_id: str, voucher_sigma: float,
              bond_pct: Optional[float] = None, expiry: Optional[datetime] = None) -> VouchRecord:
        if voucher_did == vouchee_did:
            raise VouchingError("Cannot vouch for yourself")
        if voucher_sigma < self.MIN_VOUCHER_SCORE:
            raise VouchingError(
                f"Voucher σ ({voucher_sigma:.2f}) below minimum ({self.MIN_VOUCHER_SCORE:.2f})")
        if self._creates_cycle(voucher_did, vouchee_did, session_id):
            raise VouchingError(
                f"Circular vouching detected: {vouchee_did} already vouches for "
                f"{voucher_did} in session {session_id}")
        pct = self.DEFAULT_BOND_PCT if bond_pct is None else bond_pct
        pct = max(0.0, min(1.0, pct))
        bonded = voucher_sigma * pct
        exposure = self.get_total_exposure(voucher_did, session_id)
        limit = voucher_sigma * self.max_exposure
        if exposure + bonded > limit:
            raise VouchingError(
                f"Voucher {voucher_did} would exceed max exposure "
                f"({self.max_exposure:.0%} of σ). Current: {exposure:.3f}, "
                f"requested: {bonded:.3f}, limit: {limit:.3f}")
        record = VouchRecord(vouch_id=f"vouch:{uuid.uuid4()}", voucher_did=voucher_did, vouchee_did=vouchee_did,
                             session_id=session_id, bonded_sigma_pct=pct, bonded_amount=bonded, expiry=expiry)
        self._vouches[record.vouch_id] = record
        return record

    def _live(self, v: VouchRecord, session_id: str) -> bool:
        return v.session_id == session_id and v.is_active and not v.is_expired

    def get_vouchers_for(self, agent_did: str, session_id: str) -> list[VouchRecord]:
        return self._active_vouches_for(agent_did, session_id)

    def _active_vouches_for(self, agent_did: str, session_id: str) -> list[VouchRecord]:
        return [v for v in self._vouches.values() if v.vouchee_did == agent_did and self._live(v, session_id)]

    def get_total_exposure(self, voucher_did: str, session_id: str) -> float:
        return sum(v.bonded_amount for v in self._vouches.values()
                   if v.voucher_did == voucher_did and self._live(v, session_id))

    def release_bond(self, vouch_id: str) -> None:
        if vouch_id not in self._vouches:
            raise VouchingError(f"Vouch {vouch_id} not found")
        rec = self._vouches[vouch_id]
        rec.is_active = False
        rec.released_at = datetime.now(timezone.utc)

    def release_session_bonds(self, session_id: str) -> int:
        released = 0
        for v in self._vouches.values():
            if v.session_id == session_id and v.is_active:
                v.is_active = False
                v.released_at = datetime.now(timezone.utc)
                released += 1
        return released

    def _creates_cycle(self, voucher_did: str, vouchee_did: str, session_id: str) -> bool:
        live = [v for v in self._vouches.values() if self._live(v, session_id)]
        if any(v.voucher_did == vouchee_did and v.vouchee_did == voucher_did for v in live):
            return True
        seen = {voucher_did}
        frontier = [vouchee_did]
        while frontier:
            cur = frontier.pop(0)
            if cur in seen:
                return True
            seen.add(cur)
            frontier.extend(v.vouchee_did for v in live if v.voucher_did == cur and v.vouchee_did not in seen)
        return False

    # ---- the arithmetic on the path: GPU ----------------------------------------
    def compute_sigma_eff(self, vouchee_did: str, session_id: str, vouchee_sigma: float,
                          risk_weight: float) -> float:
        """sigma_eff = sigma_L + (omega * sum of live bonded amounts), capped at 1.0
        (vouching.py:128-151), evaluated by ahv_sigma_eff_rings on one CSR row."""
        out = self.compute_sigma_eff_batch(session_id, [vouchee_did], [vouchee_sigma], [risk_weight])
        return float(out["sigma_eff"][0])

    def get_total_exposure_batch(self, voucher_dids: Sequence[str], session_id: str, now: Optional[datetime] = None,
                                 ctx: Optional[nat.Context] = None) -> np.ndarray:
        """get_total_exposure (vouching.py:157-166) for many vouchers in one launch."""
        ctx = ctx or nat.get_context()
        index = {v: i for i, v in enumerate(voucher_dids)}
        rows: list[list] = [[] for _ in voucher_dids]
        for v in self._vouches.values():
            if v.session_id == session_id:
                r = index.get(v.voucher_did)
                if r is not None:
                    rows[r].append(v)
        rowptr = np.zeros(len(voucher_dids) + 1, dtype=np.uint32)
        np.cumsum([len(r) for r in rows], out=rowptr[1:])
        flat = [v for r in rows for v in r]
        bonded = np.array([v.bonded_amount for v in flat], dtype=np.float64)
        active = np.array([1 if v.is_active else 0 for v in flat], dtype=np.uint8)
        expiry = np.array([nat.EXPIRY_NONE if v.expiry is None else _to_us(v.expiry) for v in flat], dtype=np.int64)
        return ctx.exposure_sums(rowptr, bonded, active, expiry, _to_us(now or datetime.now(timezone.utc)))

    def build_csr(self, session_id: str, agents: Sequence[str]) -> VouchGraphCSR:
        return VouchGraphCSR.from_records(self._vouches.values(), session_id, agents)

    def compute_sigma_eff_batch(self, session_id: str, agents: Sequence[str], sigmas: Sequence[float],
                                risk_weights: Sequence[float], has_consensus: Optional[Sequence[bool]] = None,
                                current_rings: Optional[Sequence[int]] = None, now: Optional[datetime] = None,
                                ctx: Optional[nat.Context] = None) -> dict:
        """All agents of a session in one launch.  Returns sigma_eff, ring, demote and the list of
        agents whose sigma_eff lies within 1e-6 (relative) of a ring threshold."""
        ctx = ctx or nat.get_context()
        g = self.build_csr(session_id, agents)
        A = len(agents)
        cons = np.zeros(A, np.uint8) if has_consensus is None else np.asarray(has_consensus, dtype=np.uint8)
        ring_in = np.full(A, 3, np.uint8) if current_rings is None else np.asarray(current_rings, dtype=np.uint8)
        now_us = _to_us(now or datetime.now(timezone.utc))
        se, ring, demote, near = ctx.sigma_eff_rings(g.rowptr, g.bonded, g.active, g.expiry, now_us,
                                                     np.asarray(sigmas, dtype=np.float64),
                                                     np.asarray(risk_weights, dtype=np.float64), cons, ring_in,
                                                     exact=self.exact_order)
        return {"sigma_eff": se, "ring": ring, "demote": demote,
                "near_threshold": [agents[i] for i in np.nonzero(near)[0]]}

    def compute_sigma_eff_many(self, requests: Sequence[tuple], now: Optional[datetime] = None,
                               ctx: Optional[nat.Context] = None) -> np.ndarray:
        """``[compute_sigma_eff(vouchee, session, sigma, omega) for ...]`` (vouching.py:128-151) for requests
        that may name different sessions, in one launch: row r of the CSR holds the records whose
        (vouchee, session) is request r's, in insertion order.  requests: (vouchee_did, session_id,
        vouchee_sigma, risk_weight)."""
        ctx = ctx or nat.get_context()
        index: dict = {}
        for r, (did, sid, _, _) in enumerate(requests):
            index.setdefault((did, sid), []).append(r)
        rows: list[list] = [[] for _ in requests]
        for v in self._vouches.values():                  # dict order == insertion order
            for r in index.get((v.vouchee_did, v.session_id), ()):
                rows[r].append(v)
        rowptr = np.zeros(len(requests) + 1, dtype=np.uint32)
        if requests:
            np.cumsum([len(r) for r in rows], out=rowptr[1:])
        flat = [v for r in rows for v in r]
        bonded = np.array([v.bonded_amount for v in flat], dtype=np.float64)
        active = np.array([1 if v.is_active else 0 for v in flat], dtype=np.uint8)
        expiry = np.array([nat.EXPIRY_NONE if v.expiry is None else _to_us(v.expiry) for v in flat], dtype=np.int64)
        A = len(requests)
        se, _, _, _ = ctx.sigma_eff_rings(rowptr, bonded, active, expiry, _to_us(now or datetime.now(timezone.utc)),
                                          np.array([q[2] for q in requests], dtype=np.float64),
                                          np.array([q[3] for q in requests], dtype=np.float64),
                                          np.zeros(A, np.uint8), np.full(A, 3, np.uint8), exact=self.exact_order)
        return se
