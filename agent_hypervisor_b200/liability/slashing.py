"""SlashingEngine with the reference's API (``liability/slashing.py:17-145``).

The penalty arithmetic and the bounded cascade run on the GPU (``ahv_slash_cascade``): the vouch
records of the session become the CSR the sigma_eff pass uses plus a voucher column, the device
walks the cascade in the reference's order and returns the slash / clip records, and this class
replays those records into the caller's ``agent_scores`` dict, the vouching engine
(``release_bond``) and ``history`` - so every observable the reference leaves behind is the same,
in the same order.  ``slash_batch`` does that for many sessions in one launch.
"""

from __future__ import annotations

import uuid
from dataclasses import dataclass, field
from datetime import datetime, timezone
from typing import Optional, Sequence

import numpy as np

from .. import _native as nat
from .vouching import _to_us


@dataclass
class SlashResult:
    """slashing.py:17-30."""

    slash_id: str
    vouchee_did: str
    vouchee_sigma_before: float
    vouchee_sigma_after: float
    voucher_clips: list["VoucherClip"]
    reason: str
    session_id: str
    timestamp: datetime = field(default_factory=lambda: datetime.now(timezone.utc))
    cascade_depth: int = 0


@dataclass
class VoucherClip:
    """slashing.py:33-41."""

    voucher_did: str
    sigma_before: float
    sigma_after: float
    risk_weight: float
    vouch_id: str


@dataclass
class SlashRequest:
    """One ``slash`` call of a batch (same arguments as ``SlashingEngine.slash``)."""

    vouchee_did: str
    session_id: str
    vouchee_sigma: float
    risk_weight: float
    reason: str
    agent_scores: dict
    cascade_depth: int = 0


class _SessionGraph:
    """Vouch records of one session as CSR rows by vouchee (insertion order) + voucher column."""

    def __init__(self, records, session_id: str, extra_dids) -> None:
        recs = [v for v in records if v.session_id == session_id]       # dict order == insertion order
        index: dict = {}
        for d in extra_dids:
            index.setdefault(d, len(index))
        for v in recs:
            index.setdefault(v.vouchee_did, len(index))
            index.setdefault(v.voucher_did, len(index))
        self.index = index
        self.dids = list(index)
        rows: list[list] = [[] for _ in index]
        for v in recs:
            rows[index[v.vouchee_did]].append(v)
        self.records = [v for r in rows for v in r]                      # CSR position -> record
        self.rowptr = np.zeros(len(index) + 1, dtype=np.uint32)
        if index:
            np.cumsum([len(r) for r in rows], out=self.rowptr[1:])
        self.voucher = np.array([index[v.voucher_did] for v in self.records], dtype=np.uint32)
        self.active = np.array([1 if v.is_active else 0 for v in self.records], dtype=np.uint8)
        self.expiry = np.array([nat.EXPIRY_NONE if v.expiry is None else _to_us(v.expiry) for v in self.records],
                               dtype=np.int64)


class SlashingEngine:
    """Joint-liability penalties (slashing.py:44-145)."""

    MAX_CASCADE_DEPTH = 2
    SIGMA_FLOOR = 0.05

    def __init__(self, vouching_engine: object) -> None:
        self._vouching = vouching_engine
        self._slash_history: list[SlashResult] = []

    def slash(self, vouchee_did: str, session_id: str, vouchee_sigma: float, risk_weight: float, reason: str,
              agent_scores: dict, cascade_depth: int = 0) -> SlashResult:
        """One slashing event (slashing.py:62-141); returns the top-level SlashResult, cascades are in
        ``history`` after it."""
        return self.slash_batch([SlashRequest(vouchee_did, session_id, vouchee_sigma, risk_weight, reason,
                                              agent_scores, cascade_depth)])[0]

    def slash_batch(self, requests: Sequence[SlashRequest], now: Optional[datetime] = None,
                    ctx: Optional[nat.Context] = None) -> list[SlashResult]:
        """Many ``slash`` calls in one launch.  Requests of one session run in the order given against the
        same state; different sessions are independent.  All requests of a session must pass the same
        ``agent_scores`` dict (as ``Hypervisor.verify_behavior`` does, core.py:260-275)."""
        ctx = ctx or nat.get_context()
        by_session: dict = {}
        for i, r in enumerate(requests):
            if not 0 <= int(r.cascade_depth) <= 255:
                raise ValueError("cascade_depth must be in 0..255")
            g = by_session.setdefault(r.session_id, {"scores": r.agent_scores, "req": []})
            if g["scores"] is not r.agent_scores:
                raise ValueError(f"requests of session {r.session_id} pass different agent_scores dicts")
            g["req"].append(i)
        if len({id(g["scores"]) for g in by_session.values()}) != len(by_session):
            raise ValueError("two sessions of one batch share an agent_scores dict: they are not independent; "
                             "call slash_batch once per session instead")
        records = list(self._vouching._vouches.values())
        graphs = []
        rowptr = [np.zeros(1, np.uint32)]
        cols = {k: [] for k in ("voucher", "active", "expiry", "scores")}
        srp, erp = [0], [0]
        ev_agent, ev_sigma, ev_omega, ev_depth, ev_req = [], [], [], [], []
        for sid, g in by_session.items():
            reqs = [requests[i] for i in g["req"]]
            sg = _SessionGraph(records, sid, list(g["scores"]) + [r.vouchee_did for r in reqs])
            a0, e0 = srp[-1], int(rowptr[-1][-1])
            rowptr.append(sg.rowptr[1:] + np.uint32(e0))
            cols["voucher"].append(sg.voucher + np.uint32(a0))
            cols["active"].append(sg.active)
            cols["expiry"].append(sg.expiry)
            cols["scores"].append(np.array([g["scores"].get(d, 0.0) for d in sg.dids], dtype=np.float64))
            for i, r in zip(g["req"], reqs):
                ev_agent.append(a0 + sg.index[r.vouchee_did]); ev_sigma.append(r.vouchee_sigma)
                ev_omega.append(r.risk_weight); ev_depth.append(int(r.cascade_depth)); ev_req.append(i)
            srp.append(a0 + len(sg.dids)); erp.append(len(ev_agent))
            graphs.append((sid, g, sg, a0, e0))
        cat = lambda xs, dt: np.concatenate(xs).astype(dt, copy=False) if xs else np.zeros(0, dt)   # noqa: E731
        o = ctx.slash_cascade(
            cat(rowptr, np.uint32), cat(cols["voucher"], np.uint32), cat(cols["active"], np.uint8),
            cat(cols["expiry"], np.int64), _to_us(now or datetime.now(timezone.utc)),
            np.array(srp, np.uint64), np.array(erp, np.uint64), np.array(ev_agent, np.uint32),
            np.array(ev_sigma, np.float64), np.array(ev_omega, np.float64), cat(cols["scores"], np.float64),
            event_depth=np.array(ev_depth, np.uint8))
        if o["status"].any():
            raise nat.NativeError("ahv_slash_cascade refused a session (status %r)" % o["status"].tolist())
        # replay, request by request (the order the reference would have run them in)
        out: list = [None] * len(requests)
        per_event: dict = {}
        made: list[dict] = []
        for s, (sid, g, sg, a0, e0) in enumerate(graphs):
            sl = o["slash_log"][erp[s] + e0:][: int(o["slash_count"][s])]
            made.append({})
            for j, rec in enumerate(sl):                      # a session's records are in history order (:119)
                per_event.setdefault(ev_req[int(rec["event"])], []).append((s, j, rec))
        for i in range(len(requests)):
            req = requests[i]
            for (s, j, rec) in per_event.get(i, ()):
                sid, g, sg, a0, e0 = graphs[s]
                cl = o["clip_log"][e0:]
                scores = g["scores"]
                vouchee = sg.dids[int(rec["agent"]) - a0]
                scores[vouchee] = 0.0                         # :88
                clips = []
                for q in cl[int(rec["clip_begin"]): int(rec["clip_begin"]) + int(rec["clip_count"])]:
                    v = sg.records[int(q["edge"]) - e0]
                    scores[v.voucher_did] = float(q["sigma_after"])       # :97
                    clips.append(VoucherClip(voucher_did=v.voucher_did, sigma_before=float(q["sigma_before"]),
                                             sigma_after=float(q["sigma_after"]), risk_weight=req.risk_weight,
                                             vouch_id=v.vouch_id))
                    self._vouching.release_bond(v.vouch_id)   # :108
                if rec["parent"] < 0:
                    reason, before = req.reason, req.vouchee_sigma
                else:
                    p = made[s][int(rec["parent"])]
                    reason, before = f"Cascade from {p.vouchee_did}: {p.reason}", float(rec["sigma_before"])   # :135
                res = SlashResult(slash_id=f"slash:{uuid.uuid4()}", vouchee_did=vouchee, vouchee_sigma_before=before,
                                  vouchee_sigma_after=0.0, voucher_clips=clips, reason=reason, session_id=sid,
                                  cascade_depth=int(rec["depth"]))
                made[s][j] = res
                self._slash_history.append(res)
                if rec["parent"] < 0:
                    out[i] = res
        return out

    @property
    def history(self) -> list[SlashResult]:
        return list(self._slash_history)
