"""LiabilityMatrix with the reference's API (``liability/__init__.py:14-139``): the per-session
edge list of voucher -> vouchee bonds.  Record keeping and the small graph walks stay host Python;
the arithmetic - exposure sums - runs on the GPU (``ahv_exposure_sums``), one row per voucher in
edge-insertion order, and ``to_vouch_csr`` hands the same edges to the sigma_eff pass grouped by
vouchee.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from .. import _native as nat


@dataclass
class LiabilityEdge:
    """liability/__init__.py:14-21."""

    voucher_did: str
    vouchee_did: str
    bonded_amount: float
    vouch_id: str


class LiabilityMatrix:
    def __init__(self, session_id: str) -> None:
        self.session_id = session_id
        self._edges: list[LiabilityEdge] = []

    # ---- record keeping (same behaviour as the reference) --------------------------------
    def add_edge(self, voucher_did: str, vouchee_did: str, bonded_amount: float, vouch_id: str) -> LiabilityEdge:
        edge = LiabilityEdge(voucher_did=voucher_did, vouchee_did=vouchee_did, bonded_amount=bonded_amount,
                             vouch_id=vouch_id)
        self._edges.append(edge)
        return edge

    def remove_edge(self, vouch_id: str) -> None:
        self._edges = [e for e in self._edges if e.vouch_id != vouch_id]

    def who_vouches_for(self, agent_did: str) -> list[LiabilityEdge]:
        return [e for e in self._edges if e.vouchee_did == agent_did]

    def who_is_vouched_by(self, agent_did: str) -> list[LiabilityEdge]:
        return [e for e in self._edges if e.voucher_did == agent_did]

    def clear(self) -> None:
        self._edges.clear()

    @property
    def edges(self) -> list[LiabilityEdge]:
        return list(self._edges)

    # ---- arithmetic on the GPU ------------------------------------------------------------
    def _csr(self, keys: Sequence[str], by: str):
        index = {k: i for i, k in enumerate(keys)}
        rows: list[list[float]] = [[] for _ in keys]
        for e in self._edges:
            r = index.get(getattr(e, by))
            if r is not None:
                rows[r].append(e.bonded_amount)
        rowptr = np.zeros(len(keys) + 1, dtype=np.uint32)
        np.cumsum([len(r) for r in rows], out=rowptr[1:])
        bonded = np.array([b for r in rows for b in r], dtype=np.float64)
        E = len(bonded)
        return rowptr, bonded, np.ones(E, np.uint8), np.full(E, nat.EXPIRY_NONE, np.int64)

    def total_exposure_batch(self, voucher_dids: Sequence[str], ctx: Optional[nat.Context] = None) -> np.ndarray:
        """Sum of bonded amounts over each voucher's out-edges, all vouchers in one launch."""
        ctx = ctx or nat.get_context()
        rowptr, bonded, active, expiry = self._csr(voucher_dids, "voucher_did")
        return ctx.exposure_sums(rowptr, bonded, active, expiry, 0)

    def total_exposure(self, voucher_did: str) -> float:
        """liability/__init__.py:64-66.  (Python's sum() of nothing is the int 0; kept.)"""
        if not any(e.voucher_did == voucher_did for e in self._edges):
            return 0
        return float(self.total_exposure_batch([voucher_did])[0])

    def to_vouch_csr(self, agents: Sequence[str]):
        """(rowptr, bonded, active, expiry) grouped by vouchee, for ahv_sigma_eff_rings."""
        return self._csr(agents, "vouchee_did")

    # ---- small graph walks (host; pointer chasing over a handful of edges) -----------------
    def cascade_path(self, agent_did: str, max_depth: int = 2) -> list[list[str]]:
        """Paths along which slashing ``agent_did`` would cascade (liability/__init__.py:68-77,101-121)."""
        found: list[list[str]] = []

        def walk(path: list[str]) -> None:
            if len(path) > max_depth + 1:
                return
            out = self.who_is_vouched_by(path[-1])
            if not out:
                if len(path) > 1:
                    found.append(list(path))
                return
            for edge in out:
                if edge.vouchee_did not in path:
                    walk(path + [edge.vouchee_did])
            if len(path) > 1:
                found.append(list(path))

        walk([agent_did])
        return found

    def has_cycle(self) -> bool:
        """liability/__init__.py:79-95,123-139 (colouring DFS over the edge list)."""
        succ: dict[str, list[str]] = {}
        nodes: set[str] = set()
        for e in self._edges:
            succ.setdefault(e.voucher_did, []).append(e.vouchee_did)
            nodes.update((e.voucher_did, e.vouchee_did))
        state: dict[str, int] = {}          # 1 = on the current path, 2 = finished
        for root in nodes:
            if root in state:
                continue
            stack = [(root, iter(succ.get(root, ())))]
            state[root] = 1
            while stack:
                node, it = stack[-1]
                nxt = next(it, None)
                if nxt is None:
                    state[node] = 2
                    stack.pop()
                elif state.get(nxt) == 1:
                    return True
                elif nxt not in state:
                    state[nxt] = 1
                    stack.append((nxt, iter(succ.get(nxt, ()))))
        return False
