"""RingEnforcer with the reference's API (``rings/enforcer.py:28-137``), evaluated by the
CUDA gate kernels.  Single calls go through the same batched entry points with n = 1;
``check_batch`` / ``classify_rings`` are the batched forms.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from .. import _native as nat
from ..models import ActionDescriptor, ExecutionRing


@dataclass
class RingCheckResult:
    """rings/enforcer.py:15-25."""

    allowed: bool
    required_ring: ExecutionRing
    agent_ring: ExecutionRing
    sigma_eff: float
    reason: str
    requires_consensus: bool = False
    requires_sre_witness: bool = False


def classify_rings(sigma_eff: Sequence[float], has_consensus: Sequence[bool], current_ring: Sequence[int],
                   ctx: Optional[nat.Context] = None):
    """(ring, should_demote) arrays: models.py:34-42 and enforcer.py:134-137 on the device.

    Uses ahv_sigma_eff_rings with empty vouch rows and omega = 0, for which
    sigma_eff + 0*0 is the input itself (then min(x, 1.0), which does not move a threshold test).
    """
    ctx = ctx or nat.get_context()
    n = len(sigma_eff)
    rowptr = np.zeros(n + 1, dtype=np.uint32)
    z = np.zeros(0)
    _, ring, demote, _ = ctx.sigma_eff_rings(rowptr, z, z.astype(np.uint8), z.astype(np.int64), 0,
                                             np.asarray(sigma_eff, dtype=np.float64), np.zeros(n),
                                             np.asarray(has_consensus, dtype=np.uint8),
                                             np.asarray(current_ring, dtype=np.uint8))
    return ring, demote


def required_rings(action_flags: Sequence[int], ctx: Optional[nat.Context] = None) -> np.ndarray:
    ctx = ctx or nat.get_context()
    n = len(action_flags)
    z8 = np.zeros(n, dtype=np.uint8)
    _, _, req = ctx.ring_check(z8, np.asarray(action_flags, dtype=np.uint8), np.zeros(n), z8, z8)
    return req


def check_batch(agent_ring, action_flags, sigma_eff, has_consensus, has_sre_witness,
                ctx: Optional[nat.Context] = None):
    """Batched RingEnforcer.check: returns (allowed u8[n], reason_code u8[n], required u8[n])."""
    ctx = ctx or nat.get_context()
    return ctx.ring_check(agent_ring, action_flags, sigma_eff, has_consensus, has_sre_witness)


def _flags_of(action) -> int:
    """AHV_ACT_* bits of an ActionDescriptor - ours or the reference's (models.py:104-132)."""
    f = getattr(action, "action_flags", None)
    if f is not None:
        return int(f)
    rev = getattr(action.reversibility, "value", action.reversibility)
    return ((nat.ACT_ADMIN if action.is_admin else 0) | (nat.ACT_READ_ONLY if action.is_read_only else 0)
            | (nat.ACT_REV_NONE if rev == "none" else 0))


class RingEnforcer:
    """4-ring privilege gate (rings/enforcer.py:28-137).  Works with this package's model types and,
    after ``install()``, with the reference's own ``ExecutionRing`` / ``ActionDescriptor``: rings
    come back as members of the enum class the caller passed in."""

    RING_1_THRESHOLD = 0.95
    RING_2_THRESHOLD = 0.60

    def __init__(self) -> None:
        self._sre_witness_callback: Optional[object] = None

    def _reason_text(self, code: int, agent_ring: int, required: int, sigma_eff: float) -> str:
        # strings of rings/enforcer.py:68,80-83,91,103-106,116-119,127
        if code == nat.REASON_NEEDS_WITNESS:
            return "Ring 0 actions require SRE Witness co-sign"
        if code == nat.REASON_RING1_SIGMA:
            return f"Ring 1 requires σ_eff > {self.RING_1_THRESHOLD}, got {sigma_eff:.3f}"
        if code == nat.REASON_NEEDS_CONSENSUS:
            return "Ring 1 non-reversible actions require consensus"
        if code == nat.REASON_RING2_SIGMA:
            return f"Ring 2 requires σ_eff > {self.RING_2_THRESHOLD}, got {sigma_eff:.3f}"
        if code == nat.REASON_RING_INSUFFICIENT:
            return f"Agent ring {agent_ring} insufficient for required ring {required}"
        return "Access granted"

    def check(self, agent_ring: ExecutionRing, action: ActionDescriptor, sigma_eff: float,
              has_consensus: bool = False, has_sre_witness: bool = False) -> RingCheckResult:
        al, code, req = check_batch([int(agent_ring.value)], [_flags_of(action)], [sigma_eff],
                                    [bool(has_consensus)], [bool(has_sre_witness)])
        code, req = int(code[0]), int(req[0])
        ring_cls = type(agent_ring) if hasattr(type(agent_ring), "RING_0_ROOT") else ExecutionRing
        return RingCheckResult(
            allowed=bool(al[0]), required_ring=ring_cls(req), agent_ring=agent_ring, sigma_eff=sigma_eff,
            reason=self._reason_text(code, int(agent_ring.value), req, sigma_eff),
            requires_consensus=(code == nat.REASON_NEEDS_CONSENSUS),
            requires_sre_witness=(code == nat.REASON_NEEDS_WITNESS))

    def check_many(self, agent_rings, actions: Sequence[ActionDescriptor], sigma_eff, has_consensus,
                   has_sre_witness) -> list[RingCheckResult]:
        al, code, req = check_batch([int(r.value) for r in agent_rings], [_flags_of(a) for a in actions],
                                    sigma_eff, has_consensus, has_sre_witness)
        return [RingCheckResult(bool(al[i]), ExecutionRing(int(req[i])), agent_rings[i], sigma_eff[i],
                                self._reason_text(int(code[i]), int(agent_rings[i].value), int(req[i]), sigma_eff[i]),
                                int(code[i]) == nat.REASON_NEEDS_CONSENSUS, int(code[i]) == nat.REASON_NEEDS_WITNESS)
                for i in range(len(al))]

    def compute_ring(self, sigma_eff: float, has_consensus: bool = False) -> ExecutionRing:
        return ExecutionRing.from_sigma_eff(sigma_eff, has_consensus)

    def should_demote(self, current_ring: ExecutionRing, sigma_eff: float) -> bool:
        _, demote = classify_rings([sigma_eff], [False], [int(current_ring.value)])
        return bool(demote[0])
