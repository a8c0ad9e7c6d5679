"""The model types the hot path needs, with the reference's names and behaviour
(``/root/reference/src/hypervisor/models.py``): ``ExecutionRing`` (:19-42),
``ReversibilityLevel`` (:45-66) and ``ActionDescriptor`` (:104-132).

``ExecutionRing.from_sigma_eff`` classifies on the GPU through the same kernel as the
batched pass (``ahv_sigma_eff_rings``); the enum values and thresholds are unchanged.
"""

from __future__ import annotations

from dataclasses import dataclass
from enum import Enum
from typing import Optional


class ExecutionRing(int, Enum):
    """Ring 0 root / 1 privileged / 2 standard / 3 sandbox (models.py:19-32)."""

    RING_0_ROOT = 0
    RING_1_PRIVILEGED = 1
    RING_2_STANDARD = 2
    RING_3_SANDBOX = 3

    @classmethod
    def from_sigma_eff(cls, sigma_eff: float, has_consensus: bool = False) -> "ExecutionRing":
        """models.py:34-42: >0.95 with consensus -> 1; >0.60 -> 2; else 3 (strict '>')."""
        from .rings.enforcer import classify_rings

        ring, _ = classify_rings([sigma_eff], [has_consensus], [3])
        return cls(int(ring[0]))


class ReversibilityLevel(str, Enum):
    """models.py:45-66."""

    FULL = "full"
    PARTIAL = "partial"
    NONE = "none"

    @property
    def risk_weight_range(self) -> tuple[float, float]:
        return {"full": (0.1, 0.3), "partial": (0.5, 0.8), "none": (0.9, 1.0)}[self.value]

    @property
    def default_risk_weight(self) -> float:
        lo, hi = self.risk_weight_range
        return (lo + hi) / 2


@dataclass
class ActionDescriptor:
    """models.py:104-132 (same fields, same defaults)."""

    action_id: str
    name: str
    execute_api: str
    undo_api: Optional[str] = None
    reversibility: ReversibilityLevel = ReversibilityLevel.NONE
    undo_window_seconds: int = 0
    compensation_method: Optional[str] = None
    is_read_only: bool = False
    is_admin: bool = False

    @property
    def risk_weight(self) -> float:
        return self.reversibility.default_risk_weight

    @property
    def action_flags(self) -> int:
        """Bit encoding handed to the device (AHV_ACT_* in include/ahv.h)."""
        from ._native import ACT_ADMIN, ACT_READ_ONLY, ACT_REV_NONE

        return ((ACT_ADMIN if self.is_admin else 0) | (ACT_READ_ONLY if self.is_read_only else 0)
                | (ACT_REV_NONE if self.reversibility == ReversibilityLevel.NONE else 0))

    @property
    def required_ring(self) -> ExecutionRing:
        """models.py:122-132, evaluated by the device gate kernel (ahv_ring_check)."""
        from .rings.enforcer import required_rings

        return ExecutionRing(int(required_rings([self.action_flags])[0]))
