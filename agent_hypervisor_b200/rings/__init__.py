from .enforcer import RingCheckResult, RingEnforcer, check_batch, classify_rings, required_rings

__all__ = ["RingCheckResult", "RingEnforcer", "check_batch", "classify_rings", "required_rings"]
