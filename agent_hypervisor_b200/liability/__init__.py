from .matrix import LiabilityEdge, LiabilityMatrix
from .vouching import VouchRecord, VouchingEngine, VouchingError, VouchGraphCSR

__all__ = ["LiabilityEdge", "LiabilityMatrix", "VouchRecord", "VouchingEngine", "VouchingError", "VouchGraphCSR"]
