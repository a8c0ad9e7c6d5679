from .vouching import VouchRecord, VouchingEngine, VouchingError, VouchGraphCSR

__all__ = ["VouchRecord", "VouchingEngine", "VouchingError", "VouchGraphCSR"]
