from .matrix import LiabilityEdge, LiabilityMatrix
from .slashing import SlashingEngine, SlashRequest, SlashResult, VoucherClip
from .vouching import VouchRecord, VouchingEngine, VouchingError, VouchGraphCSR

__all__ = ["LiabilityEdge", "LiabilityMatrix", "VouchRecord", "VouchingEngine", "VouchingError", "VouchGraphCSR",
           "SlashingEngine", "SlashRequest", "SlashResult", "VoucherClip"]
