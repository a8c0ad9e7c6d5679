from .delta import DeltaEngine, SemanticDelta, VFSChange
from .commitment import CommitmentEngine, CommitmentRecord
from .packing import ArenaBuilder, PackedSessions, canonical_preimage, classify_parent
from .batch import BatchAuditor
from .replay import DeviceReplay, ReplayPipeline

__all__ = ["DeltaEngine", "SemanticDelta", "VFSChange", "CommitmentEngine", "CommitmentRecord",
           "ArenaBuilder", "PackedSessions", "canonical_preimage", "classify_parent", "BatchAuditor", "ReplayPipeline", "DeviceReplay"]
