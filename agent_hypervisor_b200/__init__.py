"""agent_hypervisor_b200 - B200-native drop-in for agent-hypervisor's audit/Merkle and
sigma_eff/ring-gating hot path.  Class names and signatures follow
``hypervisor/__init__.py:56-82`` of the reference for the classes on the path.

Importing this package loads nothing native; the first call that needs the GPU loads
``libahv_b200.so`` and raises ``NativeUnavailable`` if it (or a CUDA device) is missing.
"""

from ._native import Context, NativeError, NativeUnavailable, get_context, load_library
from .audit import (ArenaBuilder, BatchAuditor, CommitmentEngine, CommitmentRecord, DeltaEngine, DeviceReplay, PackedSessions,
                    ReplayPipeline, SemanticDelta, VFSChange)
from .liability import (LiabilityEdge, LiabilityMatrix, SlashingEngine, SlashRequest, SlashResult, VoucherClip, VouchGraphCSR,
                        VouchingEngine, VouchingError, VouchRecord)
from .models import ActionDescriptor, ExecutionRing, ReversibilityLevel
from .rings import RingCheckResult, RingEnforcer
from .vfs import content_hashes
from .install import install, uninstall
from .integration import terminate_sessions, terminate_sessions_sync

__version__ = "0.1.0"

__all__ = [
    "Context", "NativeError", "NativeUnavailable", "get_context", "load_library",
    "ArenaBuilder", "BatchAuditor", "CommitmentEngine", "CommitmentRecord", "DeltaEngine", "PackedSessions", "ReplayPipeline", "DeviceReplay",
    "SemanticDelta", "VFSChange", "LiabilityEdge", "LiabilityMatrix", "VouchGraphCSR", "VouchingEngine", "VouchingError", "VouchRecord",
    "SlashingEngine", "SlashRequest", "SlashResult", "VoucherClip",
    "ActionDescriptor", "ExecutionRing", "ReversibilityLevel", "RingCheckResult", "RingEnforcer", "content_hashes", "install", "uninstall", "terminate_sessions", "terminate_sessions_sync",
]
