"""``install()``: put the GPU classes under an installed ``hypervisor`` package without editing it.

The reference binds the hot-path classes by name at import time (``core.py:16-29``,
``__init__.py:56-82``, ``api/server.py``), so swapping them means rebinding those names in every
module that imported them.  ``install()`` does exactly that for ``DeltaEngine``, ``SemanticDelta``,
``VFSChange``, ``CommitmentEngine``, ``VouchingEngine``, ``VouchRecord``, ``VouchingError``,
``LiabilityMatrix``, ``LiabilityEdge``, ``SlashingEngine``, ``SlashResult``, ``VoucherClip``,
``RingEnforcer`` and ``RingCheckResult``; the model enums
(``ExecutionRing``, ``ActionDescriptor``, ...) stay the reference's own - the GPU classes accept
them as they are.  ``uninstall()`` restores the originals.
"""

from __future__ import annotations

import importlib
import sys

_REPLACEMENTS = {
    "hypervisor.audit.delta": ("DeltaEngine", "SemanticDelta", "VFSChange"),
    "hypervisor.audit.commitment": ("CommitmentEngine", "CommitmentRecord"),
    "hypervisor.liability.vouching": ("VouchingEngine", "VouchRecord", "VouchingError"),
    "hypervisor.liability": ("LiabilityMatrix", "LiabilityEdge"),
    "hypervisor.liability.slashing": ("SlashingEngine", "SlashResult", "VoucherClip"),
    "hypervisor.rings.enforcer": ("RingEnforcer", "RingCheckResult"),
}
_saved: dict = {}


def install(package: str = "hypervisor") -> list[str]:
    """Rebind the hot-path classes in every loaded ``hypervisor`` module.  Returns the names rebound
    (``module.attr``).  Safe to call twice."""
    import agent_hypervisor_b200 as ours

    for mod in _REPLACEMENTS:
        importlib.import_module(mod.replace("hypervisor", package, 1))
    importlib.import_module(package)
    try:
        importlib.import_module(package + ".core")
    except Exception:      # optional dependencies of the orchestrator may be missing
        pass
    originals = {}
    for mod, names in _REPLACEMENTS.items():
        m = sys.modules[mod.replace("hypervisor", package, 1)]
        for n in names:
            cur = getattr(m, n)
            originals[n] = _saved.get((m.__name__, n), (cur,))[0] if getattr(cur, "__module__", "").startswith("agent_hypervisor_b200") else cur
    done = []
    for name, m in list(sys.modules.items()):
        if m is None or not (name == package or name.startswith(package + ".")):
            continue
        for n, orig in originals.items():
            if getattr(m, n, None) is orig:
                _saved.setdefault((name, n), (orig,))
                setattr(m, n, getattr(ours, n))
                done.append(f"{name}.{n}")
    return sorted(done)


def uninstall() -> None:
    for (name, n), (orig,) in list(_saved.items()):
        m = sys.modules.get(name)
        if m is not None:
            setattr(m, n, orig)
    _saved.clear()
