"""Caller integration: ending many sessions with ONE device pass.

``Hypervisor.terminate_session`` (core.py:192-227) computes one session's Merkle root
(``compute_merkle_root``, :204), records it (``CommitmentEngine.commit``, :206-211), releases the session's
bonds (:214), collects garbage (:217-222) and archives the session (:225) - one session per call, so a server
draining N sessions (api/server.py:281-294 in a loop) makes N tree computations.  ``terminate_sessions`` does the
same steps in the same per-session order with the N trees built by a single launch
(``BatchAuditor.roots_of_engines`` -> ``ahv_merkle_roots``).  It works on the reference's own ``Hypervisor``
object - with the GPU classes installed (``install()``) or with the reference's engines, whose stored
``delta_hash`` strings are all the tree needs.
"""

from __future__ import annotations

from typing import Optional, Sequence

from .audit.batch import BatchAuditor


async def terminate_sessions(hv, session_ids: Sequence[str], auditor: Optional[BatchAuditor] = None) -> list:
    """``[await hv.terminate_session(sid) for sid in session_ids]`` with one tree launch.  Returns the Merkle
    roots (None where auditing is off or the session has no deltas), in the order of ``session_ids``.

    Error behaviour follows the loop it replaces: an unknown session raises ``ValueError`` (core.py:297) and a
    session that cannot be terminated raises whatever ``sso.terminate()`` raises - after every session BEFORE it
    in ``session_ids`` has been terminated, committed and archived, and with the ones after it untouched."""
    auditor = auditor or BatchAuditor()
    managed, failure = [], None
    for sid in session_ids:
        try:
            managed.append((sid, hv._get_session(sid)))       # core.py:199
        except Exception as e:                        # unknown session: the ones before it are still done, then re-raise
            failure = e
            break
    # every root in ONE launch, before any state changes: compute_merkle_root (:204) reads the engines only, so
    # doing it first is invisible to the per-session steps below and a device failure leaves every session as it was
    audited = [(sid, m) for sid, m in managed if m.sso.config.enable_audit]
    roots = dict(zip((sid for sid, _ in audited),
                     auditor.roots_of_engines([m.delta_engine for _, m in audited]))) if audited else {}
    out = []
    for sid, m in managed:
        try:
            m.sso.terminate()                         # :200
        except Exception as e:                        # this session cannot be ended: stop here, like the loop would
            failure = e
            break
        root = roots.get(sid)
        if root:
            hv.commitment.commit(session_id=sid, merkle_root=root,
                                 participant_dids=[p.agent_did for p in m.sso.participants],
                                 delta_count=m.delta_engine.turn_count)                 # :206-211
        hv.vouching.release_session_bonds(sid)                                          # :214
        hv.gc.collect(session_id=sid, vfs=m.sso.vfs if hasattr(m.sso, "vfs") else None,
                      delta_engine=m.delta_engine, delta_count=m.delta_engine.turn_count)   # :217-222
        m.sso.archive()                                                                 # :225
        out.append(root)
    if failure is not None:
        raise failure
    return out


def terminate_sessions_sync(hv, session_ids: Sequence[str], auditor: Optional[BatchAuditor] = None) -> list:
    """``terminate_sessions`` for callers without an event loop (none of its steps awaits anything)."""
    coro = terminate_sessions(hv, session_ids, auditor)
    try:
        coro.send(None)
    except StopIteration as stop:
        return stop.value
    raise RuntimeError("terminate_sessions suspended unexpectedly")
