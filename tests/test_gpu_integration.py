"""SURVEY.md 8f row 3 on the real thing: the UNMODIFIED reference package (the copy under baseline/_ref, made by
__graft_entry__.build()) runs the session lifecycle of its own integration tests
(tests/integration/test_hypervisor_e2e.py:382-447: create -> join -> activate -> capture -> terminate -> verify)
for 200 sessions twice - once as it is, once with the GPU classes installed and all sessions ended by ONE
batched call - under a frozen clock and fixed session ids.  Roots, commitment records and verdicts must be equal."""

import asyncio
import itertools
import os
import sys
import uuid
from datetime import datetime, timedelta, timezone

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = os.path.join(ROOT, "baseline", "_ref", "src")
N_SESSIONS = 200


class _FrozenClock(datetime):
    """datetime whose now() ticks one second per call from a fixed start."""
    _ticks = itertools.count()

    @classmethod
    def now(cls, tz=None):
        return datetime(2026, 1, 1, tzinfo=timezone.utc) + timedelta(seconds=next(cls._ticks))

    @classmethod
    def reset(cls):
        cls._ticks = itertools.count()


@pytest.fixture()
def reference(monkeypatch):
    if not os.path.isdir(os.path.join(REF_SRC, "hypervisor")):
        pytest.skip("baseline/_ref is not present (it is staged by __graft_entry__.build() where /root/reference exists)")
    monkeypatch.syspath_prepend(REF_SRC)
    import hypervisor
    import hypervisor.audit.delta as ref_delta
    import agent_hypervisor_b200.audit.delta as our_delta
    monkeypatch.setattr(ref_delta, "datetime", _FrozenClock)
    monkeypatch.setattr(our_delta, "datetime", _FrozenClock)
    ids = itertools.count()
    monkeypatch.setattr(uuid, "uuid4", lambda: uuid.UUID(int=next(ids) + 1))
    yield hypervisor, lambda: (ids.__setstate__(0) if hasattr(ids, "__setstate__") else None)
    from agent_hypervisor_b200 import uninstall
    uninstall()


async def _lifecycle(hypervisor, batched: bool):
    """One run of the lifecycle.  Returns per-session records of everything the audit path produced."""
    from hypervisor import Hypervisor, SessionConfig
    from hypervisor.audit.delta import VFSChange
    hv = Hypervisor(max_exposure=1.0)
    sids, engines = [], []
    for s in range(N_SESSIONS):
        session = await hv.create_session(config=SessionConfig(enable_audit=(s % 17 != 5)), creator_did="did:mesh:admin")
        sid = session.sso.session_id
        await hv.join_session(sid, "did:mesh:a", sigma_raw=0.8)
        await hv.join_session(sid, f"did:mesh:b{s % 3}", sigma_raw=0.65)
        await hv.activate_session(sid)
        if s % 4 == 0:
            hv.vouching.vouch("did:mesh:a", f"did:mesh:b{s % 3}", sid, 0.8, bond_pct=0.1)
        n_deltas = [0, 1, 2, 3, 5, 8, 13, 40][s % 8]
        for i in range(n_deltas):
            session.delta_engine.capture(f"did:mesh:agent-{i % 3}",
                                         [VFSChange(path=f"/doc{i}", operation="add", content_hash=f"h{s}-{i}")])
        sids.append(sid); engines.append(session.delta_engine)
    # verify_chain before ending, with one tampered session (test_hypervisor_e2e.py:408-423)
    engines[6]._deltas[5].agent_did = "did:mesh:tampered"
    verdicts = [e.verify_chain() for e in engines]
    if batched:
        from agent_hypervisor_b200 import terminate_sessions
        roots = await terminate_sessions(hv, sids)
    else:
        roots = [await hv.terminate_session(sid) for sid in sids]
    recs = []
    for sid, e, r in zip(sids, engines, roots):
        c = hv.commitment.get_commitment(sid)
        recs.append(dict(sid=sid, root=r, hashes=[d.delta_hash for d in e._deltas], turn=e.turn_count,
                         commit=None if c is None else (c.merkle_root, tuple(c.participant_dids), c.delta_count),
                         verifies=hv.commitment.verify(sid, r) if r else None,
                         state=hv.get_session(sid).sso.state.value,
                         bonds_live=[v.is_active for v in hv.vouching._vouches.values() if v.session_id == sid]))
    return recs, verdicts, type(engines[0]).__module__


def test_lifecycle_with_installed_classes_equals_the_reference(reference, monkeypatch):
    hypervisor, _ = reference
    from agent_hypervisor_b200 import install, uninstall

    def run(batched):
        _FrozenClock.reset()
        ids = itertools.count()
        monkeypatch.setattr(uuid, "uuid4", lambda: uuid.UUID(int=next(ids) + 1))
        return asyncio.run(_lifecycle(hypervisor, batched))

    want, want_verdicts, mod_ref = run(batched=False)
    assert mod_ref.startswith("hypervisor.")                      # the reference's own DeltaEngine did this run
    done = install()
    assert "hypervisor.core.DeltaEngine" in done
    try:
        got, got_verdicts, mod_gpu = run(batched=True)
        assert mod_gpu.startswith("agent_hypervisor_b200.")       # ... and the GPU one this run
        per_session, _, _ = run(batched=False)                    # installed classes, the reference's own loop
    finally:
        uninstall()
    assert got_verdicts == want_verdicts and want_verdicts.count(False) == 1 and not want_verdicts[6]
    assert len(got) == len(want) == N_SESSIONS
    for w, g, p in zip(want, got, per_session):
        assert g == w, (w["sid"], w, g)
        assert p == w, (w["sid"], w, p)
    assert sum(1 for w in want if w["root"]) > 100 and any(w["root"] is None for w in want)
    assert all(w["state"] == "archived" for w in want) and not any(any(w["bonds_live"]) for w in want)

    # the batched call on the UNinstalled reference objects as well (their stored delta_hash strings feed the tree)
    mixed, _, mod = run(batched=True)
    assert mod.startswith("hypervisor.") and mixed == want


def test_terminate_sessions_error_behaviour_and_flush_batch(reference):
    hypervisor, _ = reference
    from agent_hypervisor_b200 import CommitmentEngine, DeltaEngine, VFSChange, terminate_sessions_sync
    from hypervisor import Hypervisor, SessionConfig

    async def make(hv, n):
        out = []
        for s in range(n):
            m = await hv.create_session(config=SessionConfig(), creator_did="did:mesh:admin")
            await hv.join_session(m.sso.session_id, "did:mesh:a", sigma_raw=0.8)
            await hv.activate_session(m.sso.session_id)
            m.delta_engine.capture("did:mesh:a", [hypervisor.audit.delta.VFSChange(path="/x", operation="add", content_hash=f"c{s}")])
            out.append(m)
        return out
    hv = Hypervisor()
    ms = asyncio.run(make(hv, 4))
    sids = [m.sso.session_id for m in ms]
    with pytest.raises(ValueError):                                # unknown session: the ones before it are done
        terminate_sessions_sync(hv, [sids[0], "session:nope", sids[1]])
    assert hv.commitment.get_commitment(sids[0]) is not None and hv.commitment.get_commitment(sids[1]) is None
    assert hv.get_session(sids[0]).sso.state.value == "archived" and hv.get_session(sids[1]).sso.state.value == "active"
    roots = terminate_sessions_sync(hv, sids[1:])
    assert roots == [m.delta_engine.compute_merkle_root() for m in ms[1:]] and all(roots)

    # CommitmentEngine.flush_batch with queued engines: one launch, records in queue order
    ce = CommitmentEngine()
    engines = []
    for s in range(6):
        e = DeltaEngine(f"session:q{s}")
        for i in range(s):                                          # session q0 stays empty
            e.capture("did:mesh:a", [VFSChange(path=f"/f{i}", operation="add", content_hash=f"h{i}")])
        engines.append(e)
    pre = ce.commit("session:done", "ab" * 32, ["did:x"], 3)
    ce.queue_for_batch(pre)
    for e in engines:
        ce.queue_engine_for_batch(e.session_id, e, ["did:mesh:a"])
    recs = ce.flush_batch()
    assert [r.session_id for r in recs] == ["session:done"] + [f"session:q{s}" for s in range(1, 6)]
    for r, e in zip(recs[1:], engines[1:]):
        assert r.merkle_root == e.compute_merkle_root() and r.delta_count == e.turn_count and ce.verify(e.session_id, r.merkle_root)
    assert ce.get_commitment("session:q0") is None and ce.flush_batch() == []
