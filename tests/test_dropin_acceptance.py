"""Drop-in acceptance: the scenarios the reference's own unit / integration tests exercise on the
classes of the hot path (tests/unit/test_audit.py, test_liability.py, test_rings.py, test_models.py,
tests/integration/test_hypervisor_e2e.py:382-447, benchmarks/bench_hypervisor.py:121-184), run
against this package's classes of the same names.  Written from the behaviour, not copied; every
hash / sum goes through the CUDA library."""

import hashlib

import pytest

pytestmark = pytest.mark.gpu


# ------------------------------------------------------------------ audit (test_audit.py)
class TestDeltaEngine:
    def setup_method(self):
        from agent_hypervisor_b200 import DeltaEngine
        self.engine = DeltaEngine("session:test-audit")

    def _capture(self, n, prefix="/f"):
        from agent_hypervisor_b200 import VFSChange
        return [self.engine.capture("did:a", [VFSChange(path=f"{prefix}{i}.txt", operation="add")]) for i in range(n)]

    def test_first_capture(self):
        from agent_hypervisor_b200 import VFSChange
        d = self.engine.capture("did:agent1", [VFSChange(path="/file.txt", operation="add", content_hash="abc123")])
        assert d.turn_id == 1 and d.parent_hash is None and len(d.delta_hash) == 64
        assert d.delta_id == "delta:1" and d.session_id == "session:test-audit"
        assert d.delta_hash == hashlib.sha256(d.preimage()).hexdigest()

    def test_chain_links(self):
        ds = self._capture(3)
        assert ds[0].parent_hash is None
        assert ds[1].parent_hash == ds[0].delta_hash and ds[2].parent_hash == ds[1].delta_hash
        assert self.engine.turn_count == 3 and self.engine.deltas is not self.engine._deltas

    def test_verify_clean_chain(self):
        self._capture(5)
        assert self.engine.verify_chain() is True

    def test_root_shape_and_empty(self):
        assert self.engine.compute_merkle_root() is None
        self._capture(4)
        root = self.engine.compute_merkle_root()
        assert isinstance(root, str) and len(root) == 64 and root == root.lower()

    def test_custom_delta_id(self):
        from agent_hypervisor_b200 import VFSChange
        d = self.engine.capture("did:a", [VFSChange(path="/x", operation="delete")], delta_id="custom-7")
        assert d.delta_id == "custom-7" and d.turn_id == 1


class TestCommitmentEngine:
    def test_commit_verify_queue(self):
        from agent_hypervisor_b200 import CommitmentEngine
        e = CommitmentEngine()
        r = e.commit("session:1", "abc123", ["did:a", "did:b"], 10)
        assert e.verify("session:1", "abc123") and not e.verify("session:1", "wrong")
        assert not e.verify("nonexistent", "abc")
        assert e.get_commitment("session:1") is r and r.committed_to == "local" and r.delta_count == 10
        e.queue_for_batch(r)
        assert e.flush_batch() == [r] and e.flush_batch() == []


# ------------------------------------------------------------------ e2e audit trail (test_hypervisor_e2e.py:382-447)
def test_audit_trail_tamper_and_idempotence():
    from agent_hypervisor_b200 import CommitmentEngine, DeltaEngine, VFSChange
    eng = DeltaEngine("session:e2e")
    for i in range(10):
        eng.capture(f"did:agent{i % 3}", [VFSChange(path=f"/doc{i}.md", operation="modify", content_hash=f"c{i}",
                                                    previous_hash=f"p{i}")])
    assert eng.turn_count == 10 and eng.verify_chain()
    r1 = eng.compute_merkle_root()
    assert r1 == eng.compute_merkle_root()                       # idempotent
    ce = CommitmentEngine()
    ce.commit("session:e2e", r1, ["did:agent0"], 10)
    assert ce.verify("session:e2e", eng.compute_merkle_root())
    eng._deltas[5].agent_did = "did:tampered"                    # tamper in the middle
    assert eng.verify_chain() is False
    assert not ce.verify("session:e2e", eng.compute_merkle_root())


# ------------------------------------------------------------------ benchmark loops (bench_hypervisor.py:145-184)
@pytest.mark.parametrize("n", [1, 10, 50, 100])
def test_benchmark_shaped_usage(n):
    from agent_hypervisor_b200 import DeltaEngine, VFSChange
    from oracle import audit_oracle as ao
    de = DeltaEngine("bench-merkle")
    for i in range(n):
        de.capture("did:mesh:a", [VFSChange(path=f"/f{i}", operation="add", content_hash=f"h{i}")])
    assert de.compute_merkle_root() == ao.merkle_root_hex([d.delta_hash for d in de.deltas])
    assert de.verify_chain()


# ------------------------------------------------------------------ liability (test_liability.py)
class TestVouchingEngine:
    def setup_method(self):
        from agent_hypervisor_b200 import VouchingEngine
        self.engine = VouchingEngine()
        self.session = "session:test-1"

    def test_default_bond(self):
        r = self.engine.vouch("did:mesh:high", "did:mesh:low", self.session, 0.8)
        assert (r.voucher_did, r.vouchee_did, r.is_active, r.bonded_sigma_pct) == ("did:mesh:high", "did:mesh:low", True, 0.20)
        assert abs(r.bonded_amount - 0.16) < 1e-9 and r.vouch_id.startswith("vouch:")

    def test_protocol_violations(self):
        from agent_hypervisor_b200 import VouchingError
        with pytest.raises(VouchingError, match="Cannot vouch for yourself"):
            self.engine.vouch("did:mesh:a", "did:mesh:a", self.session, 0.8)
        with pytest.raises(VouchingError, match="below minimum"):
            self.engine.vouch("did:mesh:low", "did:mesh:other", self.session, 0.3)
        self.engine.vouch("did:mesh:a", "did:mesh:b", self.session, 0.8)
        with pytest.raises(VouchingError, match="Circular"):
            self.engine.vouch("did:mesh:b", "did:mesh:a", self.session, 0.7)
        self.engine.vouch("did:mesh:b", "did:mesh:c", self.session, 0.8)
        # like the reference (vouching.py:214-229): the walk never revisits the would-be voucher, so
        # the 3-cycle c -> a -> b -> c is NOT caught; only direct mutual vouching is
        assert self.engine.vouch("did:mesh:c", "did:mesh:a", self.session, 0.9).is_active
        with pytest.raises(VouchingError, match="exceed max exposure"):
            for k in range(10):
                self.engine.vouch("did:mesh:x", f"did:mesh:y{k}", self.session, 0.9, bond_pct=0.2)
        with pytest.raises(VouchingError, match="not found"):
            self.engine.release_bond("vouch:nope")

    def test_sigma_eff_points(self):
        self.engine.vouch("did:mesh:high", "did:mesh:low", self.session, 0.9, bond_pct=0.5)
        assert abs(self.engine.compute_sigma_eff("did:mesh:low", self.session, 0.3, 0.2) - 0.39) < 1e-9
        assert self.engine.compute_sigma_eff("did:mesh:low", self.session, 0.8, 1.0) <= 1.0
        assert self.engine.compute_sigma_eff("did:mesh:nobody", self.session, 0.25, 0.9) == 0.25
        assert self.engine.compute_sigma_eff("did:mesh:low", "another-session", 0.3, 0.2) == 0.3

    def test_two_vouchers(self):
        self.engine.vouch("did:mesh:a", "did:mesh:low", self.session, 0.8, bond_pct=0.5)
        self.engine.vouch("did:mesh:b", "did:mesh:low", self.session, 0.6, bond_pct=0.5)
        assert abs(self.engine.compute_sigma_eff("did:mesh:low", self.session, 0.1, 0.5) - 0.45) < 1e-9

    def test_release_and_expiry(self):
        from datetime import datetime, timedelta, timezone
        self.engine.vouch("did:mesh:a", "did:mesh:b", self.session, 0.8)
        self.engine.vouch("did:mesh:a", "did:mesh:c", self.session, 0.8)
        past = datetime.now(timezone.utc) - timedelta(hours=1)
        self.engine.vouch("did:mesh:d", "did:mesh:b", self.session, 0.9, expiry=past)       # already expired
        assert abs(self.engine.compute_sigma_eff("did:mesh:b", self.session, 0.1, 1.0) - 0.26) < 1e-9
        assert self.engine.release_session_bonds(self.session) == 3
        assert self.engine.get_vouchers_for("did:mesh:b", self.session) == []
        assert self.engine.compute_sigma_eff("did:mesh:b", self.session, 0.1, 1.0) == 0.1

    def test_exposure(self):
        self.engine.vouch("did:mesh:a", "did:mesh:b", self.session, 0.8, bond_pct=0.3)
        self.engine.vouch("did:mesh:a", "did:mesh:c", self.session, 0.8, bond_pct=0.2)
        assert abs(self.engine.get_total_exposure("did:mesh:a", self.session) - 0.40) < 1e-9
        assert abs(self.engine.get_total_exposure_batch(["did:mesh:a", "did:mesh:zz"], self.session)[0] - 0.40) < 1e-9


class TestLiabilityMatrix:
    def setup_method(self):
        from agent_hypervisor_b200 import LiabilityMatrix
        self.matrix = LiabilityMatrix("session:test-1")

    def test_queries(self):
        self.matrix.add_edge("did:a", "did:b", 0.2, "v1")
        self.matrix.add_edge("did:a", "did:c", 0.3, "v2")
        assert len(self.matrix.who_vouches_for("did:b")) == 1 and len(self.matrix.who_is_vouched_by("did:a")) == 2
        assert abs(self.matrix.total_exposure("did:a") - 0.5) < 1e-9
        assert not self.matrix.has_cycle()
        self.matrix.add_edge("did:b", "did:a", 0.2, "v3")
        assert self.matrix.has_cycle()
        self.matrix.clear()
        assert len(self.matrix.edges) == 0


# ------------------------------------------------------------------ rings (test_rings.py, test_models.py)
class TestRingEnforcer:
    def setup_method(self):
        from agent_hypervisor_b200 import RingEnforcer
        self.enforcer = RingEnforcer()

    def test_gate_outcomes(self):
        from agent_hypervisor_b200 import ActionDescriptor, ExecutionRing, ReversibilityLevel
        read = ActionDescriptor(action_id="search", name="Search", execute_api="/search", is_read_only=True)
        assert self.enforcer.check(ExecutionRing.RING_3_SANDBOX, read, 0.3).allowed
        draft = ActionDescriptor("draft", "Draft", "/draft", undo_api="/draft/undo", reversibility=ReversibilityLevel.FULL)
        r = self.enforcer.check(ExecutionRing.RING_3_SANDBOX, draft, 0.7)
        assert not r.allowed and "insufficient" in r.reason.lower()
        r = self.enforcer.check(ExecutionRing.RING_2_STANDARD, draft, 0.3)
        assert not r.allowed and r.reason == "Ring 2 requires σ_eff > 0.6, got 0.300"
        delete = ActionDescriptor("delete", "Delete", "/delete", reversibility=ReversibilityLevel.NONE)
        r = self.enforcer.check(ExecutionRing.RING_1_PRIVILEGED, delete, 0.96, has_consensus=False)
        assert not r.allowed and r.requires_consensus
        assert self.enforcer.check(ExecutionRing.RING_1_PRIVILEGED, delete, 0.96, has_consensus=True).allowed
        config = ActionDescriptor("config", "Config", "/config", is_admin=True)
        r = self.enforcer.check(ExecutionRing.RING_0_ROOT, config, 1.0, has_sre_witness=False)
        assert not r.allowed and r.requires_sre_witness
        assert self.enforcer.check(ExecutionRing.RING_0_ROOT, config, 1.0, has_sre_witness=True).allowed

    def test_demotion_and_thresholds(self):
        from agent_hypervisor_b200 import ExecutionRing, ReversibilityLevel
        assert self.enforcer.should_demote(ExecutionRing.RING_2_STANDARD, sigma_eff=0.3)
        assert not self.enforcer.should_demote(ExecutionRing.RING_2_STANDARD, sigma_eff=0.7)
        f = ExecutionRing.from_sigma_eff
        assert f(0.96, has_consensus=True) == ExecutionRing.RING_1_PRIVILEGED
        assert f(0.96, has_consensus=False) == ExecutionRing.RING_2_STANDARD
        assert f(0.61) == ExecutionRing.RING_2_STANDARD and f(0.60) == ExecutionRing.RING_3_SANDBOX
        assert f(0.95, has_consensus=True) == ExecutionRing.RING_2_STANDARD and f(0.0) == ExecutionRing.RING_3_SANDBOX
        assert self.enforcer.compute_ring(0.85) == ExecutionRing.RING_2_STANDARD       # bench_hypervisor.py:125
        assert ReversibilityLevel.FULL.default_risk_weight == pytest.approx(0.2)
        assert ReversibilityLevel.PARTIAL.default_risk_weight == pytest.approx(0.65)
        assert ReversibilityLevel.NONE.default_risk_weight == pytest.approx(0.95)


# ------------------------------------------------------------------ slashing (test_slashing.py, test_scenarios.py:471-545)
class TestSlashingEngine:
    def setup_method(self):
        from agent_hypervisor_b200 import SlashingEngine, VouchingEngine
        self.vouching = VouchingEngine()
        self.slashing = SlashingEngine(self.vouching)
        self.session = "session:test-slash"

    def _slash(self, scores, sigma, omega, reason="violation"):
        return self.slashing.slash(vouchee_did="did:mesh:bad", session_id=self.session, vouchee_sigma=sigma,
                                   risk_weight=omega, reason=reason, agent_scores=scores)

    def test_vouchee_goes_to_zero(self):
        scores = {"did:mesh:bad": 0.7, "did:mesh:good": 0.9}
        self.vouching.vouch("did:mesh:good", "did:mesh:bad", self.session, 0.9)
        r = self._slash(scores, 0.7, 0.5)
        assert scores["did:mesh:bad"] == 0.0 and r.vouchee_sigma_after == 0.0 and r.vouchee_sigma_before == 0.7
        assert r.session_id == self.session and r.reason == "violation" and r.cascade_depth == 0

    def test_voucher_clip_and_bond_release(self):
        scores = {"did:mesh:bad": 0.5, "did:mesh:voucher": 0.9}
        rec = self.vouching.vouch("did:mesh:voucher", "did:mesh:bad", self.session, 0.9)
        r = self._slash(scores, 0.5, 0.5, "Hallucination")
        assert len(r.voucher_clips) == 1
        clip = r.voucher_clips[0]
        assert clip.voucher_did == "did:mesh:voucher" and clip.vouch_id == rec.vouch_id and clip.risk_weight == 0.5
        assert abs(clip.sigma_before - 0.9) < 1e-9 and abs(clip.sigma_after - 0.45) < 1e-9
        assert abs(scores["did:mesh:voucher"] - 0.45) < 1e-9
        assert rec.is_active is False and rec.released_at is not None
        assert self.vouching.get_vouchers_for("did:mesh:bad", self.session) == []

    def test_floor(self):
        from agent_hypervisor_b200 import SlashingEngine
        scores = {"did:mesh:bad": 0.1, "did:mesh:voucher": 0.06}
        self.vouching.vouch("did:mesh:voucher", "did:mesh:bad", self.session, 0.8)
        self._slash(scores, 0.1, 0.95, "Fraud")
        assert scores["did:mesh:voucher"] == SlashingEngine.SIGMA_FLOOR

    def test_all_vouchers_clipped_and_history(self):
        scores = {"did:mesh:bad": 0.4, "did:mesh:v1": 0.8, "did:mesh:v2": 0.7}
        self.vouching.vouch("did:mesh:v1", "did:mesh:bad", self.session, 0.8)
        self.vouching.vouch("did:mesh:v2", "did:mesh:bad", self.session, 0.7)
        r = self._slash(scores, 0.4, 0.3, "Mute triggered")
        assert [c.voucher_did for c in r.voucher_clips] == ["did:mesh:v1", "did:mesh:v2"]
        assert abs(scores["did:mesh:v1"] - 0.56) < 1e-9 and abs(scores["did:mesh:v2"] - 0.49) < 1e-9
        assert self.slashing.history == [r] and self.slashing.history is not self.slashing._slash_history

    def test_cascade_two_levels_and_unknown_voucher(self):
        """a <- b <- c <- d: slashing a at omega 0.95 wipes b (cascade), then c (depth 2), and clips d without
        going further; an agent missing from agent_scores is read as 0.0 and gets the floor."""
        s = self.session
        for voucher, vouchee in (("b", "a"), ("c", "b"), ("d", "c"), ("e", "d")):
            self.vouching.vouch(f"did:{voucher}", f"did:{vouchee}", s, 0.9)
        scores = {"did:a": 0.4, "did:b": 0.9, "did:c": 0.9, "did:e": 0.9}          # did:d missing
        r = self.slashing.slash("did:a", s, 0.4, 0.95, "drift", scores)
        h = self.slashing.history
        assert [(x.vouchee_did, x.cascade_depth) for x in h] == [("did:a", 0), ("did:b", 1), ("did:c", 2)]
        assert h[0] is r and h[1].reason == "Cascade from did:a: drift"
        assert h[2].reason == "Cascade from did:b: Cascade from did:a: drift"
        assert scores == {"did:a": 0.0, "did:b": 0.0, "did:c": 0.0, "did:e": 0.9, "did:d": 0.05}
        assert list(scores) == ["did:a", "did:b", "did:c", "did:e", "did:d"]
        assert len(self.vouching.get_vouchers_for("did:d", s)) == 1           # the e -> d bond was never reached

    def test_works_over_the_reference_style_engine_surface(self):
        """Only `_vouches` and `release_bond` of the vouching engine are used; release_bond errors propagate."""
        from agent_hypervisor_b200 import VouchingError
        self.vouching.vouch("did:v", "did:mesh:bad", self.session, 0.9)
        calls = []
        orig = self.vouching.release_bond
        self.vouching.release_bond = lambda vid: (calls.append(vid), orig(vid))[1]
        self._slash({"did:v": 0.8}, 0.5, 0.5)
        assert len(calls) == 1
        with pytest.raises(VouchingError):
            self.vouching.release_bond("vouch:none")


def test_commitment_verify_batch_matches_per_session_verify():
    """CommitmentEngine.verify_batch == [verify(sid, engine.compute_merkle_root())] (commitment.py:59-64 over
    delta.py:117-134): clean sessions, a tampered stored hash, an overwritten non-digest hash, an empty engine,
    an unknown session."""
    from agent_hypervisor_b200 import CommitmentEngine, DeltaEngine, VFSChange
    engines = []
    for s in range(12):
        e = DeltaEngine(f"s{s}")
        for i in range(s):                               # 0 .. 11 deltas
            e.capture("did:a", [VFSChange(path=f"/f{i}", operation="add", content_hash=f"h{s}{i}")])
        engines.append(e)
    ce = CommitmentEngine()
    for e in engines[:-1]:                               # the last session is never committed
        root = e.compute_merkle_root()
        if root is not None:
            ce.commit(e.session_id, root, ["did:a"], e.turn_count)
    engines[5]._deltas[2].delta_hash = "0" * 64          # stored leaf changed after the commit
    engines[7]._deltas[0].delta_hash = "not-a-digest"    # the reference would hash the string as it is
    want = [ce.verify(e.session_id, e.compute_merkle_root()) for e in engines]
    got = ce.verify_batch([e.session_id for e in engines], engines)
    assert got == want
    assert got == [False, True, True, True, True, False, True, False, True, True, True, False]


def test_sigma_eff_many_sessions_one_launch(liability_golden):
    """compute_sigma_eff_many over requests from different sessions == the single calls, and == the real
    engine's recorded values for the golden graphs (tests/golden/liability_golden.json)."""
    from datetime import datetime, timedelta, timezone
    import recipes
    from agent_hypervisor_b200 import VouchingEngine, VouchRecord
    now = datetime.now(timezone.utc)
    ve = VouchingEngine()
    reqs, want = [], []
    for gi, g in enumerate(liability_golden["graphs"][:4]):
        rg = recipes.random_vouch_graph(g["seed"], g["A"], g["E"])
        for k, (a, b, vs, pct, active, ek) in enumerate(rg["edges"]):
            exp = None if ek == 0 else (now + timedelta(days=3650) if ek == 1 else now - timedelta(days=3650))
            vid = f"vouch:{gi}:{k}"
            ve._vouches[vid] = VouchRecord(vouch_id=vid, voucher_did=f"did:{a}", vouchee_did=f"did:{b}", session_id=f"s{gi}",
                                           bonded_sigma_pct=pct, bonded_amount=vs * pct, expiry=exp, is_active=active)
        for i in range(g["A"]):
            reqs.append((f"did:{i}", f"s{gi}", rg["sigma"][i], rg["omega"][i]))
            want.append(float.fromhex(g["sigma_eff"][i]))
    order = list(range(len(reqs)))[::-1][::3] + list(range(len(reqs)))       # interleaved sessions, repeats
    got = ve.compute_sigma_eff_many([reqs[i] for i in order])
    assert [float(x) for x in got] == [want[i] for i in order]
    assert ve.compute_sigma_eff(*reqs[7]) == want[7]
    assert len(ve.compute_sigma_eff_many([])) == 0
