#!/usr/bin/env python3
"""Generate golden vectors by running the REAL reference (imported from
/root/reference/src) on the deterministic recipes in tests/recipes.py.

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Writes tests/golden/audit_golden.json, tests/golden/liability_golden.json and
tests/golden/slashing_golden.json (``--slashing-only`` rewrites just the last).
The reference's own tests hold no digest constants (timestamps come from
datetime.now()), so these fixtures - the reference's outputs on fixed inputs -
are what pins the oracle and the CUDA path.
"""

from __future__ import annotations

import hashlib
import json
import os
import sys
from datetime import datetime, timedelta, timezone

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference/src")

import recipes  # noqa: E402
from hypervisor.audit.delta import DeltaEngine, SemanticDelta, VFSChange  # noqa: E402
from hypervisor.audit.commitment import CommitmentEngine  # noqa: E402
from hypervisor.liability.vouching import VouchingEngine, VouchRecord  # noqa: E402
from hypervisor.models import ActionDescriptor, ExecutionRing, ReversibilityLevel  # noqa: E402
from hypervisor.rings.enforcer import RingEnforcer  # noqa: E402
import hypervisor  # noqa: E402


def build_engine(fields: list[dict], session_id=None) -> tuple[DeltaEngine, list[bytes]]:
    """Real DeltaEngine holding real SemanticDelta objects with a fixed clock.

    Mirrors DeltaEngine.capture (delta.py:85-115) except that the timestamp is
    the recipe's instead of datetime.now().  Pre-images are captured by
    re-running the reference's own json.dumps expression via compute_hash's
    inputs (we recompute them with the same expression to store their bytes).
    """
    eng = DeltaEngine(session_id or (fields[0]["session_id"] if fields else "session:empty"))
    pre = []
    for f in fields:
        eng._turn_counter += 1
        parent = eng._deltas[-1].delta_hash if eng._deltas else None
        d = SemanticDelta(
            delta_id=f["delta_id"], turn_id=f["turn_id"], session_id=f["session_id"],
            agent_did=f["agent_did"], timestamp=f["timestamp"],
            changes=[VFSChange(**c) for c in f["changes"]], parent_hash=parent)
        d.compute_hash()
        eng._deltas.append(d)
        pre.append(preimage_of(d))
    return eng, pre


def preimage_of(d: SemanticDelta) -> bytes:
    # the exact expression of delta.py:43-62
    return json.dumps(
        {
            "delta_id": d.delta_id, "turn_id": d.turn_id, "session_id": d.session_id,
            "agent_did": d.agent_did, "timestamp": d.timestamp.isoformat(),
            "changes": [
                {"path": c.path, "operation": c.operation, "content_hash": c.content_hash,
                 "previous_hash": c.previous_hash} for c in d.changes],
            "parent_hash": d.parent_hash,
        }, sort_keys=True).encode()


def audit_golden() -> dict:
    out: dict = {"reference_version": hypervisor.__version__}

    # -- SHA-256 known answers (hashlib = the reference's primitive, delta.py:63) + NIST
    out["nist"] = [
        {"msg": "", "digest": "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"},
        {"msg": "abc", "digest": "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"},
        {"msg": "abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq",
         "digest": "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"},
    ]
    for v in out["nist"]:
        assert hashlib.sha256(v["msg"].encode()).hexdigest() == v["digest"]
    out["sha_kats"] = [
        {"seed": s, "len": L, "digest": hashlib.sha256(recipes.kat_message(s, L)).hexdigest()}
        for s, L in enumerate(recipes.SHA_KAT_LENGTHS)
    ]
    out["node_kat"] = {"left": "a" * 64, "right": "b" * 64,
                       "digest": hashlib.sha256(("a" * 64 + "b" * 64).encode()).hexdigest()}

    # -- SURVEY 8c golden session
    gs = {}
    for n in [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 33, 63, 64, 65, 125, 250, 1000]:
        eng, pre = build_engine(recipes.golden_session(n), "session:golden")
        rec = {"root": eng.compute_merkle_root(), "verify": eng.verify_chain(),
               "preimages_sha": hashlib.sha256(b"".join(pre)).hexdigest(),
               "last_leaf": eng._deltas[-1].delta_hash if n else None}
        if n <= 8:
            rec["leaves"] = [d.delta_hash for d in eng._deltas]
        if n == 2:
            rec["preimage0"] = pre[0].decode()
            rec["preimage1"] = pre[1].decode()
        gs[str(n)] = rec
    out["golden_session"] = gs

    # -- benchmark-shaped session (bench_hypervisor.py:157-184 shape), fixed clock
    eng, pre = build_engine(recipes.bench_shape_session(1000))
    out["bench_shape"] = {
        "n": 1000, "root": eng.compute_merkle_root(), "verify": eng.verify_chain(),
        "len0": len(pre[0]), "len1": len(pre[1]), "len_last": len(pre[-1]),
        "leaf0": eng._deltas[0].delta_hash, "last_leaf": eng._deltas[-1].delta_hash,
        "preimages_sha": hashlib.sha256(b"".join(pre)).hexdigest()}

    # -- randomly shaped sessions (>= 1000, SURVEY 8c "Oracle at scale")
    rs = []
    for seed in range(1200):
        eng, pre = build_engine(recipes.random_session(seed))
        rs.append({"seed": seed, "n": len(pre), "root": eng.compute_merkle_root(),
                   "verify": eng.verify_chain(),
                   "last_leaf": eng._deltas[-1].delta_hash if pre else None,
                   "preimages_sha": hashlib.sha256(b"".join(pre)).hexdigest()})
    out["random_sessions"] = rs

    # -- verify_chain behaviour under tampering (delta.py:136-152 quirks)
    vk = []

    def case(name, n, mutate):
        eng, _ = build_engine(recipes.golden_session(n), "session:golden")
        mutate(eng)
        verdict = eng.verify_chain()
        vk.append({"name": name, "n": n, "verdict": verdict,
                   "hashes_after": [d.delta_hash for d in eng._deltas],
                   "root_after": eng.compute_merkle_root()})

    case("clean", 10, lambda e: None)
    case("agent_mid", 10, lambda e: setattr(e._deltas[5], "agent_did", "did:evil"))      # e2e test :421-423
    case("agent_last", 10, lambda e: setattr(e._deltas[9], "agent_did", "did:evil"))     # NOT detected
    case("agent_first", 10, lambda e: setattr(e._deltas[0], "agent_did", "did:evil"))
    case("parent_first_set", 10, lambda e: setattr(e._deltas[0], "parent_hash", "0" * 64))
    case("parent_none_mid", 10, lambda e: setattr(e._deltas[4], "parent_hash", None))
    case("parent_other_mid", 10, lambda e: setattr(e._deltas[4], "parent_hash", "nothex"))
    case("parent_upper_mid", 10, lambda e: setattr(e._deltas[4], "parent_hash", e._deltas[3].delta_hash.upper()))
    case("parent_wrong_mid", 10, lambda e: setattr(e._deltas[7], "parent_hash", "f" * 64))
    case("stored_hash_only", 10, lambda e: setattr(e._deltas[3], "delta_hash", "deadbeef"))  # healed, True
    case("single_ok", 1, lambda e: None)
    case("single_parent_set", 1, lambda e: setattr(e._deltas[0], "parent_hash", "x"))
    case("empty", 0, lambda e: None)
    out["verify_kats"] = vk

    # -- config 2 (every pre-image exactly 256 B)
    c2 = {}
    for s in [0, 1234, 99999999]:
        eng = DeltaEngine(f"session:{s:08d}")
        pads = []
        for i in range(1000):
            parent = eng._deltas[-1].delta_hash if eng._deltas else None
            base = recipes.config2_delta(s, i, 0)
            d0 = SemanticDelta(changes=[], parent_hash=parent, **{k: v for k, v in base.items() if k != "changes"})
            pad = recipes.CONFIG2_LEN - len(preimage_of(d0))
            assert pad >= 0, (s, i, pad)
            f = recipes.config2_delta(s, i, pad)
            d = SemanticDelta(changes=[], parent_hash=parent, **{k: v for k, v in f.items() if k != "changes"})
            assert len(preimage_of(d)) == recipes.CONFIG2_LEN
            d.compute_hash()
            eng._deltas.append(d)
            eng._turn_counter += 1
            pads.append(pad)
        c2[str(s)] = {"last_leaf": eng._deltas[-1].delta_hash, "leaf0": eng._deltas[0].delta_hash,
                      "root": eng.compute_merkle_root(), "verify": eng.verify_chain(),
                      "pad0": pads[0], "pad1": pads[1], "pad_last": pads[-1]}
    out["config2"] = c2

    # -- CommitmentEngine (commitment.py:42-64)
    ce = CommitmentEngine()
    ce.commit("s1", gs["4"]["root"], ["did:a"], 4)
    out["commitment"] = {"root": gs["4"]["root"], "ok": ce.verify("s1", gs["4"]["root"]),
                         "wrong": ce.verify("s1", gs["5"]["root"]), "unknown": ce.verify("nope", gs["4"]["root"])}
    return out


def liability_golden() -> dict:
    out: dict = {}
    # -- the three points the reference's own tests pin (tests/unit/test_liability.py:39-67)
    pts = []
    ve = VouchingEngine()
    ve.vouch("did:h", "did:l", "s1", 0.9, bond_pct=0.2)
    pts.append({"desc": "one voucher", "value": ve.compute_sigma_eff("did:l", "s1", 0.4, 0.5).hex()})
    ve = VouchingEngine()
    ve.vouch("did:h", "did:l", "s1", 1.0, bond_pct=0.8)
    pts.append({"desc": "capped", "value": ve.compute_sigma_eff("did:l", "s1", 0.9, 1.0).hex()})
    ve = VouchingEngine()
    ve.vouch("did:h1", "did:l", "s1", 0.9, bond_pct=0.2)
    ve.vouch("did:h2", "did:l", "s1", 0.8, bond_pct=0.2)
    pts.append({"desc": "two vouchers", "value": ve.compute_sigma_eff("did:l", "s1", 0.3, 0.5).hex()})
    out["pinned_points"] = pts

    # -- random graphs through the real VouchingEngine / RingEnforcer
    enf = RingEnforcer()
    now = datetime.now(timezone.utc)
    graphs = []
    for seed, (A, E) in enumerate([(1, 0), (5, 12), (40, 200), (300, 3000), (1000, 10000)]):
        g = recipes.random_vouch_graph(seed, A, E)
        ve = VouchingEngine()
        for k, (a, b, vs, pct, active, ek) in enumerate(g["edges"]):
            exp = None if ek == 0 else (now + timedelta(days=3650) if ek == 1 else now - timedelta(days=3650))
            r = VouchRecord(vouch_id=f"vouch:{k}", voucher_did=f"did:{a}", vouchee_did=f"did:{b}",
                            session_id="s", bonded_sigma_pct=pct, bonded_amount=vs * pct,   # vouching.py:105
                            expiry=exp, is_active=active)
            ve._vouches[r.vouch_id] = r
            # decoy record in another session: must not contribute
            if k % 17 == 0:
                ve._vouches[f"decoy:{k}"] = VouchRecord(
                    vouch_id=f"decoy:{k}", voucher_did=f"did:{a}", vouchee_did=f"did:{b}",
                    session_id="other", bonded_sigma_pct=pct, bonded_amount=vs * pct)
        se, ring, dem = [], [], []
        for i in range(A):
            x = ve.compute_sigma_eff(f"did:{i}", "s", g["sigma"][i], g["omega"][i])
            se.append(float(x).hex())
            ring.append(int(enf.compute_ring(x, g["consensus"][i]).value))
            dem.append(bool(enf.should_demote(ExecutionRing(g["current_ring"][i]), x)))
        graphs.append({"seed": seed, "A": A, "E": E, "sigma_eff": se, "ring": ring, "demote": dem})
    out["graphs"] = graphs

    # -- thresholds (tests/unit/test_models.py:15-33) and gate grid (enforcer.py:44-128)
    sig = [0.0, 0.3, 0.59, 0.5999999, 0.6, 0.6000001, 0.61, 0.85, 0.94, 0.9499999, 0.95, 0.9500001, 0.96, 1.0,
           float("nan"), -0.5, 1.5]
    out["ring_grid"] = [
        {"sigma": float(s).hex() if s == s else "nan", "consensus": c,
         "ring": int(ExecutionRing.from_sigma_eff(s, c).value)}
        for s in sig for c in (False, True)]
    checks = []
    for admin in (False, True):
        for ro in (False, True):
            for rev in (ReversibilityLevel.NONE, ReversibilityLevel.FULL, ReversibilityLevel.PARTIAL):
                act = ActionDescriptor(action_id="a", name="n", execute_api="/x", reversibility=rev,
                                       is_read_only=ro, is_admin=admin)
                for ar in range(4):
                    for s in sig:
                        for cons in (False, True):
                            for wit in (False, True):
                                r = enf.check(ExecutionRing(ar), act, s, cons, wit)
                                checks.append([int(admin), int(ro), rev.value, ar,
                                               float(s).hex() if s == s else "nan", int(cons), int(wit),
                                               int(r.allowed), int(r.required_ring.value), r.reason,
                                               int(r.requires_consensus), int(r.requires_sre_witness)])
    out["checks"] = checks
    # -- LiabilityMatrix (liability/__init__.py:24-139)
    import random
    from hypervisor.liability import LiabilityMatrix
    rng = random.Random(3)
    m = LiabilityMatrix("s")
    edges = []
    for k in range(400):
        a, b = rng.randrange(40), rng.randrange(40)
        amt = rng.choice([rng.uniform(0, 0.8), 0.0, -0.0, 1e-300, 0.1])
        edges.append((f"did:{a}", f"did:{b}", amt, f"v{k}"))
        m.add_edge(*edges[-1])
    mat = {"edges": [[e[0], e[1], float(e[2]).hex(), e[3]] for e in edges],
           "exposure": {f"did:{i}": float(m.total_exposure(f"did:{i}")).hex() for i in range(42)},
           "has_cycle": m.has_cycle()}
    m2 = LiabilityMatrix("t")
    for a, b in [("a", "b"), ("b", "c"), ("c", "d"), ("a", "e")]:
        m2.add_edge(a, b, 0.1, a + b)
    mat["acyclic"] = {"has_cycle": m2.has_cycle(), "cascade_a": m2.cascade_path("a"),
                      "cascade_a3": m2.cascade_path("a", 3), "cascade_d": m2.cascade_path("d")}
    m2.add_edge("d", "a", 0.1, "da")
    mat["cyclic"] = {"has_cycle": m2.has_cycle(), "cascade_a": m2.cascade_path("a", 4)}
    out["matrix"] = mat
    out["risk_weights"] = {lv.value: float(lv.default_risk_weight).hex() for lv in ReversibilityLevel}
    return out


def fhex(x: float) -> str:
    return "nan" if x != x else float(x).hex()


N_SLASH_CASES = 400


def slashing_golden() -> dict:
    """The real SlashingEngine (liability/slashing.py:62-141) over the real VouchingEngine on the
    deterministic cases of recipes.random_slash_case: history (pre-order, as appended at :119), clips,
    final agent_scores in dict order, and which bonds ended up released."""
    from hypervisor.liability.slashing import SlashingEngine
    now = datetime.now(timezone.utc)
    cases = []
    for seed in range(N_SLASH_CASES):
        c = recipes.random_slash_case(seed)
        ve = VouchingEngine()
        for k, (a, b, active, ek) in enumerate(c["edges"]):
            exp = None if ek == 0 else (now + timedelta(days=3650) if ek == 1 else now - timedelta(days=3650))
            ve._vouches[f"vouch:{k}"] = VouchRecord(
                vouch_id=f"vouch:{k}", voucher_did=f"did:{a}", vouchee_did=f"did:{b}", session_id="s",
                bonded_sigma_pct=0.2, bonded_amount=0.1, expiry=exp, is_active=active)
            if k % 5 == 0:      # same edge in another session: must be left alone
                ve._vouches[f"decoy:{k}"] = VouchRecord(
                    vouch_id=f"decoy:{k}", voucher_did=f"did:{a}", vouchee_did=f"did:{b}", session_id="other",
                    bonded_sigma_pct=0.2, bonded_amount=0.1)
        se = SlashingEngine(ve)
        scores = {f"did:{a}": v for a, v in c["scores"]}
        top = []
        for (v, sig, om, dep) in c["events"]:
            r = se.slash(vouchee_did=f"did:{v}", session_id="s", vouchee_sigma=sig, risk_weight=om,
                         reason="golden", agent_scores=scores, cascade_depth=dep)
            top.append(se.history.index(r))
        hist = []
        for r in se.history:
            hist.append({
                "vouchee": int(r.vouchee_did[4:]), "depth": r.cascade_depth, "sigma_before": fhex(r.vouchee_sigma_before),
                "sigma_after": fhex(r.vouchee_sigma_after), "reason": r.reason,
                "clips": [[int(cl.voucher_did[4:]), fhex(cl.sigma_before), fhex(cl.sigma_after), fhex(cl.risk_weight),
                           int(cl.vouch_id[6:])] for cl in r.voucher_clips]})
        assert all(v.is_active for k, v in ve._vouches.items() if k.startswith("decoy:"))
        cases.append({"seed": seed, "history": hist, "top_level": top,
                      "scores": [[int(k[4:]), fhex(v)] for k, v in scores.items()],
                      "active": [int(ve._vouches[f"vouch:{k}"].is_active) for k in range(len(c["edges"]))]})
    # the four cases the reference's own tests pin (tests/unit/test_slashing.py:14-89)
    pinned = []
    for (edges, sc, ev) in [
        ([("good", "bad", 0.9)], {"bad": 0.7, "good": 0.9}, ("bad", 0.7, 0.5)),
        ([("voucher", "bad", 0.9)], {"bad": 0.5, "voucher": 0.9}, ("bad", 0.5, 0.5)),
        ([("voucher", "bad", 0.8)], {"bad": 0.1, "voucher": 0.06}, ("bad", 0.1, 0.95)),
        ([("v1", "bad", 0.8), ("v2", "bad", 0.7)], {"bad": 0.4, "v1": 0.8, "v2": 0.7}, ("bad", 0.4, 0.3)),
    ]:
        ve = VouchingEngine()
        se = SlashingEngine(ve)
        for (a, b, s) in edges:
            ve.vouch(f"did:mesh:{a}", f"did:mesh:{b}", "session:test-slash", s)
        scores = {f"did:mesh:{k}": v for k, v in sc.items()}
        r = se.slash(vouchee_did=f"did:mesh:{ev[0]}", session_id="session:test-slash", vouchee_sigma=ev[1],
                     risk_weight=ev[2], reason="r", agent_scores=scores)
        pinned.append({"edges": [list(e) for e in edges], "scores_in": sc, "event": list(ev),
                       "scores_out": {k[9:]: fhex(v) for k, v in scores.items()},
                       "clips": [[cl.voucher_did[9:], fhex(cl.sigma_before), fhex(cl.sigma_after)] for cl in r.voucher_clips]})
    return {"cases": cases, "pinned": pinned, "sigma_floor": fhex(SlashingEngine.SIGMA_FLOOR),
            "max_cascade_depth": SlashingEngine.MAX_CASCADE_DEPTH}


if __name__ == "__main__":
    sg = slashing_golden()
    with open(os.path.join(HERE, "slashing_golden.json"), "w") as f:
        json.dump(sg, f, separators=(",", ":"), sort_keys=True)
    print("slashing cases:", len(sg["cases"]), "with cascades:",
          sum(1 for c in sg["cases"] if any(h["depth"] > 0 for h in c["history"])),
          "depth 2:", sum(1 for c in sg["cases"] if any(h["depth"] > 1 for h in c["history"])))
    if "--slashing-only" in sys.argv:
        sys.exit(0)
    a = audit_golden()
    with open(os.path.join(HERE, "audit_golden.json"), "w") as f:
        json.dump(a, f, indent=0, sort_keys=True)
    l = liability_golden()
    with open(os.path.join(HERE, "liability_golden.json"), "w") as f:
        json.dump(l, f, indent=0, sort_keys=True)
    print("golden root n=1000:", a["golden_session"]["1000"]["root"])
    print("config2 s=1234 last leaf:", a["config2"]["1234"]["last_leaf"])
    print("files:", os.path.getsize(os.path.join(HERE, "audit_golden.json")),
          os.path.getsize(os.path.join(HERE, "liability_golden.json")))
