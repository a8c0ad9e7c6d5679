"""GPU parity for the slashing pass (ahv_slash_cascade, reference liability/slashing.py:62-141):
golden records of the real SlashingEngine, the C restatement on large batches, and the drop-in class."""

from __future__ import annotations

from datetime import datetime, timedelta, timezone

import numpy as np
import pytest

import recipes
from helpers import check_slash_against_golden, fromhex, same_f64, slash_batch
from oracle import coracle as co
from oracle import liability_oracle as lo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from agent_hypervisor_b200 import _native as nat
    return nat.get_context()


@pytest.fixture(params=[0, 2, 1], ids=["lanes+staged", "staged", "global-only"])
def variant(request, ctx):
    """Every kernel form: lane per session for groups of small sessions + warp per session staged in shared
    memory where it fits (default); warp per session only; everything on global memory."""
    ctx.set_option("slash_variant", request.param)
    yield request.param
    ctx.set_option("slash_variant", 0)


def run_gpu(ctx, b):
    o = ctx.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                          b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"],
                          event_depth=b.get("event_depth"))
    assert not o["status"].any()
    return (o["scores"], o["touched"], o["active"], o["slash_log"], o["clip_log"], o["slash_count"], o["clip_count"])


def assert_same_outputs(got, want, b):
    g_sc, g_t, g_a, g_sl, g_cl, g_ns, g_nc = got
    w_sc, w_t, w_a, w_sl, w_cl, w_ns, w_nc = want
    assert same_f64(g_sc, w_sc) and np.array_equal(g_t, w_t) and np.array_equal(g_a, w_a)
    assert np.array_equal(g_ns, w_ns) and np.array_equal(g_nc, w_nc)
    srp, erp, rp = b["session_agent_rowptr"], b["event_rowptr"], b["rowptr"]
    for s in range(len(srp) - 1):
        e0 = int(rp[int(srp[s])])
        a = g_sl[int(erp[s]) + e0:][: int(g_ns[s])]; w = w_sl[int(erp[s]) + e0:][: int(w_ns[s])]
        for f in ("agent", "parent", "event", "depth", "clip_begin", "clip_count"):
            assert np.array_equal(a[f], w[f]), (s, f)
        assert same_f64(a["sigma_before"], w["sigma_before"])
        a = g_cl[e0:][: int(g_nc[s])]; w = w_cl[e0:][: int(w_nc[s])]
        assert np.array_equal(a["edge"], w["edge"]) and np.array_equal(a["slash"], w["slash"])
        assert same_f64(a["sigma_before"], w["sigma_before"]) and same_f64(a["sigma_after"], w["sigma_after"])


def test_golden_cases_one_launch(ctx, slashing_golden, variant):
    gold = slashing_golden["cases"]
    cases = [recipes.random_slash_case(g["seed"]) for g in gold]
    b = slash_batch(cases)
    check_slash_against_golden(cases, gold, b, run_gpu(ctx, b))
    one = slash_batch(cases[7:8])                 # and a batch of one session
    check_slash_against_golden(cases[7:8], gold[7:8], one, run_gpu(ctx, one))


def test_reference_unit_test_cases(ctx, slashing_golden):
    """tests/unit/test_slashing.py:14-89 of the reference, as recorded from the real engine."""
    for p in slashing_golden["pinned"]:
        names = list(p["scores_in"])
        idx = {n: i for i, n in enumerate(names)}
        c = dict(n_agents=len(names), edges=[(idx[a], idx[v], True, 0) for a, v, _ in p["edges"]],
                 scores=[(idx[n], v) for n, v in p["scores_in"].items()],
                 events=[(idx[p["event"][0]], p["event"][1], p["event"][2], 0)])
        b = slash_batch([c])
        sc, _, act, _, cl, _, nc = run_gpu(ctx, b)
        for n, v in p["scores_out"].items():
            assert sc[idx[n]] == fromhex(v)
        assert not act.any() and int(nc[0]) == len(p["clips"])
        for q, (n, x, y) in zip(cl, p["clips"]):
            assert names[int(b["voucher"][int(q["edge"])])] == n and q["sigma_before"] == fromhex(x) and q["sigma_after"] == fromhex(y)


def big_batch(seed, n_sessions, giant=None):
    """Random sessions straight as arrays (fast to build): A agents and E records each, duplicates and
    cycles included; ``giant`` = (A, E) appends one session with a 60 % hub."""
    rng = np.random.default_rng(seed)
    A_s = rng.integers(1, 24, n_sessions)
    E_s = (A_s * rng.choice([0.0, 1.0, 3.0, 8.0], n_sessions)).astype(np.int64)
    if giant:
        A_s = np.append(A_s, giant[0]); E_s = np.append(E_s, giant[1])
    S = len(A_s)
    srp = np.zeros(S + 1, np.uint64); srp[1:] = np.cumsum(A_s)
    A, E = int(srp[-1]), int(E_s.sum())
    sess_of_edge = np.repeat(np.arange(S), E_s)
    a_lo = srp[:-1].astype(np.int64)[sess_of_edge]
    width = A_s[sess_of_edge]
    hub = rng.random(E) < 0.6
    vouchee = a_lo + np.where(hub, 0, rng.integers(0, 1 << 30, E) % width)
    few = np.minimum(width, 3)
    voucher = a_lo + np.where(rng.random(E) < 0.3, rng.integers(0, 1 << 30, E) % few, rng.integers(0, 1 << 30, E) % width)
    order = np.argsort(vouchee, kind="stable")          # rows by vouchee, insertion order inside a row
    vouchee, voucher = vouchee[order], voucher[order]
    rowptr = np.zeros(A + 1, np.uint32)
    np.add.at(rowptr, vouchee + 1, 1)
    rowptr = np.cumsum(rowptr).astype(np.uint32)
    n_ev = rng.integers(1, 4, S)
    erp = np.zeros(S + 1, np.uint64); erp[1:] = np.cumsum(n_ev)
    sess_of_ev = np.repeat(np.arange(S), n_ev)
    NE = int(erp[-1])
    ev_agent = srp[:-1].astype(np.int64)[sess_of_ev] + np.where(rng.random(NE) < 0.7, 0, rng.integers(0, 1 << 30, NE) % A_s[sess_of_ev])
    scores = rng.random(A)
    scores[rng.random(A) < 0.1] = 0.0
    scores[rng.random(A) < 0.02] = np.nan
    kind = rng.choice([0, 1, 2], E, p=[0.7, 0.2, 0.1])
    now = 1_800_000_000_000_000
    expiry = np.where(kind == 0, np.iinfo(np.int64).max, np.where(kind == 1, now + 10 ** 15, now - 10 ** 15)).astype(np.int64)
    return dict(rowptr=rowptr, voucher=voucher.astype(np.uint32), active=(rng.random(E) < 0.9).astype(np.uint8),
                expiry=expiry, now=now, session_agent_rowptr=srp, event_rowptr=erp, event_agent=ev_agent.astype(np.uint32),
                event_sigma=rng.random(NE), event_omega=rng.choice(recipes.SLASH_OMEGAS, NE),
                event_depth=rng.choice([0, 0, 0, 1, 2, 3], NE).astype(np.uint8), scores=scores)


def run_c(b):
    return co.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                            b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"],
                            b["event_depth"])


def test_large_batch_and_giant_session_vs_c_restatement(ctx, variant):
    b = big_batch(11, 30000, giant=(5000, 200000))
    got, want = run_gpu(ctx, b), run_c(b)
    assert int(want[6].sum()) > 100000 and int(want[5].max()) > 50          # real work, real cascades
    assert_same_outputs(got, want, b)
    # sessions around the staging limit (6 KB of state per warp): both sides of it in one batch
    for A, E in ((40, 330), (41, 330), (8, 354), (8, 355), (300, 100), (1, 361)):
        eb = big_batch(100 + A + E, 9, giant=(A, E))
        assert_same_outputs(run_gpu(ctx, eb), run_c(eb), eb)
    small = big_batch(12, 257)
    py = lo.slash_cascade(small["rowptr"], small["voucher"], small["active"], small["expiry"], small["now"],
                          small["session_agent_rowptr"], small["event_rowptr"], small["event_agent"], small["event_sigma"],
                          small["event_omega"], small["scores"], small["event_depth"])
    assert_same_outputs(run_gpu(ctx, small), py, small)


@pytest.mark.parametrize("slice_bytes", [0, 1024, 2064, 7040, 10752])
def test_lane_slice_sizes(ctx, slice_bytes):
    """k_slash_lanes with every size of its per-warp shared-memory slice (0 = chosen from the batch's average
    session): the runs of sessions a warp stages are cut differently, the records must not change.  The
    mid-size workload shape (32 agents x 256 records a session, where the auto choice is the 7040-byte slice) and a
    ragged batch whose sessions fall on both sides of every slice."""
    import workloads
    ctx.set_option("slash_slice", slice_bytes)
    try:
        mid = workloads.slash_sessions(600, 32, 256)
        mid["event_depth"] = None
        got = run_gpu(ctx, mid)
        want = co.slash_cascade(*[mid[k] for k in ("rowptr", "voucher", "active", "expiry", "now", "session_agent_rowptr",
                                                   "event_rowptr", "event_agent", "event_sigma", "event_omega", "scores")], None)
        assert int(want[6].sum()) > 30000
        assert_same_outputs(got, want, mid)
        b = big_batch(21, 4000, giant=(90, 700))
        assert_same_outputs(run_gpu(ctx, b), run_c(b), b)
    finally:
        ctx.set_option("slash_slice", 0)
    with pytest.raises(Exception):
        ctx.set_option("slash_slice", 1000)


def test_edge_shapes_and_refusals(ctx):
    from agent_hypervisor_b200 import _native as nat
    z32, z64 = np.zeros(0, np.uint32), np.zeros(0, np.uint64)
    # no sessions at all; sessions without agents, records or events
    o = ctx.slash_cascade(np.zeros(1, np.uint32), z32, np.zeros(0, np.uint8), np.zeros(0, np.int64), 0, np.zeros(1, np.uint64),
                          np.zeros(1, np.uint64), z32, np.zeros(0), np.zeros(0), np.zeros(0))
    assert len(o["slash_count"]) == 0
    o = ctx.slash_cascade(np.zeros(3, np.uint32), z32, np.zeros(0, np.uint8), np.zeros(0, np.int64), 0,
                          np.array([0, 0, 2, 2], np.uint64), np.array([0, 0, 1, 1], np.uint64), np.array([1], np.uint32),
                          np.array([0.5]), np.array([0.95]), np.array([0.7, 0.8]))
    assert list(o["slash_count"]) == [0, 1, 0] and list(o["clip_count"]) == [0, 0, 0]
    assert list(o["scores"]) == [0.7, 0.0] and list(o["touched"]) == [0, 1]
    # an event naming another session's agent, a record whose voucher is in another session: flagged, skipped
    o = ctx.slash_cascade(np.array([0, 1, 1], np.uint32), np.array([1], np.uint32), np.ones(1, np.uint8),
                          np.full(1, np.iinfo(np.int64).max), 0, np.array([0, 1, 2], np.uint64), np.array([0, 1, 2], np.uint64),
                          np.array([0, 0], np.uint32), np.array([0.5, 0.5]), np.array([0.5, 0.5]), np.array([0.9, 0.9]))
    assert list(o["status"]) == [nat.SLASH_BAD_VOUCHER, nat.SLASH_BAD_EVENT]
    assert list(o["scores"]) == [0.0, 0.9] and list(o["active"]) == [1]
    with pytest.raises(nat.NativeError):          # session_agent_rowptr must end at n_agents
        ctx.slash_cascade(np.zeros(3, np.uint32), z32, np.zeros(0, np.uint8), np.zeros(0, np.int64), 0,
                          np.array([0, 1], np.uint64), np.array([0, 0], np.uint64), z32, np.zeros(0), np.zeros(0), np.zeros(2))


def build_engine(case, session="s", engine=None):
    from agent_hypervisor_b200 import VouchingEngine, VouchRecord
    now = datetime.now(timezone.utc)
    ve = engine or VouchingEngine()
    for k, (a, v, active, ek) in enumerate(case["edges"]):
        exp = None if ek == 0 else (now + timedelta(days=3650) if ek == 1 else now - timedelta(days=3650))
        vid = f"vouch:{session}:{k}"
        ve._vouches[vid] = VouchRecord(vouch_id=vid, voucher_did=f"did:{a}", vouchee_did=f"did:{v}", session_id=session,
                                       bonded_sigma_pct=0.2, bonded_amount=0.1, expiry=exp, is_active=active)
    return ve


def check_engine_state(case, gold, ve, history, scores, session="s"):
    assert len(history) == len(gold["history"])
    for r, h in zip(history, gold["history"]):
        assert r.vouchee_did == f"did:{h['vouchee']}" and r.cascade_depth == h["depth"] and r.session_id == session
        assert same_f64(r.vouchee_sigma_before, fromhex(h["sigma_before"])) and r.vouchee_sigma_after == 0.0
        assert r.reason == h["reason"] and r.slash_id.startswith("slash:")
        assert len(r.voucher_clips) == len(h["clips"])
        for cl, hc in zip(r.voucher_clips, h["clips"]):
            assert cl.voucher_did == f"did:{hc[0]}" and cl.vouch_id == f"vouch:{session}:{hc[4]}"
            assert same_f64(cl.sigma_before, fromhex(hc[1])) and same_f64(cl.sigma_after, fromhex(hc[2]))
            assert same_f64(cl.risk_weight, fromhex(hc[3]))
    assert [k for k in scores] == [f"did:{a}" for a, _ in gold["scores"]]              # same keys, same order
    assert same_f64([scores[f"did:{a}"] for a, _ in gold["scores"]], [fromhex(v) for _, v in gold["scores"]])
    assert [int(ve._vouches[f"vouch:{session}:{k}"].is_active) for k in range(len(case["edges"]))] == gold["active"]


def test_dropin_slashing_engine_matches_reference_records(slashing_golden):
    """agent_hypervisor_b200.SlashingEngine.slash, call by call, against the real engine's history, clips,
    agent_scores dict (values and key order) and released bonds."""
    from agent_hypervisor_b200 import SlashingEngine
    for g in slashing_golden["cases"][:120]:
        c = recipes.random_slash_case(g["seed"])
        ve = build_engine(c)
        se = SlashingEngine(ve)
        scores = {f"did:{a}": v for a, v in c["scores"]}
        tops = []
        for (v, sig, om, dep) in c["events"]:
            r = se.slash(vouchee_did=f"did:{v}", session_id="s", vouchee_sigma=sig, risk_weight=om, reason="golden",
                         agent_scores=scores, cascade_depth=dep)
            tops.append(next(i for i, h in enumerate(se.history) if h is r))
        assert tops == g["top_level"]
        check_engine_state(c, g, ve, se.history, scores)


def test_dropin_slash_batch_many_sessions(slashing_golden):
    """slash_batch: 150 sessions of one vouching engine in one launch, requests interleaved across sessions."""
    from agent_hypervisor_b200 import SlashingEngine, SlashRequest, VouchingEngine
    gold = slashing_golden["cases"][100:250]
    cases = [recipes.random_slash_case(g["seed"]) for g in gold]
    ve = VouchingEngine()
    scores = []
    for i, c in enumerate(cases):
        build_engine(c, session=f"s{i}", engine=ve)
        scores.append({f"did:{a}": v for a, v in c["scores"]})
    se = SlashingEngine(ve)
    reqs, owner = [], []
    for k in range(3):                            # round-robin: event k of every session
        for i, c in enumerate(cases):
            if k < len(c["events"]):
                v, sig, om, dep = c["events"][k]
                reqs.append(SlashRequest(f"did:{v}", f"s{i}", sig, om, "golden", scores[i], dep)); owner.append(i)
    res = se.slash_batch(reqs)
    assert all(r is not None and r.session_id == f"s{i}" for r, i in zip(res, owner))
    hist = se.history
    # history is in request order; per session it is the reference's history
    pos = {id(r): j for j, r in enumerate(hist)}
    assert [pos[id(r)] for r in res] == sorted(pos[id(r)] for r in res)
    for i, (c, g) in enumerate(zip(cases, gold)):
        check_engine_state(c, g, ve, [h for h in hist if h.session_id == f"s{i}"], scores[i], session=f"s{i}")
    with pytest.raises(ValueError):               # one dict for two sessions: not independent
        shared = {}
        se.slash_batch([SlashRequest("did:0", "s0", 0.5, 0.5, "x", shared), SlashRequest("did:0", "s1", 0.5, 0.5, "x", shared)])
