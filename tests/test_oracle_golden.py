"""Pins BOTH oracle restatements (hashlib-Python and plain C) against the
golden vectors produced by the real reference (tests/golden/make_golden.py).
CPU only."""

import hashlib

import numpy as np
import pytest

import recipes
from helpers import chain_preimages, config2_session, hexes, pack_sessions
from oracle import audit_oracle as ao
from oracle import coracle as co
from oracle import liability_oracle as lo


def test_c_sha256_nist_and_kats(audit_golden):
    for v in audit_golden["nist"]:
        assert co.sha256(v["msg"].encode()).hex() == v["digest"]
    for v in audit_golden["sha_kats"]:
        m = recipes.kat_message(v["seed"], v["len"])
        assert co.sha256(m).hex() == v["digest"], v["len"]
        assert hashlib.sha256(m).hexdigest() == v["digest"]
    nk = audit_golden["node_kat"]
    assert co.sha256((nk["left"] + nk["right"]).encode()).hex() == nk["digest"]


def test_c_sha256_batch_random_lengths():
    rng = np.random.default_rng(7)
    lens = np.concatenate([np.arange(0, 200), rng.integers(0, 5000, 300)])
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    arena = rng.integers(0, 256, int(offs[-1]), dtype=np.uint8)
    got = co.sha256_batch(arena, offs)
    ref = ao.leaf_digests(arena, offs)
    assert np.array_equal(got, ref)
    assert co.count_blocks(offs) == sum((int(l) + 9 + 63) // 64 for l in lens)


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 33, 63, 64, 65, 125, 250, 1000])
def test_golden_session(audit_golden, n):
    g = audit_golden["golden_session"][str(n)]
    sess = chain_preimages(recipes.golden_session(n))
    assert hashlib.sha256(b"".join(p for p, _ in sess)).hexdigest() == g["preimages_sha"]
    arena, offs, par, kind, rp = pack_sessions([sess])
    # python oracle
    d = ao.leaf_digests(arena, offs)
    assert ao.merkle_root_hex(hexes(d)) == g["root"]
    if n:
        assert hexes(d)[-1] == g["last_leaf"]
    if "leaves" in g:
        assert hexes(d) == g["leaves"]
    v, fb, _ = ao.verify_chains(arena, offs, par, kind, rp)
    assert bool(v[0]) == g["verify"]
    # C oracle
    dc = co.sha256_batch(arena, offs)
    assert np.array_equal(dc, d)
    roots, empty = co.merkle_roots(dc, rp)
    if n == 0:
        assert empty[0] == 1 and g["root"] is None
    else:
        assert empty[0] == 0 and bytes(roots[0]).hex() == g["root"]
    vc, fbc, dd = co.verify_chains(arena, offs, par, kind, rp)
    assert bool(vc[0]) == g["verify"] and fbc[0] == -1
    assert np.array_equal(dd, d)


def test_golden_preimage_text(audit_golden):
    g = audit_golden["golden_session"]["2"]
    sess = chain_preimages(recipes.golden_session(2))
    assert sess[0][0].decode() == g["preimage0"]
    assert sess[1][0].decode() == g["preimage1"]
    assert len(sess[0][0]) == 254 and len(sess[1][0]) == 316


def test_bench_shape(audit_golden):
    g = audit_golden["bench_shape"]
    sess = chain_preimages(recipes.bench_shape_session(g["n"]))
    assert (len(sess[0][0]), len(sess[1][0]), len(sess[-1][0])) == (g["len0"], g["len1"], g["len_last"])
    arena, offs, par, kind, rp = pack_sessions([sess])
    d = co.sha256_batch(arena, offs)
    assert hexes(d)[0] == g["leaf0"] and hexes(d)[-1] == g["last_leaf"]
    roots, _ = co.merkle_roots(d, rp)
    assert bytes(roots[0]).hex() == g["root"]


def test_random_sessions_batched(audit_golden):
    """>= 1000 randomly shaped real SemanticDelta sessions."""
    rs = audit_golden["random_sessions"]
    sessions = []
    for r in rs:
        sess = chain_preimages(recipes.random_session(r["seed"]))
        assert len(sess) == r["n"]
        assert hashlib.sha256(b"".join(p for p, _ in sess)).hexdigest() == r["preimages_sha"], r["seed"]
        sessions.append(sess)
    arena, offs, par, kind, rp = pack_sessions(sessions)
    d = co.sha256_batch(arena, offs)
    roots, empty = co.merkle_roots(d, rp)
    pr, pe = ao.merkle_roots(d, rp)
    assert np.array_equal(roots, pr) and np.array_equal(empty, pe)
    v, fb, _ = co.verify_chains(arena, offs, par, kind, rp)
    for s, r in enumerate(rs):
        if r["n"] == 0:
            assert empty[s] == 1 and r["root"] is None
        else:
            assert bytes(roots[s]).hex() == r["root"], r["seed"]
            assert hexes(d[int(rp[s + 1]) - 1])[0] == r["last_leaf"]
        assert bool(v[s]) == r["verify"]


def _tampered(name, n):
    """Rebuild the tamper cases of make_golden.py at pre-image level."""
    fields = recipes.golden_session(n)
    clean = chain_preimages(fields)
    pre = [p for p, _ in clean]
    par = [q for _, q in clean]
    leaf = [hashlib.sha256(p).hexdigest() for p in pre]

    def re_ser(i, **kw):
        f = dict(fields[i]); f.update(kw)
        pre[i] = ao.canonical_preimage(f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
                                       f["timestamp"].isoformat(), f["changes"], par[i])

    def set_parent(i, val):
        par[i] = val
        re_ser(i)

    if name == "agent_mid": re_ser(5, agent_did="did:evil")
    elif name == "agent_last": re_ser(9, agent_did="did:evil")
    elif name == "agent_first": re_ser(0, agent_did="did:evil")
    elif name == "parent_first_set": set_parent(0, "0" * 64)
    elif name == "parent_none_mid": set_parent(4, None)
    elif name == "parent_other_mid": set_parent(4, "nothex")
    elif name == "parent_upper_mid": set_parent(4, leaf[3].upper())
    elif name == "parent_wrong_mid": set_parent(7, "f" * 64)
    elif name == "single_parent_set": set_parent(0, "x")
    stored = list(leaf)
    if name == "stored_hash_only":
        stored[3] = "deadbeef"
    return pre, par, stored


def test_verify_kats(audit_golden):
    for k in audit_golden["verify_kats"]:
        pre, par, stored = _tampered(k["name"], k["n"])
        verdict, first_bad, recomputed = ao.verify_chain_session(pre, par)
        assert verdict == k["verdict"], k["name"]
        # state parity: the reference overwrote delta_hash for indices <= first_bad (all if ok)
        after = list(stored)
        for i, h in enumerate(recomputed):
            after[i] = h
        assert after == k["hashes_after"], k["name"]
        assert ao.merkle_root_hex(after) == k["root_after"], k["name"]
        # arena-level forms agree (python + C)
        arena, offs, parents, kind, rp = pack_sessions([list(zip(pre, par))])
        v, fb, _ = ao.verify_chains(arena, offs, parents, kind, rp)
        vc, fbc, _ = co.verify_chains(arena, offs, parents, kind, rp)
        assert bool(v[0]) == k["verdict"] and bool(vc[0]) == k["verdict"], k["name"]
        assert fb[0] == first_bad and fbc[0] == first_bad, k["name"]


@pytest.mark.parametrize("s", [0, 1234, 99999999])
def test_config2_recipe(audit_golden, s):
    g = audit_golden["config2"][str(s)]
    sess = config2_session(s)
    arena, offs, par, kind, rp = pack_sessions([sess])
    assert np.all(np.diff(offs.astype(np.int64)) == 256)
    d = co.sha256_batch(arena, offs)
    assert hexes(d)[0] == g["leaf0"] and hexes(d)[-1] == g["last_leaf"]
    roots, _ = co.merkle_roots(d, rp)
    assert bytes(roots[0]).hex() == g["root"]


def test_commitment(audit_golden):
    c = audit_golden["commitment"]
    assert ao.commitment_verify(c["root"], c["root"]) == c["ok"]
    assert ao.commitment_verify(c["root"], "0" * 64) == c["wrong"]
    assert ao.commitment_verify(None, c["root"]) == c["unknown"]


# ------------------------------------------------------------------ liability
def graph_to_csr(g, A):
    """Insertion-ordered CSR by vouchee (what the host packer builds)."""
    rows = [[] for _ in range(A)]
    for (a, b, vs, pct, active, ek) in g["edges"]:
        rows[b].append((vs * pct, active, ek))
    rowptr = np.zeros(A + 1, dtype=np.uint32)
    bonded, act, exp = [], [], []
    now = 1_800_000_000_000_000
    for i, r in enumerate(rows):
        rowptr[i + 1] = rowptr[i] + len(r)
        for (bd, ac, ek) in r:
            bonded.append(bd); act.append(1 if ac else 0)
            exp.append(lo.EXPIRY_NONE if ek == 0 else (now + 10**15 if ek == 1 else now - 10**15))
    return (rowptr, np.array(bonded, dtype=np.float64), np.array(act, dtype=np.uint8),
            np.array(exp, dtype=np.int64), now)


def test_sigma_eff_graphs(liability_golden):
    for gr in liability_golden["graphs"]:
        g = recipes.random_vouch_graph(gr["seed"], gr["A"], gr["E"])
        rowptr, bonded, act, exp, now = graph_to_csr(g, gr["A"])
        args = (rowptr, bonded, act, exp, now, np.array(g["sigma"]), np.array(g["omega"]),
                np.array(g["consensus"], dtype=np.uint8), np.array(g["current_ring"], dtype=np.uint8))
        for impl in (lo.sigma_eff_rings, co.sigma_eff_rings):
            se, ring, dem, near = impl(*args)
            want = np.array([float.fromhex(x) for x in gr["sigma_eff"]])
            assert np.array_equal(se, want), (gr["seed"], impl.__module__)   # bit-identical
            assert list(ring) == gr["ring"]
            assert [bool(x) for x in dem] == gr["demote"]


def test_sigma_eff_pinned_points(liability_golden):
    # tests/unit/test_liability.py:39-67 : 0.4+0.5*0.18, capped 1.0, 0.3+0.5*(0.18+0.16)
    vals = [float.fromhex(p["value"]) for p in liability_golden["pinned_points"]]
    rowptr = np.array([0, 1, 2, 4], dtype=np.uint32)
    bonded = np.array([0.9 * 0.2, 1.0 * 0.8, 0.9 * 0.2, 0.8 * 0.2])
    ones = np.ones(4, np.uint8)
    exp = np.full(4, lo.EXPIRY_NONE, dtype=np.int64)
    for impl in (lo.sigma_eff_rings, co.sigma_eff_rings):
        se, *_ = impl(rowptr, bonded, ones, exp, 0, np.array([0.4, 0.9, 0.3]), np.array([0.5, 1.0, 0.5]),
                      np.zeros(3, np.uint8), np.full(3, 3, np.uint8))
        assert list(se) == vals
    assert abs(vals[0] - 0.49) < 1e-9 and vals[1] == 1.0 and abs(vals[2] - 0.47) < 1e-9


def _sig(x):
    return float("nan") if x == "nan" else float.fromhex(x)


def test_ring_grid_and_checks(liability_golden):
    for r in liability_golden["ring_grid"]:
        assert lo.ring_from_sigma(_sig(r["sigma"]), r["consensus"]) == r["ring"]
    rows = liability_golden["checks"]
    flags = np.array([(lo.ACT_ADMIN if c[0] else 0) | (lo.ACT_READ_ONLY if c[1] else 0)
                      | (lo.ACT_REV_NONE if c[2] == "none" else 0) for c in rows], dtype=np.uint8)
    ar = np.array([c[3] for c in rows], dtype=np.uint8)
    sg = np.array([_sig(c[4]) for c in rows])
    cons = np.array([c[5] for c in rows], dtype=np.uint8)
    wit = np.array([c[6] for c in rows], dtype=np.uint8)
    for impl in (lo.ring_check, co.ring_check):
        al, re, rq = impl(ar, flags, sg, cons, wit)
        assert list(al) == [c[7] for c in rows]
        assert list(rq) == [c[8] for c in rows]
        for code, c in zip(re, rows):
            assert (code == lo.REASON_NEEDS_CONSENSUS) == bool(c[10])
            assert (code == lo.REASON_NEEDS_WITNESS) == bool(c[11])
            assert (code == lo.REASON_GRANTED) == (c[9] == "Access granted")


def test_exposure_sums_match_python_sum_and_reference(liability_golden):
    """total_exposure / get_total_exposure are Python sum(): compensated since CPython 3.12."""
    import sys
    g = liability_golden["matrix"]
    names = list(g["exposure"])
    idx = {n: i for i, n in enumerate(names)}
    rows = [[] for _ in names]
    for a, b, amt, vid in g["edges"]:
        if a in idx:
            rows[idx[a]].append(float.fromhex(amt))
    rowptr = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.uint32)
    bonded = np.array([x for r in rows for x in r])
    ones = np.ones(len(bonded), np.uint8); none = np.full(len(bonded), lo.EXPIRY_NONE, np.int64)
    want_ref = np.array([float.fromhex(g["exposure"][n]) for n in names])      # produced by the reference on CPython 3.12
    for impl in (lo.exposure_sums, co.exposure_sums):
        assert np.array_equal(impl(rowptr, bonded, ones, none, 0, True), want_ref)
        plain = impl(rowptr, bonded, ones, none, 0, False)
        want_plain = []
        for r in rows:
            f = 0.0
            for x in r:
                f += x
            want_plain.append(f)
        assert np.array_equal(plain, np.array(want_plain))
    if sys.version_info >= (3, 12):
        assert [sum(r) for r in rows] == list(want_ref)
    # ill-conditioned rows where the two summations differ visibly
    rng = np.random.default_rng(0)
    vals = np.concatenate([[1e16, 1.0, -1e16, 1.0], rng.uniform(-1, 1, 500) * 10.0 ** rng.integers(-8, 8, 500)])
    rp = np.array([0, 4, len(vals)], np.uint32)
    o = np.ones(len(vals), np.uint8); nn = np.full(len(vals), lo.EXPIRY_NONE, np.int64)
    a = lo.exposure_sums(rp, vals, o, nn, 0, True); b = co.exposure_sums(rp, vals, o, nn, 0, True)
    assert np.array_equal(a, b) and a[0] == 2.0
    assert lo.exposure_sums(rp, vals, o, nn, 0, False)[0] == 1.0          # plain summation loses the first 1.0
    assert np.array_equal(lo.exposure_sums(rp, vals, o, nn, 0, False), co.exposure_sums(rp, vals, o, nn, 0, False))


def test_slashing_cases_match_reference(slashing_golden):
    """Python restatement of SlashingEngine.slash vs the real engine's history, clips, final scores and
    released bonds on 400 random sessions - each alone, and all of them as one batch."""
    from helpers import check_slash_against_golden, slash_batch
    gold = slashing_golden["cases"]
    cases = [recipes.random_slash_case(g["seed"]) for g in gold]
    assert float.fromhex(slashing_golden["sigma_floor"]) == lo.SIGMA_FLOOR
    assert slashing_golden["max_cascade_depth"] == lo.MAX_CASCADE_DEPTH
    assert sum(1 for g in gold if any(h["depth"] == 2 for h in g["history"])) > 50
    for c, g in list(zip(cases, gold))[:60]:
        b = slash_batch([c])
        out = lo.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                               b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"], b["event_depth"])
        check_slash_against_golden([c], [g], b, out)
    b = slash_batch(cases)
    out = lo.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                           b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"], b["event_depth"])
    check_slash_against_golden(cases, gold, b, out)
    out_c = co.slash_cascade(b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"],
                             b["event_rowptr"], b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"], b["event_depth"])
    check_slash_against_golden(cases, gold, b, out_c)       # the plain-C restatement, same records


def test_slashing_pinned_by_reference_tests(slashing_golden):
    """The four cases of the reference's tests/unit/test_slashing.py:14-89, through the restatement."""
    for p in slashing_golden["pinned"]:
        names = list(p["scores_in"])
        idx = {n: i for i, n in enumerate(names)}
        c = dict(n_agents=len(names), edges=[(idx[a], idx[b], True, 0) for a, b, _ in p["edges"]],
                 scores=[(idx[n], v) for n, v in p["scores_in"].items()],
                 events=[(idx[p["event"][0]], p["event"][1], p["event"][2], 0)])
        from helpers import fromhex, slash_batch
        b = slash_batch([c])
        sc, *_rest, cl, ns, nc = lo.slash_cascade(
            b["rowptr"], b["voucher"], b["active"], b["expiry"], b["now"], b["session_agent_rowptr"], b["event_rowptr"],
            b["event_agent"], b["event_sigma"], b["event_omega"], b["scores"])
        for n, v in p["scores_out"].items():
            assert sc[idx[n]] == fromhex(v)
        assert [[names[int(b["voucher"][int(q["edge"])])], float(q["sigma_before"]), float(q["sigma_after"])] for q in cl[: int(nc[0])]] \
            == [[n, fromhex(x), fromhex(y)] for n, x, y in p["clips"]]
    # and the literal numbers those tests assert
    assert abs(fromhex(slashing_golden["pinned"][1]["scores_out"]["voucher"]) - 0.45) < 1e-9
    assert abs(fromhex(slashing_golden["pinned"][3]["scores_out"]["v1"]) - 0.56) < 1e-9
    assert fromhex(slashing_golden["pinned"][2]["scores_out"]["voucher"]) >= 0.05
