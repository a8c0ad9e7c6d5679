"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the golden
vectors produced by the real reference.  Bit-exact for digests, roots, verdicts, rings;
sigma_eff bit-identical (well inside the 1e-6 relative tolerance north_star allows)."""

import hashlib

import numpy as np
import pytest

import recipes
import workloads
from helpers import SIGMA_LONG_CUT, chain_preimages, config2_session, hexes, pack_sessions, sigma_expected_default_path
from oracle import audit_oracle as ao
from oracle import coracle as co
from oracle import liability_oracle as lo
from agent_hypervisor_b200 import _native as nat

pytestmark = pytest.mark.gpu

SIGMA_REL_TOL = 1e-6   # BASELINE.json north_star; we additionally assert bit equality


@pytest.fixture(scope="module")
def ctx():
    from agent_hypervisor_b200 import get_context
    return get_context()


def _tight(msgs):
    offs = np.zeros(len(msgs) + 1, dtype=np.uint64)
    np.cumsum([len(m) for m in msgs], out=offs[1:])
    arena = np.frombuffer(b"".join(msgs), dtype=np.uint8) if offs[-1] else np.zeros(0, np.uint8)
    return arena, offs


# ------------------------------------------------------------------ C ABI error behaviour
def test_c_abi_error_codes(ctx):
    import ctypes as C
    from agent_hypervisor_b200 import NativeError
    lib, h = ctx.lib, ctx.h
    vp = C.c_void_p
    out = np.zeros(64, np.uint8)
    a = np.zeros(64, np.uint8)
    assert lib.ahv_sha256_batch(h, vp(a.ctypes.data), None, 1, 0, vp(out.ctypes.data)) == -1       # no offsets, no fixed_len
    assert b"offsets is NULL" in lib.ahv_last_error()
    assert lib.ahv_sha256_batch(h, vp(a.ctypes.data), None, 1, 64, None) == -1                      # null output
    assert lib.ahv_sha256_batch(None, vp(a.ctypes.data), None, 1, 64, vp(out.ctypes.data)) == -1    # null context
    assert lib.ahv_sha256_batch(h, None, None, 0, 0, None) == 0                                     # empty batch is fine
    rp = np.array([0, 2, 1], np.uint64)                                                             # decreasing rowptr
    assert lib.ahv_merkle_roots(h, vp(a.ctypes.data), vp(rp.ctypes.data), 2, vp(out.ctypes.data), None) == -1
    d = ctx.dev_alloc(256)
    try:
        with pytest.raises(NativeError, match="16-byte aligned"):
            ctx.sha256_batch_dev(d, 64, 0, 1, 64, d + 8)
        with pytest.raises(NativeError, match="n_sessions \\* uniform_leaves"):
            ctx.commit_dev(d, 128, 0, 64, 0, 2, 3, 1, 1, d, d, d)
    finally:
        ctx.dev_free(d)
    with pytest.raises(NativeError, match="unknown option"):
        ctx.set_option("no_such_option", 1)
    with pytest.raises(NativeError, match="top_max"):
        ctx.set_option("top_max", 1)
    c2 = C.c_void_p()
    assert lib.ahv_create(4096, C.byref(c2)) == -1 and b"out of range" in lib.ahv_last_error()


# ------------------------------------------------------------------ SHA-256
def test_sha256_known_answers(ctx, audit_golden):
    msgs = [v["msg"].encode() for v in audit_golden["nist"]]
    want = [v["digest"] for v in audit_golden["nist"]]
    for v in audit_golden["sha_kats"]:
        msgs.append(recipes.kat_message(v["seed"], v["len"]))
        want.append(v["digest"])
    nk = audit_golden["node_kat"]
    msgs.append((nk["left"] + nk["right"]).encode()); want.append(nk["digest"])
    arena, offs = _tight(msgs)           # tight packing -> every byte alignment occurs
    assert hexes(ctx.sha256_batch(arena, offs)) == want


@pytest.mark.parametrize("sort", [0, 1])
def test_sha256_random_lengths_vs_oracle(ctx, sort):
    rng = np.random.default_rng(11)
    lens = np.concatenate([np.arange(0, 260), rng.integers(0, 3000, 3000), [16383, 16384, 16385, 70000]])
    rng.shuffle(lens)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    arena = rng.integers(0, 256, int(offs[-1]), dtype=np.uint8)
    ctx.set_option("sort", sort)
    try:
        got = ctx.sha256_batch(arena, offs)
    finally:
        ctx.set_option("sort", 1)
    assert np.array_equal(got, co.sha256_batch(arena, offs))


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("flen", [64, 128, 256, 320, 384, 100, 55, 1])
def test_sha256_fixed_len(ctx, flen, variant):
    rng = np.random.default_rng(flen)
    n = 2500 + flen % 7            # partial last warp
    arena = rng.integers(0, 256, n * flen, dtype=np.uint8)
    ctx.set_option("fixed_variant", variant)
    try:
        got = ctx.sha256_batch(arena, None, fixed_len=flen, n=n)
    finally:
        ctx.set_option("fixed_variant", 0)
    offs = (np.arange(n + 1) * flen).astype(np.uint64)
    assert np.array_equal(got, co.sha256_batch(arena, offs))


def test_config3_shape_vs_oracle(ctx):
    """Config-3 shape (log-uniform 64 B..16 KB, tightly packed, every alignment), 2x10^5 messages = 0.6 GB:
    hash + 200 trees + verify against the C oracle, sorted and unsorted launch order."""
    n = 200_000
    arena, offs = workloads.config3_arena(n, seed=0xA6E27)
    want = co.sha256_batch(arena, offs)
    got = ctx.sha256_batch(arena, offs)
    assert np.array_equal(got, want)
    rp = (np.arange(201, dtype=np.uint64) * 1000)
    par = np.zeros((n, 32), np.uint8); par[1:] = want[:-1]
    kind = np.ones(n, np.uint8); kind[::1000] = 0
    par[123_456] ^= 1                                        # one broken link: session 123, position 456
    res = ctx.verify_chains(arena, offs, par, kind, rp, want_roots=True)
    wr, _ = co.merkle_roots(want, rp)
    assert np.array_equal(res["roots"], wr) and np.array_equal(res["digests"], want)
    assert list(np.nonzero(res["verdict"] == 0)[0]) == [123] and res["first_bad"][123] == 456
    lens = np.diff(offs.astype(np.int64))
    assert lens.min() >= 64 and lens.max() <= 16384 and 2000 < lens.mean() < 4000


def test_config3_one_million_messages_vs_oracle(ctx):
    """BASELINE configs[2] shape at 10^6 messages (2.9 GB, log-uniform 64 B..16 KB, tightly packed so every byte
    alignment occurs), everything against the C oracle: digests with the length-sorted and the unsorted launch
    order, 10^3 trees x 10^3 leaves and ONE 10^6-leaf tree, verify with two broken links."""
    n = 1_000_000
    arena, offs = workloads.config3_arena(n, seed=0xA6E27)
    want = co.sha256_batch(arena, offs)
    d_a, d_o, d_d = ctx.dev_alloc(arena.nbytes + 64), ctx.dev_alloc(offs.nbytes), ctx.dev_alloc(32 * n)
    try:
        ctx.h2d(d_a, arena); ctx.h2d(d_o, offs)
        got = np.empty((n, 32), np.uint8)
        for sort in (1, 0):
            ctx.set_option("sort", sort)
            ctx.memset(d_d, 0, 32 * n)
            ctx.sha256_batch_dev(d_a, arena.nbytes, d_o, n, 0, d_d)
            ctx.d2h(got, d_d); ctx.sync()
            assert np.array_equal(got, want), f"sort={sort}"
    finally:
        ctx.set_option("sort", 1)
        for p in (d_a, d_o, d_d):
            ctx.dev_free(p)
    rp = np.arange(1001, dtype=np.uint64) * np.uint64(1000)
    r1, _ = ctx.merkle_roots(want, rp)
    w1, _ = co.merkle_roots(want, rp)
    assert np.array_equal(r1, w1)
    r2, _ = ctx.merkle_roots(want, np.array([0, n], np.uint64))
    w2, _ = co.merkle_roots(want, np.array([0, n], np.uint64))
    assert np.array_equal(r2, w2)
    par = np.zeros((n, 32), np.uint8); par[1:] = want[:-1]
    kind = np.ones(n, np.uint8); kind[::1000] = 0
    par[654_321, 7] ^= 0x40; kind[999_999] = 2                       # session 654 at 321; the very last delta's parent "other"
    res = ctx.verify_chains(arena, offs, par, kind, rp, expected_roots=w1)
    v, fb, _ = co.verify_chains(arena, offs, par, kind, rp)
    assert np.array_equal(res["verdict"], v) and np.array_equal(res["first_bad"], fb) and res["root_match"].all()
    assert list(np.nonzero(v == 0)[0]) == [654, 999] and fb[654] == 321 and fb[999] == 999


def test_sha256_empty_batch(ctx):
    assert ctx.sha256_batch(np.zeros(0, np.uint8), np.zeros(1, np.uint64)).shape == (0, 32)


def test_sha256_one_matches_hashlib(ctx):
    """ahv_sha256_one (message inside the launch, digest through mapped host memory) at the padding edges, at
    its size limit and past it (where it falls back to the batch path)."""
    rng = np.random.default_rng(9)
    for n in (0, 1, 3, 55, 56, 57, 63, 64, 65, 119, 120, 127, 128, 255, 256, 330, 1000, 3583, 3584, 3585, 5000, 70000):
        m = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert ctx.sha256_one(m) == hashlib.sha256(m).digest(), n
    for _ in range(200):                          # back to back: the mapped result slot is reused every call
        m = rng.integers(0, 256, int(rng.integers(0, 600)), dtype=np.uint8).tobytes()
        assert ctx.sha256_one(m) == hashlib.sha256(m).digest()


def test_vfs_content_hashes(ctx):
    """session/sso.py:214-216 for a batch of file contents (incl. empty, unicode, large)."""
    from agent_hypervisor_b200 import content_hashes
    rng = np.random.default_rng(12)
    contents = ["", "hello", "héllo ☃ \U0001F600", "x" * 55, "y" * 64, "z" * 1_000_000]
    contents += ["".join(chr(int(c)) for c in rng.integers(32, 0x2000, int(n))) for n in rng.integers(0, 3000, 200)]
    assert content_hashes(contents) == [hashlib.sha256(c.encode()).hexdigest() for c in contents]
    assert content_hashes([]) == []


# ------------------------------------------------------------------ Merkle
@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 33, 63, 64, 65, 125, 250, 1000])
def test_golden_session_commit(ctx, audit_golden, n):
    g = audit_golden["golden_session"][str(n)]
    arena, offs, par, kind, rp = pack_sessions([chain_preimages(recipes.golden_session(n))])
    digs, roots, empty = ctx.commit(arena, offs, rp)
    if n == 0:
        assert empty[0] == 1 and g["root"] is None
        return
    assert empty[0] == 0 and roots[0].tobytes().hex() == g["root"]
    assert hexes(digs)[-1] == g["last_leaf"]
    if "leaves" in g:
        assert hexes(digs) == g["leaves"]
    r2, e2 = ctx.merkle_roots(digs, rp)           # tree alone over the same digests
    assert r2[0].tobytes().hex() == g["root"] and e2[0] == 0


@pytest.mark.parametrize("top_max,compress", [(32, 1), (256, 1), (2, 1), (64, 0), (128, 0)])
def test_merkle_sizes_vs_oracle(ctx, top_max, compress):
    """Every size through both tree-finish kernels (warp per session <=64, CTA per session <=256),
    the global level steps, and both compression formulations; ragged sessions."""
    ctx.set_option("top_max", top_max); ctx.set_option("compress", compress)
    try:
        _merkle_sizes(ctx)
    finally:
        ctx.set_option("top_max", 0); ctx.set_option("compress", 1)


def _merkle_sizes(ctx):
    rng = np.random.default_rng(5)
    sizes = list(range(0, 70)) + [127, 128, 129, 255, 256, 257, 258, 511, 512, 513, 1000, 1023, 1024, 1025, 2049, 5000]
    rng.shuffle(sizes)
    rp = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    digs = rng.integers(0, 256, (int(rp[-1]), 32), dtype=np.uint8)
    roots, empty = ctx.merkle_roots(digs, rp)
    wr, we = co.merkle_roots(digs, rp)
    assert np.array_equal(empty, we) and np.array_equal(roots, wr)
    # a subset checked against the hashlib restatement of delta.py:117-134 as well
    pr, pe = ao.merkle_roots(digs[: int(rp[40])], rp[:41])
    assert np.array_equal(roots[:40], pr) and np.array_equal(empty[:40], pe)


def test_merkle_many_small_sessions(ctx):
    """More sessions than the one-CTA row-pointer scan is used for (8 tiles of 4096): per-level node counts + the
    tiled scan, every global level down to width 2, the thread-per-session finish - ragged (0..12 leaves, empties
    included) and uniform - against the C oracle."""
    rng = np.random.default_rng(8)
    S = 40_001
    sizes = rng.integers(0, 13, S)
    rp = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    digs = rng.integers(0, 256, (int(rp[-1]), 32), dtype=np.uint8)
    roots, empty = ctx.merkle_roots(digs, rp)
    wr, we = co.merkle_roots(digs, rp)
    assert np.array_equal(empty, we) and np.array_equal(roots, wr)
    rp3 = (np.arange(S + 1, dtype=np.uint64) * 3)
    d3 = digs[: 3 * S]
    r3, e3 = ctx.merkle_roots(d3, rp3)
    w3, _ = co.merkle_roots(d3, rp3)
    assert np.array_equal(r3, w3) and not e3.any()


def test_merkle_single_big_tree(ctx):
    rng = np.random.default_rng(6)
    n = 100_003
    digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    rp = np.array([0, n], dtype=np.uint64)
    roots, _ = ctx.merkle_roots(digs, rp)
    wr, _ = co.merkle_roots(digs, rp)
    assert np.array_equal(roots, wr)


def test_random_sessions_commit_and_verify(ctx, audit_golden):
    rs = audit_golden["random_sessions"]
    sessions = [chain_preimages(recipes.random_session(r["seed"])) for r in rs]
    arena, offs, par, kind, rp = pack_sessions(sessions)
    digs, roots, empty = ctx.commit(arena, offs, rp)
    res = ctx.verify_chains(arena, offs, par, kind, rp, expected_roots=roots)
    for s, r in enumerate(rs):
        if r["n"] == 0:
            assert empty[s] == 1
            assert res["root_match"][s] == 0      # no commitment for an empty session
        else:
            assert roots[s].tobytes().hex() == r["root"], r["seed"]
            assert res["root_match"][s] == 1
        assert bool(res["verdict"][s]) == r["verify"] and res["first_bad"][s] == -1
    assert np.array_equal(res["digests"], digs)
    assert np.array_equal(digs, co.sha256_batch(arena, offs))


# ------------------------------------------------------------------ verify_chain semantics
def test_verify_kats(ctx, audit_golden):
    from test_oracle_golden import _tampered
    sessions, meta = [], []
    for k in audit_golden["verify_kats"]:
        pre, par, stored = _tampered(k["name"], k["n"])
        sessions.append(list(zip(pre, par)))
        meta.append((k, pre, par))
    arena, offs, parents, kind, rp = pack_sessions(sessions)
    res = ctx.verify_chains(arena, offs, parents, kind, rp)
    for s, (k, pre, par) in enumerate(meta):
        verdict, first_bad, _ = ao.verify_chain_session(pre, par)
        assert bool(res["verdict"][s]) == k["verdict"] == verdict, k["name"]
        assert res["first_bad"][s] == first_bad, k["name"]


def test_verify_random_tampering_vs_oracle(ctx):
    rng = np.random.default_rng(21)
    sessions = []
    for s in range(300):
        n = int(rng.integers(0, 30))
        sess = chain_preimages(recipes.golden_session(n))
        pre = [p for p, _ in sess]; par = [q for _, q in sess]
        for _ in range(int(rng.integers(0, 3))):
            if n == 0: break
            i = int(rng.integers(0, n)); what = int(rng.integers(0, 4))
            if what == 0: pre[i] = pre[i] + b" "
            elif what == 1: par[i] = None
            elif what == 2: par[i] = "e" * 64
            else: par[i] = "zz"
        sessions.append(list(zip(pre, par)))
    arena, offs, parents, kind, rp = pack_sessions(sessions)
    res = ctx.verify_chains(arena, offs, parents, kind, rp)
    v, fb, d = co.verify_chains(arena, offs, parents, kind, rp)
    assert np.array_equal(res["verdict"], v) and np.array_equal(res["first_bad"], fb)
    assert np.array_equal(res["digests"], d)


# ------------------------------------------------------------------ config 2 shape (fixed 256 B, uniform sessions)
@pytest.mark.parametrize("variant", [0, 1, 2])
def test_config2_small_vs_golden_and_oracle(ctx, audit_golden, variant):
    S, L = 40, 1000
    first = 1200
    arena = np.zeros(S * L * 256, np.uint8)
    offs, hexpos, rp = workloads.config2_fill(arena, first, S, L)
    digs, par, kind = ctx.chain_capture(arena, offs, hexpos, rp)          # patches arena in place
    g = audit_golden["config2"]["1234"]
    s = 1234 - first
    assert digs[s * L + L - 1].tobytes().hex() == g["last_leaf"]
    assert digs[s * L].tobytes().hex() == g["leaf0"]
    want_sess = config2_session(1234)
    assert arena[s * L * 256:(s + 1) * L * 256].tobytes() == b"".join(p for p, _ in want_sess)
    ctx.set_option("fixed_variant", variant)
    try:
        d2, roots, empty = ctx.commit(arena, None, None, fixed_len=256, uniform_leaves=L, n_sessions=S)
        res = ctx.verify_chains(arena, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S,
                                expected_roots=roots)
    finally:
        ctx.set_option("fixed_variant", 0)
    assert roots[s].tobytes().hex() == g["root"]
    assert np.array_equal(d2, digs) and np.array_equal(d2, co.sha256_batch(arena, offs))
    wr, _ = co.merkle_roots(d2, rp)
    assert np.array_equal(roots, wr)
    assert res["verdict"].all() and (res["first_bad"] == -1).all() and res["root_match"].all()
    # the generic (offsets + rowptr) path must agree with the fixed/uniform fast path
    d3, r3, _ = ctx.commit(arena, offs, rp)
    assert np.array_equal(d3, d2) and np.array_equal(r3, roots)
    # tamper: one byte in the middle of session 7, and the LAST delta of session 9 (not detectable)
    bad = arena.copy()
    bad[(7 * L + 500) * 256 + 20] ^= 1
    bad[(9 * L + L - 1) * 256 + 20] ^= 1
    res = ctx.verify_chains(bad, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S,
                            expected_roots=roots)
    assert list(np.nonzero(res["verdict"] == 0)[0]) == [7] and res["first_bad"][7] == 501
    assert list(np.nonzero(res["root_match"] == 0)[0]) == [7, 9]      # the root compare does see both


def test_audit_sessions_pipelined_and_ragged(ctx, audit_golden):
    """ahv_audit_sessions (commit + verify on one upload): pipelined uniform path and ragged path."""
    S, L = 1500, 100                              # 1500 sessions -> 2 pipelined chunks, last one partial
    arena = ctx.host_alloc(S * L * 256)
    offs, hexpos, rp = workloads.config2_fill(arena, 7, S, L)
    digs, par, kind = ctx.chain_capture(arena, offs, hexpos, rp)
    bad = np.array(arena)
    bad[(700 * L + 50) * 256 + 30] ^= 4           # a middle delta of session 700
    bad[(1499 * L + L - 1) * 256 + 30] ^= 4       # the last delta of the last session (links still fine)
    for data, bad_sessions in ((arena, []), (bad, [700])):
        o = ctx.audit_sessions(data, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S)
        want_d = co.sha256_batch(np.asarray(data), offs)
        want_r, _ = co.merkle_roots(want_d, rp)
        v, fb, _ = co.verify_chains(np.asarray(data), offs, par, kind, rp)
        assert np.array_equal(o["digests"], want_d) and np.array_equal(o["roots"], want_r)
        assert np.array_equal(o["verdict"], v) and np.array_equal(o["first_bad"], fb)
        assert list(np.nonzero(o["verdict"] == 0)[0]) == bad_sessions
        assert o["root_match"].all() and not o["empty"].any()      # compared with the roots of the same upload
    # against STORED commitments (the clean upload's roots): the tampered archive fails exactly where it differs,
    # including the last delta of session 1499 that verify_chain cannot see (audit/commitment.py:59-64)
    clean_r, _ = co.merkle_roots(co.sha256_batch(np.asarray(arena), offs), rp)
    o = ctx.audit_sessions(bad, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S, expected_roots=clean_r)
    assert list(np.nonzero(o["root_match"] == 0)[0]) == [700, 1499] and list(np.nonzero(o["verdict"] == 0)[0]) == [700]
    o = ctx.audit_sessions(arena, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S, expected_roots=clean_r)
    assert o["root_match"].all()
    ctx.host_free(arena)
    rs = audit_golden["random_sessions"][:300]
    sessions = [chain_preimages(recipes.random_session(r["seed"])) for r in rs]
    a2, o2, p2, k2, rp2 = pack_sessions(sessions)
    o = ctx.audit_sessions(a2, o2, p2, k2, rp2)
    for s, r in enumerate(rs):
        assert bool(o["empty"][s]) == (r["n"] == 0)
        if r["n"]:
            assert o["roots"][s].tobytes().hex() == r["root"] and o["root_match"][s] == 1
        assert o["verdict"][s] == 1 and o["first_bad"][s] == -1
    # ragged (unpipelined) path with stored commitments: golden roots, one of them wrong; an empty session never matches
    exp = np.zeros((len(rs), 32), np.uint8)
    for s, r in enumerate(rs):
        if r["n"]:
            exp[s] = np.frombuffer(bytes.fromhex(r["root"]), np.uint8)
    wrong = next(s for s, r in enumerate(rs) if r["n"] > 1)
    exp[wrong, 5] ^= 0x10
    o = ctx.audit_sessions(a2, o2, p2, k2, rp2, expected_roots=exp)
    assert [int(x) for x in o["root_match"]] == [1 if (r["n"] and s != wrong) else 0 for s, r in enumerate(rs)]
    with pytest.raises(ValueError):
        ctx.audit_sessions(a2, o2, p2, k2, rp2, expected_roots=exp[:-1])


def test_parents_read_from_preimages(ctx):
    """ahv_extract_parents_dev and the parent_pos form of ahv_audit_sessions: same side arrays / same
    results as the host-built parent_digests + parent_kind, for None, digest, non-digest and odd parents."""
    from datetime import datetime, timezone
    from agent_hypervisor_b200.audit.packing import ArenaBuilder, PARENT_POS_NONE
    from agent_hypervisor_b200.audit.delta import SemanticDelta
    t0 = datetime(2026, 1, 1, tzinfo=timezone.utc)
    h = "ab" * 32
    odd_parents = [None, h, h.upper(), h[:-1], h + "0", h[:-1] + "g", "", "x", h[:-1] + '"', "\u00e9" * 8,
                   "0" * 63 + "\u00e9", '", "parent_hash": "' + h]
    b = ArenaBuilder()
    rng = np.random.default_rng(5)
    for s in range(40):
        for i in range(int(rng.integers(0, 9))):
            ph = odd_parents[int(rng.integers(0, len(odd_parents)))]
            # ids that look like the parent field must not confuse anything: positions come from the emitter
            b.add_delta(SemanticDelta(delta_id='d", "parent_hash": "%s' % h if i % 3 == 0 else f"d{i}", turn_id=i,
                                      session_id=f"s{s}", agent_did="did:a", timestamp=t0, changes=[], parent_hash=ph))
        b.end_session()
    p = b.finish()
    assert p.parent_pos is not None and set(p.parent_kind) == {nat.PARENT_NONE, nat.PARENT_HEX, nat.PARENT_OTHER}
    assert np.array_equal(p.parent_pos == PARENT_POS_NONE, p.parent_kind == nat.PARENT_NONE)
    par, kind = ctx.extract_parents(p.arena, p.offsets, p.parent_pos)
    assert np.array_equal(kind, p.parent_kind) and np.array_equal(par, p.parents)
    # positions that point nowhere sensible are "other", never a read outside the pre-image
    wild = p.parent_pos.copy()
    wild[::2] = 0; wild[1::4] = 0xFFFFFFFE; wild[3::4] = (np.diff(p.offsets.astype(np.int64))[3::4] - 3).astype(np.uint32)
    _, k2 = ctx.extract_parents(p.arena, p.offsets, wild)
    assert set(k2) <= {nat.PARENT_OTHER}
    a = ctx.audit_sessions(p.arena, p.offsets, p.parents, p.parent_kind, p.session_rowptr)
    bpos = ctx.audit_sessions(p.arena, p.offsets, None, None, p.session_rowptr, parent_pos=p.parent_pos)
    v, fb, _ = co.verify_chains(p.arena, p.offsets, p.parents, p.parent_kind, p.session_rowptr)
    assert np.array_equal(bpos["verdict"], v) and np.array_equal(bpos["first_bad"], fb)
    for k in ("digests", "roots", "empty", "verdict", "first_bad", "root_match"):
        assert np.array_equal(a[k], bpos[k]), k
    # pipelined uniform path: positions instead of 33 bytes per delta, a broken link in session 900
    S, L = 1500, 64
    arena = ctx.host_alloc(S * L * 256)
    offs, hexpos, rp = workloads.config2_fill(arena, 3, S, L)
    _, par, kind = ctx.chain_capture(arena, offs, hexpos, rp)
    arena[(900 * L + 20) * 256 + 30] ^= 1
    want = ctx.audit_sessions(arena, None, par, kind, None, fixed_len=256, uniform_leaves=L, n_sessions=S)
    got = ctx.audit_sessions(arena, None, None, None, None, fixed_len=256, uniform_leaves=L, n_sessions=S,
                             parent_pos=hexpos)
    for k in ("digests", "roots", "empty", "verdict", "first_bad", "root_match"):
        assert np.array_equal(want[k], got[k]), k
    assert list(np.nonzero(got["verdict"] == 0)[0]) == [900] and got["first_bad"][900] == 21
    ctx.host_free(arena)
    with pytest.raises(nat.NativeError):
        ctx.audit_sessions(p.arena, p.offsets, None, None, p.session_rowptr)


def test_audit_sessions_ragged_pipelined(ctx):
    """ahv_audit_sessions on a batch large enough to be pipelined (>= 64 sessions, >= 8 MB) with variable-length
    messages and ragged sessions (empty ones included): chunked upload / per-chunk relative row pointers give
    the same digests, roots, verdicts and first-bad indices as the C restatement; then the same arena through
    the offsets + parent_pos form against the fixed-size fast path."""
    rng = np.random.default_rng(77)
    S = 2500
    counts = rng.integers(0, 41, S); counts[[0, 17, S - 1]] = 0
    rp = np.zeros(S + 1, np.uint64); rp[1:] = np.cumsum(counts)
    n = int(rp[-1])
    lens = rng.integers(1, 500, n)
    offs = np.zeros(n + 1, np.uint64); offs[1:] = np.cumsum(lens)
    arena = rng.integers(0, 256, int(offs[-1]), dtype=np.uint8)
    assert arena.nbytes >= (8 << 20)
    digs = co.sha256_batch(arena, offs)
    par = np.zeros((n, 32), np.uint8); par[1:] = digs[:-1]
    kind = np.ones(n, np.uint8)
    first = rp[:-1][counts > 0].astype(np.int64)
    kind[first] = 0; par[first] = 0                     # first delta of a session: parent None
    bad_sessions = [5, 1200, S - 2]
    for s in bad_sessions:                              # break one link in the middle of these sessions
        if counts[s] >= 3:
            par[int(rp[s]) + 2, 0] ^= 1
    o = ctx.audit_sessions(arena, offs, par, kind, rp)
    want_r, want_e = co.merkle_roots(digs, rp)
    v, fb, _ = co.verify_chains(arena, offs, par, kind, rp)
    assert np.array_equal(o["digests"], digs) and np.array_equal(o["roots"], want_r) and np.array_equal(o["empty"], want_e)
    assert np.array_equal(o["verdict"], v) and np.array_equal(o["first_bad"], fb)
    assert np.array_equal(o["root_match"], 1 - want_e)
    assert set(np.nonzero(o["verdict"] == 0)[0]) == {s for s in bad_sessions if counts[s] >= 3}
    # fixed-size content given as a ragged batch (offsets + row pointers + parent positions)
    S2, L = 1500, 64
    a2 = ctx.host_alloc(S2 * L * 256)
    o2, hexpos, rp2 = workloads.config2_fill(a2, 11, S2, L)
    ctx.chain_capture(a2, o2, hexpos, rp2)
    a2[(700 * L + 9) * 256 + 30] ^= 2
    fast = ctx.audit_sessions(a2, None, None, None, None, fixed_len=256, uniform_leaves=L, n_sessions=S2, parent_pos=hexpos)
    slow = ctx.audit_sessions(a2, o2, None, None, rp2, parent_pos=hexpos)
    for k in ("digests", "roots", "empty", "verdict", "first_bad", "root_match"):
        assert np.array_equal(fast[k], slow[k]), k
    assert list(np.nonzero(slow["verdict"] == 0)[0]) == [700] and slow["first_bad"][700] == 10
    ctx.host_free(a2)


def test_chain_capture_vs_python_chain(ctx):
    sessions = [recipes.golden_session(n) for n in (0, 1, 2, 5, 33)]
    blobs, offs, hexpos, rp = [], [0], [], [0]
    for fields in sessions:
        for i, f in enumerate(fields):
            pre = ao.canonical_preimage(f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
                                        f["timestamp"].isoformat(), f["changes"], None if i == 0 else "0" * 64)
            hexpos.append(0xFFFFFFFF if i == 0 else pre.index(b'"parent_hash": "') + 16)
            blobs.append(pre); offs.append(offs[-1] + len(pre))
        rp.append(len(hexpos))
    arena = np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()
    digs, par, kind = ctx.chain_capture(arena, np.array(offs, np.uint64), np.array(hexpos, np.uint32),
                                        np.array(rp, np.uint64))
    want = [chain_preimages(f) for f in sessions]
    assert arena.tobytes() == b"".join(p for s in want for p, _ in s)
    wa, wo, wp, wk, wr = pack_sessions(want)
    assert np.array_equal(par, wp) and np.array_equal(kind, wk)
    assert np.array_equal(digs, co.sha256_batch(wa, wo))


@pytest.mark.parametrize("prefix", [1, 0, 2])
def test_chain_capture_parent_field_anywhere(ctx, prefix):
    """Chained capture over ragged messages with the 64-character parent field at every kind of position: in the
    first block, straddling block boundaries, reaching the last byte, at all four byte alignments, absent in the
    middle of a chain - with and without the parallel pre-pass over the blocks in front of the field
    (option chain_prefix).  Checked against hashlib run link by link."""
    rng = np.random.default_rng(77 + prefix)
    S = 150
    lens, hps, rp = [], [], [0]
    for s in range(S):
        L = int(rng.integers(0, 12))
        for i in range(L):
            n = int(rng.choice([64, 65, 100, 119, 120, 127, 128, 191, 192, 256, 300, 332, 640, 1000, 2048]))
            if i == 0 or rng.random() < 0.1:
                hp = 0xFFFFFFFF
            else:
                hp = int(rng.choice([0, 1, 2, 3, 60, 63, 64, 100, n - 64, max(0, n - 65), int(rng.integers(0, n - 63))]))
                hp = min(hp, n - 64)
            lens.append(n); hps.append(hp)
        rp.append(len(lens))
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    arena = rng.integers(32, 127, int(offs[-1]), dtype=np.uint8)
    want = arena.copy()
    digs_w = np.zeros((len(lens), 32), np.uint8)
    for s in range(S):
        prev = None
        for i in range(rp[s], rp[s + 1]):
            o, n, hp = int(offs[i]), lens[i], hps[i]
            if i > rp[s] and hp != 0xFFFFFFFF:
                want[o + hp: o + hp + 64] = np.frombuffer(prev.hex().encode(), np.uint8)
            prev = hashlib.sha256(want[o: o + n].tobytes()).digest()
            digs_w[i] = np.frombuffer(prev, np.uint8)
    ctx.set_option("chain_prefix", prefix)
    try:
        got = arena.copy()
        digs, par, kind = ctx.chain_capture(got, offs, np.array(hps, np.uint32), np.array(rp, np.uint64))
    finally:
        ctx.set_option("chain_prefix", 1)
    assert np.array_equal(digs, digs_w)
    assert np.array_equal(got, want)
    is_first = np.zeros(len(lens), bool); is_first[np.array(rp[:-1])[np.diff(rp) > 0]] = True
    patched = (~is_first) & (np.array(hps, np.uint64) != 0xFFFFFFFF)
    assert np.array_equal(kind != 0, patched)


# ------------------------------------------------------------------ drop-in classes
def test_delta_engine_dropin(ctx):
    from agent_hypervisor_b200 import DeltaEngine, VFSChange
    from agent_hypervisor_b200.audit.packing import canonical_preimage
    e = DeltaEngine("session:test-audit")
    assert e.compute_merkle_root() is None and e.verify_chain()
    for i in range(10):
        d = e.capture("did:a", [VFSChange(path=f"/f{i}.txt", operation="add", content_hash=f"h{i}")])
        assert d.turn_id == i + 1
        assert d.delta_hash == hashlib.sha256(canonical_preimage(d)).hexdigest()
    ds = e.deltas
    assert ds[0].parent_hash is None and all(ds[i].parent_hash == ds[i - 1].delta_hash for i in range(1, 10))
    root = e.compute_merkle_root()
    assert root == ao.merkle_root_hex([d.delta_hash for d in ds]) and len(root) == 64
    assert e.verify_chain() and e.turn_count == 10
    assert e.compute_merkle_root() == root                            # idempotent (e2e test :443-447)
    e._deltas[9].agent_did = "did:evil"                               # last delta: NOT detected
    assert e.verify_chain()
    e._deltas[5].agent_did = "did:tampered"                           # e2e test :421-423
    stale6 = e._deltas[6].delta_hash
    assert not e.verify_chain()
    assert e._deltas[5].delta_hash == hashlib.sha256(canonical_preimage(e._deltas[5])).hexdigest()
    assert e._deltas[6].delta_hash == hashlib.sha256(canonical_preimage(e._deltas[6])).hexdigest()  # idx == first_bad
    assert stale6 == e._deltas[6].delta_hash                           # delta 6 itself was not changed
    # a delta_hash overwritten with a non-digest string still yields the reference's root
    e2 = DeltaEngine("s2")
    for i in range(5):
        e2.capture("did:a", [])
    e2._deltas[2].delta_hash = "deadbeef"
    assert e2.compute_merkle_root() == ao.merkle_root_hex([d.delta_hash for d in e2._deltas])
    e3 = DeltaEngine("s3"); e3.capture("did:a", [])
    assert e3.compute_merkle_root() == e3._deltas[0].delta_hash


def test_batch_auditor(ctx):
    from agent_hypervisor_b200 import BatchAuditor, CommitmentEngine, DeltaEngine, VFSChange
    engines = []
    for s in range(20):
        e = DeltaEngine(f"session:{s}")
        for i in range(s):
            e.capture("did:a", [VFSChange(path=f"/f{i}", operation="add")])
        engines.append(e)
    ba, ce = BatchAuditor(), CommitmentEngine()
    roots = ba.commit_engines(engines, ce, ["did:a"])
    for e, r in zip(engines, roots):
        assert r == ao.merkle_root_hex([d.delta_hash for d in e._deltas])
        assert (ce.get_commitment(e.session_id) is None) == (r is None)
    engines[7]._deltas[3].agent_did = "did:evil"
    assert ba.verify_engines(engines) == [s != 7 for s in range(20)]
    res = ba.verify(ba.pack(engines), roots)
    assert [bool(x) for x in res["root_match"]] == [s not in (0, 7) for s in range(20)]


# ------------------------------------------------------------------ liability / rings
def test_sigma_eff_golden_graphs(ctx, liability_golden):
    from test_oracle_golden import graph_to_csr
    for gr in liability_golden["graphs"]:
        g = recipes.random_vouch_graph(gr["seed"], gr["A"], gr["E"])
        rowptr, bonded, act, exp, now = graph_to_csr(g, gr["A"])
        want = np.array([float.fromhex(x) for x in gr["sigma_eff"]])
        # default path: rows of up to SIGMA_LONG_CUT records bit-identical to the real engine, longer ones within 1e-6
        se, ring, dem, near = ctx.sigma_eff_rings(rowptr, bonded, act, exp, now, g["sigma"], g["omega"],
                                                  g["consensus"], g["current_ring"])
        short = np.diff(rowptr.astype(np.int64)) <= SIGMA_LONG_CUT
        assert np.allclose(se, want, rtol=SIGMA_REL_TOL, atol=0) and np.array_equal(se[short], want[short])
        listed = near != 0
        assert [r for r, n in zip(ring, listed) if not n] == [r for r, n in zip(gr["ring"], listed) if not n]
        # exact mode: every row in the reference's order
        ctx.set_option("sigma_exact", 1)
        try:
            se, ring, dem, near = ctx.sigma_eff_rings(rowptr, bonded, act, exp, now, g["sigma"], g["omega"],
                                                      g["consensus"], g["current_ring"])
        finally:
            ctx.set_option("sigma_exact", 0)
        assert np.array_equal(se, want)
        assert list(ring) == gr["ring"] and [bool(x) for x in dem] == gr["demote"]


def _check_sigma_default_path(ctx, args, want):
    """The default sigma_eff path against the oracle: rows up to the cut bit-identical; longer rows within the
    1e-6 relative tolerance of north_star AND bit-identical to the fixed reduction shape the kernel documents;
    rings / demotion equal to the reference's wherever the agent is not on the near-threshold list."""
    got = ctx.sigma_eff_rings(*args)
    exp = sigma_expected_default_path(want, *args)
    long_rows = exp[4]
    short = np.ones(len(want[0]), bool); short[long_rows] = False
    assert np.array_equal(got[0][short], want[0][short], equal_nan=True)
    fin = np.isfinite(want[0])
    rel = np.abs(got[0][fin] - want[0][fin]) / np.maximum(np.abs(want[0][fin]), 1e-300)
    assert rel.max() <= SIGMA_REL_TOL
    assert np.array_equal(got[0], exp[0], equal_nan=True)          # the documented shape, bit for bit
    for g_, e_ in zip(got[1:], exp[1:4]):
        assert np.array_equal(g_, e_)
    listed = (got[3] != 0) | (want[3] != 0)
    assert np.array_equal(got[1][~listed], want[1][~listed]) and np.array_equal(got[2][~listed], want[2][~listed])
    return got, long_rows, float(rel.max())


@pytest.mark.parametrize("mode", ["default", "stream", "exact", "row_per_thread"])
def test_sigma_eff_config4_shape_vs_oracle(ctx, mode):
    g = workloads.config4_graph(50_000, 500_000)
    # a few very long rows (beyond one shared-memory chunk) and NaN / inf / -0.0 amounts
    g["bonded"][:20] = [float("nan"), float("inf"), -0.0, 0.0, 1e308, 1e308, -1e308, 5e-324, 1.0, -1.0] * 2
    args = (g["rowptr"], g["bonded"], g["active"], g["expiry"], g["now"], g["sigma"], g["omega"], g["consensus"],
            g["current_ring"])
    want = co.sigma_eff_rings(*args)
    if mode in ("default", "stream"):
        ctx.set_option("sigma_variant", 2 if mode == "stream" else 1)
        try:
            _, long_rows, _ = _check_sigma_default_path(ctx, args, want)
        finally:
            ctx.set_option("sigma_variant", 1)
        assert len(long_rows) > 10
        return
    ctx.set_option("sigma_exact", 1)
    ctx.set_option("sigma_variant", 0 if mode == "row_per_thread" else 1)
    try:
        got = ctx.sigma_eff_rings(*args)
    finally:
        ctx.set_option("sigma_variant", 1)
        ctx.set_option("sigma_exact", 0)
    assert np.array_equal(got[0], want[0], equal_nan=True)  # sequential fp64, no FMA: bit-identical
    for a, b in zip(got[1:], want[1:]):
        assert np.array_equal(a, b)                        # rings, demote flags, near-threshold list


def test_sigma_eff_long_and_empty_rows(ctx):
    """Rows far longer than one shared-memory chunk (2048 edges), rows around the parallel-sum cut, several
    long rows inside one chunk, empty rows, a partial last CTA."""
    rng = np.random.default_rng(4)
    deg = np.concatenate([[5000, 0, 0, 3000, 1, 2047, 2048, 2049, 0, 9000, 512, 513, 514, 600, 520, 0, 513, 767, 768, 769, 31, 32, 33, 34, 64, 65, 15, 16, 17],
                          rng.integers(0, 30, 700), [513, 40000, 513]])
    A = len(deg)
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint32)
    E = int(rowptr[-1])
    bonded = rng.uniform(0, 1e-3, E)
    active = (rng.random(E) < 0.8).astype(np.uint8)
    expiry = np.where(rng.random(E) < 0.5, lo.EXPIRY_NONE, rng.integers(0, 2000, E)).astype(np.int64)
    args = (rowptr, bonded, active, expiry, 1000, rng.random(A), rng.choice([0.2, 0.65, 0.95], A),
            (rng.random(A) < 0.5).astype(np.uint8), rng.integers(1, 4, A).astype(np.uint8))
    want = co.sigma_eff_rings(*args)
    _check_sigma_default_path(ctx, args, want)
    for cut, chunk in ((4, 2048), (32, 1296), (100, 1024), (512, 1536), (4096, 1440)):      # other cuts / chunk sizes: same contract
        ctx.set_option("sigma_long_cut", cut)
        ctx.set_option("sigma_chunk", chunk)
        try:
            got = ctx.sigma_eff_rings(*args)
        finally:
            ctx.set_option("sigma_long_cut", SIGMA_LONG_CUT)
            ctx.set_option("sigma_chunk", 1440)
        exp = sigma_expected_default_path(want, *args, cut=cut)
        assert np.array_equal(got[0], exp[0]) and all(np.array_equal(a, b) for a, b in zip(got[1:], exp[1:4]))
    # rows per tile (the host entry picks 16 for this graph's average row of ~100 records): every size, same bits
    exp = sigma_expected_default_path(want, *args)
    for rows in (1, 2, 8, 32, 64, 128):
        ctx.set_option("sigma_rows", rows)
        try:
            got = ctx.sigma_eff_rings(*args)
            ctx.set_option("sigma_exact", 1)
            got_exact = ctx.sigma_eff_rings(*args)
        finally:
            ctx.set_option("sigma_exact", 0)
            ctx.set_option("sigma_rows", 0)
        assert np.array_equal(got[0], exp[0]) and all(np.array_equal(a, b) for a, b in zip(got[1:], exp[1:4])), rows
        assert all(np.array_equal(a, b) for a, b in zip(got_exact, want)), rows
    with pytest.raises(Exception):
        ctx.set_option("sigma_rows", 3)
    ctx.set_option("sigma_exact", 1)
    try:
        for variant in (1, 0, 2):
            ctx.set_option("sigma_variant", variant)
            got = ctx.sigma_eff_rings(*args)
            for a, b in zip(got, want):
                assert np.array_equal(a, b)
    finally:
        ctx.set_option("sigma_variant", 1)
        ctx.set_option("sigma_exact", 0)
    # the streaming form (persistent CTAs, two-stage pipeline): same sums, same bits as the tiled one
    ctx.set_option("sigma_variant", 2)
    try:
        for _ in range(3):                                   # consecutive launches alternate between two tile counters
            got = ctx.sigma_eff_rings(*args)
            exp = sigma_expected_default_path(want, *args)
            assert np.array_equal(got[0], exp[0]) and all(np.array_equal(a, b) for a, b in zip(got[1:], exp[1:4]))
    finally:
        ctx.set_option("sigma_variant", 1)


def test_sigma_eff_config4_full_size(ctx):
    """BASELINE configs[3] at its stated size: 10^6 agents, 10^7 vouch records (Zipf-like in-degree, longest row
    14 017), against the C oracle: exact mode bit-identical everywhere; default mode as _check_sigma_default_path."""
    g = workloads.config4_graph(1_000_000, 10_000_000)
    args = (g["rowptr"], g["bonded"], g["active"], g["expiry"], g["now"], g["sigma"], g["omega"], g["consensus"],
            g["current_ring"])
    assert len(g["sigma"]) == 1_000_000 and int(g["rowptr"][-1]) == 10_000_000
    want = co.sigma_eff_rings(*args)
    _, long_rows, rel = _check_sigma_default_path(ctx, args, want)
    assert len(long_rows) > 300 and int(np.diff(g["rowptr"].astype(np.int64)).max()) > 10_000
    ctx.set_option("sigma_exact", 1)
    try:
        got = ctx.sigma_eff_rings(*args)
    finally:
        ctx.set_option("sigma_exact", 0)
    assert np.array_equal(got[0], want[0], equal_nan=True)
    for a, b in zip(got[1:], want[1:]):
        assert np.array_equal(a, b)


def test_liability_matrix_and_exposure(ctx, liability_golden):
    from agent_hypervisor_b200 import LiabilityMatrix, VouchingEngine
    g = liability_golden["matrix"]
    m = LiabilityMatrix("s")
    for a, b, amt, vid in g["edges"]:
        m.add_edge(a, b, float.fromhex(amt), vid)
    names = list(g["exposure"])
    got = m.total_exposure_batch(names)
    for i, nm in enumerate(names):
        want = float.fromhex(g["exposure"][nm])
        assert got[i] == want and float(m.total_exposure(nm)) == want, nm
    assert m.total_exposure("did:nobody") == 0
    rng = np.random.default_rng(0)
    deg = np.concatenate([[4, 0, 3000, 1], rng.integers(0, 40, 600)])
    rp = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint32)
    vals = rng.uniform(-1, 1, int(rp[-1])) * 10.0 ** rng.integers(-8, 8, int(rp[-1]))
    vals[:4] = [1e16, 1.0, -1e16, 1.0]
    act = (rng.random(len(vals)) < 0.9).astype(np.uint8); act[:4] = 1
    exp = np.where(rng.random(len(vals)) < 0.7, lo.EXPIRY_NONE, rng.integers(0, 2000, len(vals))).astype(np.int64)
    exp[:4] = lo.EXPIRY_NONE
    for comp in (True, False):                      # CPython >= 3.12 sum() and the plain loop
        gpu = ctx.exposure_sums(rp, vals, act, exp, 1000, compensated=comp)
        assert np.array_equal(gpu, co.exposure_sums(rp, vals, act, exp, 1000, comp))
        assert gpu[0] == (2.0 if comp else 1.0)
    # VouchingEngine.get_total_exposure for many vouchers == the per-voucher host scan
    ve = VouchingEngine(max_exposure=1.0)
    for k in range(60):
        ve.vouch(f"did:h{k % 7}", f"did:l{k}", "s1", 0.9, bond_pct=0.01 + 0.001 * k)
    list(ve._vouches.values())[5].is_active = False
    vs = [f"did:h{i}" for i in range(8)]
    batch = ve.get_total_exposure_batch(vs, "s1")
    for i, v in enumerate(vs):
        assert batch[i] == float(ve.get_total_exposure(v, "s1"))
    # sigma_eff straight from the matrix's vouchee-grouped CSR
    agents = [f"did:{i}" for i in range(40)]
    rowptr, bonded, active, expiry = m.to_vouch_csr(agents)
    se, *_ = ctx.sigma_eff_rings(rowptr, bonded, active, expiry, 0, np.full(40, 0.1), np.full(40, 0.5),
                                 np.zeros(40, np.uint8), np.full(40, 3, np.uint8))
    want, *_ = co.sigma_eff_rings(rowptr, bonded, active, expiry, 0, np.full(40, 0.1), np.full(40, 0.5),
                                  np.zeros(40, np.uint8), np.full(40, 3, np.uint8))
    assert np.array_equal(se, want)


def test_ring_check_grid_and_dropin(ctx, liability_golden):
    from agent_hypervisor_b200 import ActionDescriptor, ExecutionRing, ReversibilityLevel, RingEnforcer, VouchingEngine
    rows = liability_golden["checks"]
    sig = lambda x: float("nan") if x == "nan" else float.fromhex(x)
    flags = [(1 if c[0] else 0) | (2 if c[1] else 0) | (4 if c[2] == "none" else 0) for c in rows]
    al, re, rq = ctx.ring_check([c[3] for c in rows], flags, [sig(c[4]) for c in rows], [c[5] for c in rows],
                                [c[6] for c in rows])
    assert list(al) == [c[7] for c in rows] and list(rq) == [c[8] for c in rows]
    enf = RingEnforcer()
    for c in rows[::37]:
        act = ActionDescriptor("a", "n", "/x", reversibility=ReversibilityLevel(c[2]), is_read_only=bool(c[1]),
                               is_admin=bool(c[0]))
        r = enf.check(ExecutionRing(c[3]), act, sig(c[4]), bool(c[5]), bool(c[6]))
        assert (int(r.allowed), int(r.required_ring.value), r.reason, int(r.requires_consensus),
                int(r.requires_sre_witness)) == (c[7], c[8], c[9], c[10], c[11])
        assert act.required_ring.value == c[8]
    for r in liability_golden["ring_grid"]:
        assert enf.compute_ring(sig(r["sigma"]), r["consensus"]).value == r["ring"]
    assert enf.should_demote(ExecutionRing.RING_1_PRIVILEGED, 0.99)        # computed without consensus
    assert not enf.should_demote(ExecutionRing.RING_2_STANDARD, 0.85)
    # reference's own pinned points (tests/unit/test_liability.py:39-67)
    pts = [float.fromhex(p["value"]) for p in liability_golden["pinned_points"]]
    ve = VouchingEngine(); ve.vouch("did:h", "did:l", "s1", 0.9, bond_pct=0.2)
    assert ve.compute_sigma_eff("did:l", "s1", 0.4, 0.5) == pts[0]
    ve = VouchingEngine(); ve.vouch("did:h", "did:l", "s1", 1.0, bond_pct=0.8)
    assert ve.compute_sigma_eff("did:l", "s1", 0.9, 1.0) == pts[1] == 1.0
    ve = VouchingEngine(); ve.vouch("did:h1", "did:l", "s1", 0.9, bond_pct=0.2); ve.vouch("did:h2", "did:l", "s1", 0.8, bond_pct=0.2)
    assert ve.compute_sigma_eff("did:l", "s1", 0.3, 0.5) == pts[2]
    out = ve.compute_sigma_eff_batch("s1", ["did:l", "did:nobody"], [0.3, 0.6], [0.5, 0.5])
    assert out["sigma_eff"][0] == pts[2] and out["sigma_eff"][1] == 0.6 and out["near_threshold"] == ["did:nobody"]
    # the same DID asked for twice: both entries get its vouch rows (as two calls of the reference would)
    out = ve.compute_sigma_eff_batch("s1", ["did:l", "did:x", "did:l"], [0.3, 0.1, 0.3], [0.5, 0.5, 0.5])
    assert out["sigma_eff"][0] == pts[2] and out["sigma_eff"][2] == pts[2] and out["sigma_eff"][1] == 0.1
    # a drop-in row longer than the fast path's cut stays in the reference's order (VouchingEngine.exact_order)
    ve = VouchingEngine(max_exposure=1.0)
    amounts = []
    for k in range(80):
        r = ve.vouch(f"did:h{k}", "did:many", "s1", 0.5 + 0.005 * k, bond_pct=0.001 + 0.0001 * k)
        amounts.append(r.bonded_amount)
    total = 0.0
    for x in amounts:
        total += x                                                        # vouching.py:146-148
    assert ve.compute_sigma_eff("did:many", "s1", 0.01, 0.9) == min(0.01 + (0.9 * total), 1.0)


def test_sharded_single_tree_matches_whole_tree(ctx):
    """one big tree cut at 2^k-aligned leaves across 8 emulated ranks (sub-roots, lifted tail,
    top finish) gives the same root as hashing the whole tree, on the CUDA path."""
    from agent_hypervisor_b200 import sharding
    rng = np.random.default_rng(8)
    n = 1_000_003
    digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    whole, _ = ctx.merkle_roots(digs, np.array([0, n], np.uint64))
    want, _ = co.merkle_roots(digs, np.array([0, n], np.uint64))
    assert np.array_equal(whole, want)
    for world, k in ((8, None), (3, 10), (8, 17)):
        plan = sharding.plan_tree(n, world, k)
        subs = [sharding.local_subroots(digs[(lo << plan["k"]): min(hi << plan["k"], n)], lo << plan["k"], plan, r)
                for r, (lo, hi) in enumerate(plan["ranges"])]
        assert sharding.finish_tree(np.concatenate(subs)) == bytes(want[0]), (world, k)


# ------------------------------------------------------------------ full size (BASELINE config 2), properties
@pytest.mark.parametrize("S", [10_000, 12_500])
def test_config2_full_size_properties(ctx, S):
    """BASELINE configs[1] (10k sessions x 1k x 256 B) and the per-GPU share of configs[4] (12 500 sessions of the 10^8
    deltas over 8 GPUs) on the device: chain -> commit -> verify, checked through size-independent properties plus
    a sample of sessions against the oracle."""
    L, FL = 1000, 256
    n = S * L
    arena = np.empty(n * FL, np.uint8)
    offs, hexpos, rp = workloads.config2_fill(arena, 0, S, L)
    d = {k: ctx.dev_alloc(v) for k, v in dict(arena=arena.nbytes, offs=offs.nbytes, hexpos=hexpos.nbytes,
                                               rp=rp.nbytes, digs=32 * n, par=32 * n, kind=n, roots=32 * S,
                                               roots2=32 * S, empty=S, verdict=S, fb=8 * S, match=S, digs2=32 * n).items()}
    try:
        ctx.h2d(d["arena"], arena); ctx.h2d(d["offs"], offs); ctx.h2d(d["hexpos"], hexpos); ctx.h2d(d["rp"], rp)
        ctx.chain_capture_dev(d["arena"], arena.nbytes, d["offs"], d["hexpos"], d["rp"], S, n, d["digs"], d["par"], d["kind"])
        ctx.commit_dev(d["arena"], arena.nbytes, 0, FL, 0, S, n, L, L, d["digs2"], d["roots"], d["empty"])
        ctx.verify_chains_dev(d["arena"], arena.nbytes, 0, FL, d["par"], d["kind"], 0, S, n, L, L, d["roots"],
                              d["verdict"], d["fb"], 0, d["roots2"], d["match"])
        # generic path (offsets + rowptr, length-sorted var kernel, ragged tree) on the same data
        roots3 = ctx.dev_alloc(32 * S)
        ctx.commit_dev(d["arena"], arena.nbytes, d["offs"], 0, d["rp"], S, n, 0, L, 0, roots3, d["empty"])
        h = {k: np.empty(sz, np.uint8) for k, sz in dict(digs=32 * n, digs2=32 * n, roots=32 * S, roots2=32 * S,
                                                         verdict=S, match=S, roots3=32 * S).items()}
        fb = np.empty(S, np.int64)
        for k in ("digs", "digs2", "roots", "roots2", "verdict", "match"):
            ctx.d2h(h[k], d[k])
        ctx.d2h(h["roots3"], roots3); ctx.d2h(fb, d["fb"]); ctx.d2h(arena, d["arena"])
        ctx.sync()
        ctx.dev_free(roots3)
    finally:
        for p in d.values():
            ctx.dev_free(p)
    assert np.array_equal(h["digs"], h["digs2"])                    # capture digests == commit digests
    assert h["verdict"].all() and (fb == -1).all() and h["match"].all()
    assert np.array_equal(h["roots"], h["roots2"]) and np.array_equal(h["roots"], h["roots3"])
    assert len({h["roots"][i * 32:(i + 1) * 32].tobytes() for i in range(S)}) == S   # all sessions distinct
    rng = np.random.default_rng(1)
    for s in rng.integers(0, S, 12):                                # sample vs the C oracle
        sl = slice(int(s) * L * FL, (int(s) + 1) * L * FL)
        o = (np.arange(L + 1) * FL).astype(np.uint64)
        dg = co.sha256_batch(arena[sl], o)
        assert np.array_equal(dg.reshape(-1), h["digs"][int(s) * L * 32:(int(s) + 1) * L * 32])
        r, _ = co.merkle_roots(dg, np.array([0, L], np.uint64))
        assert np.array_equal(r[0], h["roots"][int(s) * 32:(int(s) + 1) * 32])


# ------------------------------------------------------------------ guards and end-of-allocation reads
def test_understated_max_session_leaves_is_reported_not_corrupting(ctx):
    """A _dev tree entry sizes its in-CTA finish from the caller's max_session_leaves.  Too small a value used to
    overrun the finish kernels' shared-memory level; now the sessions that do not fit get zero roots, the others
    stay right, and the next ahv_sync reports it once."""
    from agent_hypervisor_b200 import NativeError
    rng = np.random.default_rng(3)
    for sizes, claimed in (([1000, 3, 70, 500, 64], 64), ([300] * 20000 + [50000], 300), ([100000, 2], 4)):
        rp = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
        n, S = int(rp[-1]), len(sizes)
        digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        want, _ = co.merkle_roots(digs, rp)
        d_dig, d_rp, d_roots, d_empty = ctx.dev_alloc(32 * n), ctx.dev_alloc(rp.nbytes), ctx.dev_alloc(32 * S), ctx.dev_alloc(S)
        try:
            ctx.h2d(d_dig, digs); ctx.h2d(d_rp, rp)
            for entry in (ctx.merkle_roots_dev, ctx.merkle_roots_noalloc_dev):
                if entry == ctx.merkle_roots_noalloc_dev and n > 200_000:
                    continue
                ctx.memset(d_roots, 0xA5, 32 * S)
                entry(d_dig, d_rp, S, n, claimed, d_roots, d_empty)
                got = np.empty((S, 32), np.uint8)
                ctx.d2h(got, d_roots)
                with pytest.raises(NativeError, match="max_session_leaves was smaller"):
                    ctx.sync()
                ctx.sync()                                            # reported once
                bad = [s for s in range(S) if not np.array_equal(got[s], want[s])]
                assert bad and all(not got[s].any() for s in bad)     # zero roots, never garbage
                assert all(sizes[s] > claimed for s in bad)           # only sessions larger than what was claimed
                entry(d_dig, d_rp, S, n, max(sizes), d_roots, d_empty)     # the true maximum: everything right
                ctx.d2h(got, d_roots); ctx.sync()
                assert np.array_equal(got, want)
        finally:
            for p in (d_dig, d_rp, d_roots, d_empty):
                ctx.dev_free(p)


def test_host_entries_reject_decreasing_offsets(ctx):
    import ctypes as C
    vp = C.c_void_p
    a = np.zeros(4096, np.uint8)
    out = np.zeros((4, 32), np.uint8)
    offs = np.array([0, 100, 50, 200, 300], np.uint64)                # one decreasing pair inside
    assert ctx.lib.ahv_sha256_batch(ctx.h, vp(a.ctypes.data), vp(offs.ctypes.data), 4, 0, vp(out.ctypes.data)) == -1
    assert b"non-decreasing" in ctx.lib.ahv_last_error()
    rp = np.array([0, 4], np.uint64)
    roots = np.zeros((1, 32), np.uint8)
    assert ctx.lib.ahv_commit(ctx.h, vp(a.ctypes.data), vp(offs.ctypes.data), 0, vp(rp.ctypes.data), 1, 0, None,
                              vp(roots.ctypes.data), None) == -1
    # the pipelined audit call checks its offsets chunk by chunk: a bad pair in a late chunk, and offsets that run
    # past the arena's end before coming back, are both refused (and a good batch of the same shape passes)
    S, L, FL = 12000, 1, 1024                                          # one-delta sessions: no parent, every chain verifies
    n = S * L
    big = ctx.host_alloc(n * FL); big[:] = 32
    good = (np.arange(n + 1, dtype=np.uint64) * FL)
    rp2 = (np.arange(S + 1, dtype=np.uint64) * L)
    pos = np.full(n, 0xFFFFFFFF, np.uint32)
    o = dict(roots=np.zeros(32 * S, np.uint8), verdict=np.zeros(S, np.uint8))

    def call(offsets):
        return ctx.lib.ahv_audit_sessions(ctx.h, vp(big.ctypes.data), vp(offsets.ctypes.data), 0, None, None, vp(pos.ctypes.data),
                                          vp(rp2.ctypes.data), S, 0, None, None, vp(o["roots"].ctypes.data), None,
                                          vp(o["verdict"].ctypes.data), None, None)
    assert call(good) == 0 and o["verdict"].all()
    bad = good.copy(); bad[n - 100] = bad[n - 101] - 1
    assert call(bad) == -1 and b"non-decreasing" in ctx.lib.ahv_last_error()
    bad = good.copy(); bad[50:60] += np.uint64(1 << 40)
    assert call(bad) == -1 and b"non-decreasing" in ctx.lib.ahv_last_error()
    assert call(good) == 0 and o["verdict"].all()                      # the context is still usable
    ctx.host_free(big)


def test_reads_stop_at_the_end_of_the_allocation(ctx):
    """compute-sanitizer is not available on this pool, so out-of-bounds safety of the unaligned tail reads is
    checked the blunt way: inputs are placed so that their last byte is the LAST byte of a device allocation
    (a read past it would leave the allocation; with a 2 MiB-granular allocator that is the next, unmapped page)."""
    rng = np.random.default_rng(12)
    BLOCK = 2 << 20
    for trial in range(6):
        lens = np.concatenate([rng.integers(0, 300, 400), [1, 2, 3, 55, 56, 63, 64, 65, 119, 120, 121 + trial]])
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        total = int(offs[-1])
        arena = rng.integers(0, 256, total, dtype=np.uint8)
        blk = ctx.dev_alloc(BLOCK)
        d_offs, d_out = ctx.dev_alloc(offs.nbytes), ctx.dev_alloc(32 * len(lens))
        try:
            base = blk + BLOCK - total                                # the arena ends exactly where the block ends
            ctx.lib.ahv_memcpy_h2d(ctx.h, base, arena.ctypes.data, total)
            ctx.h2d(d_offs, offs)
            for sort in (0, 1):
                ctx.set_option("sort", sort)
                ctx.sha256_batch_dev(base, total, d_offs, len(lens), 0, d_out)
                got = np.empty((len(lens), 32), np.uint8)
                ctx.d2h(got, d_out); ctx.sync()
                assert np.array_equal(got, co.sha256_batch(arena, offs))
        finally:
            ctx.set_option("sort", 1)
            for p in (blk, d_offs, d_out):
                ctx.dev_free(p)
    # the sigma_eff pass (16-byte bulk copies + ordinary loads for the last partial group): edge arrays that end with
    # their allocations
    A = 5000
    deg = rng.integers(0, 24, A); deg[-1] = 13; deg[100] = 700
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint32)
    E = int(rowptr[-1])
    E += (-E) % 2                                                      # keep the 8-byte arrays 16-byte aligned at their start
    rowptr[-1] = E
    g = dict(bonded=rng.random(E), active=(rng.random(E) < 0.9).astype(np.uint8), expiry=np.full(E, lo.EXPIRY_NONE, np.int64))
    blocks = {k: ctx.dev_alloc(BLOCK) for k in g}
    starts = {k: blocks[k] + BLOCK - ((g[k].nbytes + 15) // 16) * 16 for k in g}        # 16-byte aligned starts, ...
    starts["bonded"] = blocks["bonded"] + BLOCK - g["bonded"].nbytes                      # ... these two end exactly with the block
    starts["expiry"] = blocks["expiry"] + BLOCK - g["expiry"].nbytes
    sigma, omega = rng.random(A), rng.choice([0.2, 0.65, 0.95], A)
    cons, ring = np.zeros(A, np.uint8), np.full(A, 3, np.uint8)
    small = {k: ctx.dev_alloc(max(v.nbytes, 16)) for k, v in dict(rowptr=rowptr, sigma=sigma, omega=omega, cons=cons, ring=ring).items()}
    outs = [ctx.dev_alloc(8 * A), ctx.dev_alloc(A), ctx.dev_alloc(A), ctx.dev_alloc(A)]
    try:
        for k in g:
            ctx.lib.ahv_memcpy_h2d(ctx.h, starts[k], g[k].ctypes.data, g[k].nbytes)
        for k, v in dict(rowptr=rowptr, sigma=sigma, omega=omega, cons=cons, ring=ring).items():
            ctx.h2d(small[k], v)
        ctx.sigma_eff_rings_dev(small["rowptr"], starts["bonded"], starts["active"], starts["expiry"], 1000, small["sigma"],
                                small["omega"], small["cons"], small["ring"], A, *outs)
        got = np.empty(A, np.float64)
        ctx.d2h(got, outs[0]); ctx.sync()
        want = co.sigma_eff_rings(rowptr, g["bonded"], g["active"], g["expiry"], 1000, sigma, omega, cons, ring)
        exp = sigma_expected_default_path(want, rowptr, g["bonded"], g["active"], g["expiry"], 1000, sigma, omega, cons, ring)
        assert np.array_equal(got, exp[0])
    finally:
        for p in list(blocks.values()) + list(small.values()) + outs:
            ctx.dev_free(p)
