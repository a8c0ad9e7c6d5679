"""CPU-only: compiles the per-thread device functions (sha256_core.cuh, liability_core.cuh) as
plain C++ and checks them against the golden vectors, so arithmetic errors are caught before a
GPU run.  The emulation library is test infrastructure; the product never loads it."""

import ctypes as C
import hashlib
import os
import subprocess

import numpy as np
import pytest

import recipes
from oracle import liability_oracle as lo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "hostemu", "hostemu.cpp")
OUT = os.path.join(ROOT, "tests", "_build", "libahv_hostemu.so")
u8p = C.POINTER(C.c_uint8)


@pytest.fixture(scope="module")
def emu():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-x", "c++", SRC, "-o", OUT],
                   check=True)
    L = C.CDLL(OUT)
    L.emu_sigma_row.restype = C.c_double
    L.emu_sigma_row.argtypes = [C.POINTER(C.c_double), u8p, C.POINTER(C.c_int64), C.c_int, C.c_int64,
                                C.c_double, C.c_double]
    L.emu_ring.argtypes = [C.c_double, C.c_int]
    L.emu_gate.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int)]
    L.emu_near.argtypes = [C.c_double]
    return L


def _sha(L, buf, off, ln):
    out = np.zeros(32, np.uint8)
    L.emu_sha256(C.cast(buf.ctypes.data + off, u8p), ln, out.ctypes.data_as(u8p))
    return out.tobytes()


def test_device_sha256_all_alignments(emu, audit_golden):
    for v in audit_golden["sha_kats"]:
        m = recipes.kat_message(v["seed"], v["len"])
        for a in (0, 1, 2, 3, 4, 7, 8, 13, 16):
            buf = np.zeros(len(m) + 64, np.uint8)
            buf[a:a + len(m)] = np.frombuffer(m, np.uint8)
            assert _sha(emu, buf, a, len(m)).hex() == v["digest"], (v["len"], a)
    rng = np.random.default_rng(3)
    buf = rng.integers(0, 256, 5000, dtype=np.uint8)
    for ln in range(0, 200):
        for a in range(0, 17):
            assert _sha(emu, buf, a, ln) == hashlib.sha256(buf[a:a + ln].tobytes()).digest()


def test_device_merkle_node(emu, audit_golden):
    nk = audit_golden["node_kat"]
    out = np.zeros(32, np.uint8)
    for l, r in [(hashlib.sha256(b"x").digest(), hashlib.sha256(b"y").digest()),
                 (bytes(range(32)), bytes(range(255, 223, -1))), (b"\x00" * 32, b"\xff" * 32),
                 (bytes.fromhex("0123456789abcdef" * 4), bytes.fromhex("fedcba9876543210" * 4))]:
        la, ra = np.frombuffer(l, np.uint8).copy(), np.frombuffer(r, np.uint8).copy()
        emu.emu_node(la.ctypes.data_as(u8p), ra.ctypes.data_as(u8p), out.ctypes.data_as(u8p))
        assert out.tobytes() == hashlib.sha256((l.hex() + r.hex()).encode()).digest()
    # "a"*64 + "b"*64 is hex of 0xaa.. and 0xbb..
    la = np.full(32, 0xAA, np.uint8); ra = np.full(32, 0xBB, np.uint8)
    emu.emu_node(la.ctypes.data_as(u8p), ra.ctypes.data_as(u8p), out.ctypes.data_as(u8p))
    assert out.tobytes().hex() == nk["digest"]


def test_device_liability_arithmetic(emu, liability_golden):
    for r in liability_golden["ring_grid"]:
        s = float("nan") if r["sigma"] == "nan" else float.fromhex(r["sigma"])
        assert emu.emu_ring(s, int(r["consensus"])) == r["ring"]
    req = C.c_int()
    for c in liability_golden["checks"]:
        flags = (1 if c[0] else 0) | (2 if c[1] else 0) | (4 if c[2] == "none" else 0)
        s = float("nan") if c[4] == "nan" else float.fromhex(c[4])
        code = emu.emu_gate(c[3], flags, s, c[5], c[6], C.byref(req))
        assert (code == 0) == bool(c[7]) and req.value == c[8]
        assert (code == 3) == bool(c[10]) and (code == 1) == bool(c[11])
    # sequential row sums are bit-identical to the reference's loop
    rng = np.random.default_rng(5)
    for _ in range(200):
        n = int(rng.integers(0, 40))
        bonded = rng.uniform(0, 0.8, n); active = (rng.random(n) < 0.9).astype(np.uint8)
        expiry = np.where(rng.random(n) < 0.7, lo.EXPIRY_NONE, rng.integers(0, 2000, n)).astype(np.int64)
        sigma, omega = float(rng.random()), float(rng.choice([0.2, 0.65, 0.95]))
        want, *_ = lo.sigma_eff_rings(np.array([0, n], np.uint32), bonded, active, expiry, 1000,
                                      [sigma], [omega], [0], [3])
        got = emu.emu_sigma_row(bonded.ctypes.data_as(C.POINTER(C.c_double)), active.ctypes.data_as(u8p),
                                expiry.ctypes.data_as(C.POINTER(C.c_int64)), n, 1000, sigma, omega)
        assert got == want[0]
    assert emu.emu_near(0.6) and emu.emu_near(0.95) and emu.emu_near(0.6000003) and not emu.emu_near(0.61)


def test_compress_variants_agree(emu):
    """The FMA-pipe formulations (adds as IMAD, sigmas as IMAD.WIDE/HI) are the same function."""
    rng = np.random.default_rng(9)
    for _ in range(200):
        blk = rng.integers(0, 2**32, 16, dtype=np.uint32)
        out = np.zeros(40, np.uint32)
        emu.emu_compress_variants(blk.ctypes.data_as(C.POINTER(C.c_uint32)), out.ctypes.data_as(C.POINTER(C.c_uint32)))
        for v in (1, 2, 3, 4):    # IMAD adds, WIDE/HI sigmas, scaled-sum, scaled-sum with K table
            assert np.array_equal(out[:8], out[8 * v:8 * v + 8]), v


def test_scaled_policy_matches_hashlib(emu):
    """The scaled-sum compression policy (what the kernels run by default) end to end."""
    rng = np.random.default_rng(17)
    buf = rng.integers(0, 256, 40000, dtype=np.uint8)
    out = np.zeros(32, np.uint8)
    for ln in list(range(0, 130)) + [255, 256, 257, 1000, 4096, 16384]:
        for a in (0, 1, 3, 16):
            emu.emu_sha256_scaled(C.cast(buf.ctypes.data + a, u8p), ln, out.ctypes.data_as(u8p))
            assert out.tobytes() == hashlib.sha256(buf[a:a + ln].tobytes()).digest(), (ln, a)
    for _ in range(20):
        l = rng.integers(0, 256, 32, dtype=np.uint8); r = rng.integers(0, 256, 32, dtype=np.uint8)
        emu.emu_node_scaled(l.ctypes.data_as(u8p), r.ctypes.data_as(u8p), out.ctypes.data_as(u8p))
        assert out.tobytes() == hashlib.sha256((l.tobytes().hex() + r.tobytes().hex()).encode()).digest()
        m = rng.integers(0, 256, 256, dtype=np.uint8)
        for scaled in (0, 1):
            emu.emu_sha256_fixed256(m.ctypes.data_as(u8p), scaled, out.ctypes.data_as(u8p))
            assert out.tobytes() == hashlib.sha256(m.tobytes()).digest()


def test_search_formulation_is_the_same_function(emu):
    rng = np.random.default_rng(23)
    for _ in range(100):
        blk = rng.integers(0, 2**32, 16, dtype=np.uint32)
        out = np.zeros(16, np.uint32)
        emu.emu_compress_search(blk.ctypes.data_as(C.POINTER(C.c_uint32)), out.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert np.array_equal(out[:8], out[8:])


def test_slash_lane_walker_matches_reference_records(emu, slashing_golden):
    """The code k_slash_lanes runs per lane (slashing_core.cuh: 16-bit staged runs, serial cascade with the
    clip cache, write-back), by one host thread, against the real SlashingEngine's recorded history; sessions
    too large for a slice are the ones the warp kernel would take."""
    from helpers import check_slash_against_golden, slash_batch
    from agent_hypervisor_b200 import _native as nat
    gold = slashing_golden["cases"]
    cases = [recipes.random_slash_case(g["seed"]) for g in gold]
    b = slash_batch(cases)
    A, E, S, NE = len(b["scores"]), len(b["voucher"]), len(cases), len(b["event_agent"])
    o = dict(scores=np.array(b["scores"], copy=True), touched=np.zeros(A, np.uint8), active=np.array(b["active"], copy=True),
             slash_log=np.zeros(NE + E, nat.SLASH_DT), clip_log=np.zeros(E, nat.CLIP_DT), slash_count=np.zeros(S, np.uint32),
             clip_count=np.zeros(S, np.uint32), status=np.zeros(S, np.uint8))
    sb = nat.SlashBatch()
    sb.n_agents, sb.now, sb.n_sessions = A, int(b["now"]), S
    keep = {k: np.ascontiguousarray(b[k]) for k in ("rowptr", "voucher", "expiry", "session_agent_rowptr", "event_rowptr",
                                                    "event_agent", "event_sigma", "event_omega", "event_depth")}
    for name, arr in list(keep.items()) + list(o.items()):
        setattr(sb, name, arr.ctypes.data)
    big = np.zeros(S, np.uint32)
    emu.emu_slash_lanes.restype = C.c_uint64
    n_big = emu.emu_slash_lanes(C.byref(sb), big.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert n_big == 0 and not o["status"].any()            # every golden session is small enough for a lane
    check_slash_against_golden(cases, gold, b, (o["scores"], o["touched"], o["active"], o["slash_log"], o["clip_log"],
                                                o["slash_count"], o["clip_count"]))
    # long rows overflow the 12-entry clip cache and read the clip records back; big sessions are listed
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import coracle as co
    rng = np.random.default_rng(3)
    A2 = [3, 40, 2, 700, 5]; E2 = [60, 300, 0, 3000, 200]                    # the 700 x 3000 session cannot fit
    srp = np.concatenate([[0], np.cumsum(A2)]).astype(np.uint64)
    vouchee = np.concatenate([np.sort(rng.integers(0, a, e)) + int(srp[i]) for i, (a, e) in enumerate(zip(A2, E2))])
    voucher = np.concatenate([rng.integers(0, a, e) + int(srp[i]) for i, (a, e) in enumerate(zip(A2, E2))]).astype(np.uint32)
    rowptr = np.zeros(int(srp[-1]) + 1, np.uint32); np.add.at(rowptr, vouchee + 1, 1); rowptr = np.cumsum(rowptr).astype(np.uint32)
    Et = int(rowptr[-1])
    b2 = dict(rowptr=rowptr, voucher=voucher, active=np.ones(Et, np.uint8), expiry=np.full(Et, np.iinfo(np.int64).max),
              now=0, session_agent_rowptr=srp, event_rowptr=np.arange(6, dtype=np.uint64), event_agent=srp[:-1].astype(np.uint32),
              event_sigma=np.full(5, 0.5), event_omega=np.full(5, 0.95), scores=rng.random(int(srp[-1])))
    want = co.slash_cascade(*[b2[k] for k in ("rowptr", "voucher", "active", "expiry", "now", "session_agent_rowptr",
                                              "event_rowptr", "event_agent", "event_sigma", "event_omega", "scores")])
    o2 = dict(scores=np.array(b2["scores"], copy=True), touched=np.zeros(len(b2["scores"]), np.uint8), active=np.array(b2["active"], copy=True),
              slash_log=np.zeros(5 + Et, nat.SLASH_DT), clip_log=np.zeros(Et, nat.CLIP_DT), slash_count=np.zeros(5, np.uint32),
              clip_count=np.zeros(5, np.uint32), status=np.zeros(5, np.uint8))
    sb2 = nat.SlashBatch()
    sb2.n_agents, sb2.now, sb2.n_sessions = len(b2["scores"]), 0, 5
    keep2 = {k: np.ascontiguousarray(b2[k]) for k in ("rowptr", "voucher", "expiry", "session_agent_rowptr", "event_rowptr",
                                                      "event_agent", "event_sigma", "event_omega")}
    for name, arr in list(keep2.items()) + list(o2.items()):
        setattr(sb2, name, arr.ctypes.data)
    sb2.event_depth = None
    n_big = emu.emu_slash_lanes(C.byref(sb2), big.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert n_big == 1 and big[0] == 3
    for s in (0, 1, 2, 4):
        assert o2["slash_count"][s] == want[5][s] and o2["clip_count"][s] == want[6][s]
        a0, a1 = int(srp[s]), int(srp[s + 1])
        assert np.array_equal(o2["scores"][a0:a1], want[0][a0:a1]) and np.array_equal(o2["touched"][a0:a1], want[1][a0:a1])
        e0, e1 = int(rowptr[a0]), int(rowptr[a1])
        assert np.array_equal(o2["active"][e0:e1], want[2][e0:e1])
        assert o2["clip_log"][e0:e0 + int(want[6][s])].tobytes() == want[4][e0:e0 + int(want[6][s])].tobytes()
    assert int(want[6][0]) > 12                                              # the cache did overflow


def test_rolled_node_bodies_match_hashlib(emu):
    """The code-size-reduced compression bodies of the tree kernels (rounds rolled, constants by index) give
    the reference's node: sha256((left.hex() + right.hex()).encode()) (audit/delta.py:130)."""
    rng = np.random.default_rng(21)
    for _ in range(300):
        l = rng.integers(0, 256, 32, dtype=np.uint8); r = rng.integers(0, 256, 32, dtype=np.uint8)
        out = np.zeros(32, np.uint8)
        for variant in (16, 32, 48):              # straight-line rounds before the rolled part
            emu.emu_node_rolled(l.ctypes.data_as(u8p), r.ctypes.data_as(u8p), out.ctypes.data_as(u8p), variant)
            assert out.tobytes() == hashlib.sha256((l.tobytes().hex() + r.tobytes().hex()).encode()).digest()
