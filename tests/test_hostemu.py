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
