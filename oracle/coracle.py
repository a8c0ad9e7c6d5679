"""ctypes wrapper over ``oracle/_build/libahv_oracle.so`` (TEST INFRASTRUCTURE).

The C source is ``oracle/c/ahv_oracle.c`` (plain C restatement, each function
citing the reference lines it follows).  ``build()`` runs ``make -C oracle``.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libahv_oracle.so")
_lib = None

_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)


def build() -> str:
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_sha256.argtypes = [_u8p, C.c_uint64, _u8p]
        L.orc_sha256_batch.argtypes = [_u8p, _u64p, C.c_uint64, _u8p]
        L.orc_count_blocks.argtypes = [_u64p, C.c_uint64]
        L.orc_count_blocks.restype = C.c_uint64
        L.orc_merkle_roots.argtypes = [_u8p, _u64p, C.c_uint64, _u8p, _u8p]
        L.orc_merkle_roots.restype = C.c_int
        L.orc_verify_chains.argtypes = [_u8p, _u64p, _u8p, _u8p, _u64p, C.c_uint64, _u8p, _i64p, _u8p]
        L.orc_sigma_eff_rings.argtypes = [
            _u32p, _f64p, _u8p, _i64p, C.c_int64, _f64p, _f64p, _u8p, _u8p, C.c_uint64,
            _f64p, _u8p, _u8p, _u8p,
        ]
        L.orc_exposure_sums.argtypes = [_u32p, _f64p, _u8p, _i64p, C.c_int64, C.c_uint64, C.c_int, _f64p]
        L.orc_ring_check.argtypes = [_u8p, _u8p, _f64p, _u8p, _u8p, C.c_uint64, _u8p, _u8p, _u8p]
        L.orc_slash_cascade.argtypes = [_u32p, _u32p, _u8p, _i64p, C.c_int64, _u64p, _u64p, _u32p, _f64p, _f64p, _u8p, C.c_uint64,
                                        _f64p, _u8p, C.c_void_p, C.c_void_p, _u32p, _u32p]
        L.orc_num_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_threads(n: int) -> None:
    lib().orc_set_threads(int(n))


def sha256(msg: bytes) -> bytes:
    m = np.frombuffer(msg, dtype=np.uint8) if len(msg) else np.zeros(1, dtype=np.uint8)
    m = _c(m, np.uint8)
    out = np.empty(32, dtype=np.uint8)
    lib().orc_sha256(_p(m, _u8p), len(msg), _p(out, _u8p))
    return out.tobytes()


def sha256_batch(arena, offsets) -> np.ndarray:
    arena = _c(arena, np.uint8)
    if arena.size == 0:
        arena = np.zeros(1, dtype=np.uint8)
    offsets = _c(offsets, np.uint64)
    n = len(offsets) - 1
    out = np.empty((n, 32), dtype=np.uint8)
    lib().orc_sha256_batch(_p(arena, _u8p), _p(offsets, _u64p), n, _p(out, _u8p))
    return out


def count_blocks(offsets) -> int:
    offsets = _c(offsets, np.uint64)
    return int(lib().orc_count_blocks(_p(offsets, _u64p), len(offsets) - 1))


def merkle_roots(digests, rowptr):
    digests = _c(digests, np.uint8).reshape(-1)
    if digests.size == 0:
        digests = np.zeros(32, dtype=np.uint8)
    rowptr = _c(rowptr, np.uint64)
    S = len(rowptr) - 1
    roots = np.zeros((S, 32), dtype=np.uint8)
    empty = np.zeros(S, dtype=np.uint8)
    rc = lib().orc_merkle_roots(_p(digests, _u8p), _p(rowptr, _u64p), S, _p(roots, _u8p), _p(empty, _u8p))
    if rc:
        raise MemoryError("orc_merkle_roots")
    return roots, empty


def verify_chains(arena, offsets, parents, parent_kind, rowptr, want_digests=True):
    arena = _c(arena, np.uint8)
    if arena.size == 0:
        arena = np.zeros(1, dtype=np.uint8)
    offsets = _c(offsets, np.uint64)
    n = len(offsets) - 1
    parents = _c(parents, np.uint8).reshape(-1)
    if parents.size == 0:
        parents = np.zeros(32, dtype=np.uint8)
    kind = _c(parent_kind, np.uint8)
    if kind.size == 0:
        kind = np.zeros(1, dtype=np.uint8)
    rowptr = _c(rowptr, np.uint64)
    S = len(rowptr) - 1
    verdict = np.zeros(S, dtype=np.uint8)
    first_bad = np.zeros(S, dtype=np.int64)
    digs = np.zeros((max(n, 1), 32), dtype=np.uint8) if want_digests else None
    lib().orc_verify_chains(
        _p(arena, _u8p), _p(offsets, _u64p), _p(parents, _u8p), _p(kind, _u8p), _p(rowptr, _u64p), S,
        _p(verdict, _u8p), _p(first_bad, _i64p), _p(digs, _u8p) if want_digests else None,
    )
    return verdict, first_bad, (digs[:n] if want_digests else None)


def sigma_eff_rings(rowptr, bonded, active, expiry, now, sigma, omega, consensus, current_ring):
    rowptr = _c(rowptr, np.uint32)
    A = len(rowptr) - 1
    bonded = _c(bonded, np.float64)
    active = _c(active, np.uint8)
    expiry = _c(expiry, np.int64)
    if bonded.size == 0:
        bonded = np.zeros(1); active = np.zeros(1, np.uint8); expiry = np.zeros(1, np.int64)
    sigma = _c(sigma, np.float64); omega = _c(omega, np.float64)
    consensus = _c(consensus, np.uint8); current_ring = _c(current_ring, np.uint8)
    o_s = np.empty(A, np.float64); o_r = np.empty(A, np.uint8)
    o_d = np.empty(A, np.uint8); o_n = np.empty(A, np.uint8)
    lib().orc_sigma_eff_rings(
        _p(rowptr, _u32p), _p(bonded, _f64p), _p(active, _u8p), _p(expiry, _i64p), int(now),
        _p(sigma, _f64p), _p(omega, _f64p), _p(consensus, _u8p), _p(current_ring, _u8p), A,
        _p(o_s, _f64p), _p(o_r, _u8p), _p(o_d, _u8p), _p(o_n, _u8p),
    )
    return o_s, o_r, o_d, o_n


def ring_check(agent_ring, action_flags, sigma, consensus, witness):
    agent_ring = _c(agent_ring, np.uint8); action_flags = _c(action_flags, np.uint8)
    sigma = _c(sigma, np.float64); consensus = _c(consensus, np.uint8); witness = _c(witness, np.uint8)
    n = len(agent_ring)
    al = np.empty(n, np.uint8); re = np.empty(n, np.uint8); rq = np.empty(n, np.uint8)
    lib().orc_ring_check(_p(agent_ring, _u8p), _p(action_flags, _u8p), _p(sigma, _f64p),
                         _p(consensus, _u8p), _p(witness, _u8p), n, _p(al, _u8p), _p(re, _u8p), _p(rq, _u8p))
    return al, re, rq


def exposure_sums(rowptr, bonded, active, expiry, now, compensated):
    rowptr = _c(rowptr, np.uint32)
    R = len(rowptr) - 1
    bonded = _c(bonded, np.float64); active = _c(active, np.uint8); expiry = _c(expiry, np.int64)
    if bonded.size == 0:
        bonded = np.zeros(1); active = np.zeros(1, np.uint8); expiry = np.zeros(1, np.int64)
    out = np.zeros(R, np.float64)
    lib().orc_exposure_sums(_p(rowptr, _u32p), _p(bonded, _f64p), _p(active, _u8p), _p(expiry, _i64p), int(now), R,
                            1 if compensated else 0, _p(out, _f64p))
    return out


def slash_cascade(rowptr, voucher, active, expiry, now, session_agent_rowptr, event_rowptr, event_agent, event_sigma,
                  event_omega, scores, event_depth=None):
    """Same contract as liability_oracle.slash_cascade (copies in, copies out)."""
    from .liability_oracle import CLIP_DT, SLASH_DT
    rowptr = _c(rowptr, np.uint32); voucher = _c(voucher, np.uint32)
    active = np.array(active, dtype=np.uint8, copy=True); expiry = _c(expiry, np.int64)
    srp = _c(session_agent_rowptr, np.uint64); erp = _c(event_rowptr, np.uint64)
    ea = _c(event_agent, np.uint32); es = _c(event_sigma, np.float64); eo = _c(event_omega, np.float64)
    scores = np.array(scores, dtype=np.float64, copy=True)
    S = len(srp) - 1
    E = int(rowptr[-1])
    touched = np.zeros(max(len(scores), 1), np.uint8)
    sl = np.zeros(len(ea) + E + 1, dtype=SLASH_DT); cl = np.zeros(E + 1, dtype=CLIP_DT)
    ns = np.zeros(max(S, 1), np.uint32); nc = np.zeros(max(S, 1), np.uint32)
    if voucher.size == 0:
        voucher = np.zeros(1, np.uint32); active = np.zeros(1, np.uint8); expiry = np.zeros(1, np.int64)
    if ea.size == 0:
        ea = np.zeros(1, np.uint32); es = np.zeros(1); eo = np.zeros(1)
    sc = scores if scores.size else np.zeros(1)
    lib().orc_slash_cascade(_p(rowptr, _u32p), _p(voucher, _u32p), _p(active, _u8p), _p(expiry, _i64p), int(now),
                            _p(srp, _u64p), _p(erp, _u64p), _p(ea, _u32p), _p(es, _f64p), _p(eo, _f64p),
                            None if event_depth is None else _p(_c(event_depth, np.uint8), _u8p), S,
                            _p(sc, _f64p), _p(touched, _u8p), sl.ctypes.data, cl.ctypes.data, _p(ns, _u32p), _p(nc, _u32p))
    return scores, touched[: len(scores)], active[:E], sl[: len(event_agent) + E], cl[:E], ns[:S], nc[:S]
