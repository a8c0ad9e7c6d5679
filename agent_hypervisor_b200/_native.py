"""ctypes binding of ``libahv_b200.so`` (C ABI declared in ``include/ahv.h``).

There is NO CPU fallback: if the shared library is missing, or no CUDA device is
present, every entry point raises ``NativeUnavailable`` - loudly, on first use.
"""

from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_PKG, "libahv_b200.so")

PARENT_NONE, PARENT_HEX, PARENT_OTHER = 0, 1, 2
ACT_ADMIN, ACT_READ_ONLY, ACT_REV_NONE = 1, 2, 4
REASON_GRANTED, REASON_NEEDS_WITNESS, REASON_RING1_SIGMA = 0, 1, 2
REASON_NEEDS_CONSENSUS, REASON_RING2_SIGMA, REASON_RING_INSUFFICIENT = 3, 4, 5
EXPIRY_NONE = np.iinfo(np.int64).max


# ahv_slash_record / ahv_clip_record (include/ahv.h)
SLASH_DT = np.dtype([("agent", "<u4"), ("parent", "<i4"), ("event", "<u4"), ("depth", "<u4"),
                     ("clip_begin", "<u4"), ("clip_count", "<u4"), ("sigma_before", "<f8")])
CLIP_DT = np.dtype([("edge", "<u4"), ("slash", "<u4"), ("sigma_before", "<f8"), ("sigma_after", "<f8")])
SLASH_BAD_EVENT, SLASH_BAD_VOUCHER = 1, 2


class SlashBatch(C.Structure):
    """ahv_slash_batch (include/ahv.h)."""
    _fields_ = [("n_agents", C.c_uint64), ("rowptr", C.c_void_p), ("voucher", C.c_void_p), ("active", C.c_void_p),
                ("expiry", C.c_void_p), ("now", C.c_int64), ("n_sessions", C.c_uint64),
                ("session_agent_rowptr", C.c_void_p), ("event_rowptr", C.c_void_p), ("event_agent", C.c_void_p),
                ("event_sigma", C.c_void_p), ("event_omega", C.c_void_p), ("event_depth", C.c_void_p),
                ("scores", C.c_void_p),
                ("touched", C.c_void_p), ("slash_log", C.c_void_p), ("clip_log", C.c_void_p),
                ("slash_count", C.c_void_p), ("clip_count", C.c_void_p), ("status", C.c_void_p)]


class NativeUnavailable(RuntimeError):
    """libahv_b200.so could not be loaded or no CUDA device is usable."""


class NativeError(RuntimeError):
    """An ahv_* call returned an error code."""


_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)
_vp = C.c_void_p
_u64 = C.c_uint64

# name -> (restype, argtypes); must list every symbol declared in include/ahv.h
SIGNATURES = {
    "ahv_abi_version": (C.c_int, []),
    "ahv_last_error": (C.c_char_p, []),
    "ahv_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "ahv_destroy": (None, [_vp]),
    "ahv_sync": (C.c_int, [_vp]),
    "ahv_device_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), _u64p]),
    "ahv_stream": (_vp, [_vp]),
    "ahv_launch_count": (_u64, [_vp]),
    "ahv_dev_alloc": (C.c_int, [_vp, _u64, C.POINTER(_vp)]),
    "ahv_dev_free": (C.c_int, [_vp, _vp]),
    "ahv_host_alloc": (C.c_int, [_vp, _u64, C.POINTER(_vp)]),
    "ahv_host_alloc_flags": (C.c_int, [_vp, _u64, C.c_uint32, C.POINTER(_vp)]),
    "ahv_host_free": (C.c_int, [_vp, _vp]),
    "ahv_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, _u64]),
    "ahv_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, _u64]),
    "ahv_memset_dev": (C.c_int, [_vp, _vp, C.c_int, _u64]),
    "ahv_upload_async": (C.c_int, [_vp, _vp, _vp, _u64]),
    "ahv_upload_fence": (C.c_int, [_vp]),
    "ahv_timer_start": (C.c_int, [_vp, C.c_int]),
    "ahv_timer_stop": (C.c_int, [_vp, C.c_int]),
    "ahv_timer_elapsed_ms": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float)]),
    "ahv_timer_wait": (C.c_int, [_vp, C.c_int]),
    "ahv_sha256_batch": (C.c_int, [_vp, _vp, _vp, _u64, _u64, _vp]),
    "ahv_sha256_one": (C.c_int, [_vp, C.c_char_p, _u64, _vp]),
    "ahv_sha256_batch_dev": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _u64, _vp]),
    "ahv_merkle_roots": (C.c_int, [_vp, _vp, _vp, _u64, _vp, _vp]),
    "ahv_merkle_roots_dev": (C.c_int, [_vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp]),
    "ahv_merkle_roots_noalloc_dev": (C.c_int, [_vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp]),
    "ahv_commit": (C.c_int, [_vp, _vp, _vp, _u64, _vp, _u64, _u64, _vp, _vp, _vp]),
    "ahv_commit_dev": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _vp, _u64, _u64, _u64, _u64, _vp, _vp, _vp]),
    "ahv_verify_chains": (C.c_int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ahv_verify_chains_dev": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _vp, _vp, _vp, _u64, _u64, _u64, _u64, _vp,
                                        _vp, _vp, _vp, _vp, _vp]),
    "ahv_audit_sessions": (C.c_int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp,
                                     _vp]),
    "ahv_slash_cascade": (C.c_int, [_vp, _vp]),
    "ahv_slash_cascade_dev": (C.c_int, [_vp, _vp]),
    "ahv_extract_parents_dev": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _vp, _u64, _vp, _vp]),
    "ahv_chain_capture": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp]),
    "ahv_chain_capture_dev": (C.c_int, [_vp, _vp, _u64, _vp, _vp, _vp, _u64, _u64, _vp, _vp, _vp]),
    "ahv_pack_deltas": (C.c_int, [_vp, _vp, _vp, _u64, _u64p, _vp, C.c_uint32]),
    "ahv_pack_deltas_dev": (C.c_int, [_vp, _vp, _vp, _vp, _u64, _u64p, _vp, C.c_uint32]),
    "ahv_sigma_eff_rings": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, _u64,
                                      _vp, _vp, _vp, _vp]),
    "ahv_sigma_eff_rings_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, _u64,
                                          _vp, _vp, _vp, _vp]),
    "ahv_exposure_sums": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int64, _u64, C.c_int, _vp]),
    "ahv_exposure_sums_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int64, _u64, C.c_int, _vp]),
    "ahv_ring_check": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp]),
    "ahv_ring_check_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp]),
    "ahv_comm_unique_id": (C.c_int, [_vp]),
    "ahv_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "ahv_comm_destroy": (C.c_int, [_vp]),
    "ahv_comm_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ahv_comm_all_gather_dev": (C.c_int, [_vp, _vp, _vp, _u64]),
    "ahv_comm_all_gather": (C.c_int, [_vp, _vp, _vp, _u64]),
    "ahv_gather_stride": (_u64, [_u64]),
    "ahv_gather_results_dev": (C.c_int, [_vp, _vp, _vp, _u64, _vp]),
    "ahv_int_peak": (C.c_int, [_vp, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]),
    "ahv_set_option": (C.c_int, [_vp, C.c_char_p, C.c_int64]),
}

_lib = None
_lib_lock = threading.Lock()


def load_library():
    """dlopen the library and bind every symbol (no CUDA call is made here)."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(SO_PATH):
                raise NativeUnavailable(
                    f"{SO_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(nvcc, sm_100a). There is no CPU fallback.")
            try:
                lib = C.CDLL(SO_PATH)
            except OSError as e:  # pragma: no cover
                raise NativeUnavailable(f"cannot load {SO_PATH}: {e}") from e
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)       # AttributeError if a declared symbol is missing
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


def _arr(a, dtype) -> Optional[np.ndarray]:
    if a is None:
        return None
    return np.ascontiguousarray(a, dtype=dtype)


class Context:
    """One per GPU per process (wraps ahv_ctx*)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = _vp()
        rc = self.lib.ahv_create(int(device), C.byref(h))
        if rc != 0:
            msg = self.lib.ahv_last_error().decode()
            raise NativeUnavailable(f"ahv_create({device}) failed ({rc}): {msg}")
        self.h = h
        self.device = int(device)
        self._lock = threading.Lock()   # one host thread per call
        self._options: dict[str, int] = {}  # what set_option was last given (to restore after a per-call override)
        self._pinned: dict[int, int] = {}   # numpy data address -> pinned allocation

    # -- plumbing -----------------------------------------------------------
    def _check(self, rc: int, what: str):
        if rc != 0:
            raise NativeError(f"{what} failed ({rc}): {self.lib.ahv_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.ahv_destroy(self.h)
            self.h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        self._check(self.lib.ahv_sync(self.h), "ahv_sync")

    def device_info(self):
        sm, khz, mem = C.c_int(), C.c_int(), _u64()
        self._check(self.lib.ahv_device_info(self.h, C.byref(sm), C.byref(khz), C.byref(mem)), "ahv_device_info")
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "total_mem": mem.value}

    def stream(self) -> int:
        return int(self.lib.ahv_stream(self.h) or 0)

    def launch_count(self) -> int:
        return int(self.lib.ahv_launch_count(self.h))

    def set_option(self, key: str, value: int):
        self._check(self.lib.ahv_set_option(self.h, key.encode(), int(value)), f"ahv_set_option({key})")
        self._options[key] = int(value)

    def dev_alloc(self, nbytes: int) -> int:
        p = _vp()
        self._check(self.lib.ahv_dev_alloc(self.h, int(nbytes), C.byref(p)), "ahv_dev_alloc")
        return int(p.value or 0)

    def dev_free(self, dptr: int):
        self._check(self.lib.ahv_dev_free(self.h, _vp(dptr)), "ahv_dev_free")

    def host_alloc(self, nbytes: int, dtype=np.uint8, write_combined: bool = False) -> np.ndarray:
        """Pinned host array (freed with host_free(arr)).  write_combined: upload-only buffer."""
        p = _vp()
        self._check(self.lib.ahv_host_alloc_flags(self.h, int(nbytes), 1 if write_combined else 0, C.byref(p)),
                    "ahv_host_alloc")
        buf = (C.c_uint8 * int(nbytes)).from_address(p.value)
        a = np.frombuffer(buf, dtype=np.uint8).view(dtype)
        a.flags.writeable = True
        self._pinned[a.ctypes.data] = p.value
        return a

    def host_free(self, arr: np.ndarray):
        p = self._pinned.pop(arr.ctypes.data, None)
        if p is not None:
            self._check(self.lib.ahv_host_free(self.h, _vp(p)), "ahv_host_free")

    def h2d(self, dptr: int, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        self._check(self.lib.ahv_memcpy_h2d(self.h, _vp(dptr), _vp(arr.ctypes.data), arr.nbytes), "ahv_memcpy_h2d")
        return arr   # keep alive until sync

    def d2h(self, arr: np.ndarray, dptr: int):
        self._check(self.lib.ahv_memcpy_d2h(self.h, _vp(arr.ctypes.data), _vp(dptr), arr.nbytes), "ahv_memcpy_d2h")

    def upload_async(self, dptr: int, arr: np.ndarray, nbytes: Optional[int] = None):
        """Host -> device on the context's second stream (overlaps the main stream's kernels); see upload_fence."""
        self._check(self.lib.ahv_upload_async(self.h, _vp(dptr), _vp(arr.ctypes.data), int(arr.nbytes if nbytes is None else nbytes)),
                    "ahv_upload_async")

    def upload_fence(self):
        self._check(self.lib.ahv_upload_fence(self.h), "ahv_upload_fence")

    def pack_deltas_dev(self, cols_struct, d_offsets: int, d_arena: int, capacity: int, d_parent_pos: int, want_bytes: bool = False):
        """ahv_pack_deltas_dev; ``cols_struct`` is the ctypes struct of audit.packing with DEVICE pointers."""
        nb = _u64()
        self._check(self.lib.ahv_pack_deltas_dev(self.h, C.byref(cols_struct), _vp(d_offsets), _vp(d_arena), int(capacity),
                                                 C.byref(nb) if want_bytes else None, _vp(d_parent_pos), 0), "ahv_pack_deltas_dev")
        return int(nb.value) if want_bytes else None

    def memset(self, dptr: int, value: int, nbytes: int):
        self._check(self.lib.ahv_memset_dev(self.h, _vp(dptr), int(value), int(nbytes)), "ahv_memset_dev")

    def timer_start(self, slot: int):
        self._check(self.lib.ahv_timer_start(self.h, slot), "ahv_timer_start")

    def timer_stop(self, slot: int):
        self._check(self.lib.ahv_timer_stop(self.h, slot), "ahv_timer_stop")

    def timer_ms(self, slot: int) -> float:
        ms = C.c_float()
        self._check(self.lib.ahv_timer_elapsed_ms(self.h, slot, C.byref(ms)), "ahv_timer_elapsed_ms")
        return float(ms.value)

    def int_peak(self, mode: int, iters: int = 2000):
        v, ms = C.c_double(), C.c_float()
        self._check(self.lib.ahv_int_peak(self.h, mode, iters, C.byref(v), C.byref(ms)), "ahv_int_peak")
        return float(v.value), float(ms.value)

    # -- multi-GPU: NCCL communicator owned by the context (csrc/ahv_comm.cu) ---
    def comm_unique_id(self) -> bytes:
        buf = C.create_string_buffer(128)
        self._check(self.lib.ahv_comm_unique_id(C.cast(buf, _vp)), "ahv_comm_unique_id")
        return buf.raw

    def comm_init(self, world: int, rank: int, unique_id: Optional[bytes] = None):
        buf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        self._check(self.lib.ahv_comm_init(self.h, int(world), int(rank), C.cast(buf, _vp) if buf is not None else None),
                    "ahv_comm_init")

    def comm_destroy(self):
        self._check(self.lib.ahv_comm_destroy(self.h), "ahv_comm_destroy")

    def comm_info(self) -> dict:
        w, r, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.ahv_comm_info(self.h, C.byref(w), C.byref(r), C.byref(v)), "ahv_comm_info")
        return {"world": w.value, "rank": r.value, "nccl_version": v.value}

    def comm_all_gather(self, local: np.ndarray) -> np.ndarray:
        """All-gather equal-sized host arrays: returns [world, *local.shape] (stages through the device)."""
        local = np.ascontiguousarray(local)
        world = self.comm_info()["world"]
        out = np.empty((world,) + local.shape, dtype=local.dtype)
        with self._lock:
            self._check(self.lib.ahv_comm_all_gather(self.h, _ptr(local), _ptr(out), local.nbytes), "ahv_comm_all_gather")
        return out

    def comm_all_gather_dev(self, d_send: int, d_recv: int, bytes_per_rank: int):
        self._check(self.lib.ahv_comm_all_gather_dev(self.h, _vp(d_send), _vp(d_recv), int(bytes_per_rank)),
                    "ahv_comm_all_gather_dev")

    def gather_stride(self, n_sessions: int) -> int:
        return int(self.lib.ahv_gather_stride(int(n_sessions)))

    def gather_results_dev(self, d_roots: int, d_verdict: int, n_sessions: int, d_all: int):
        """One all-gather of this rank's roots + verdicts into d_all (world * gather_stride(S) bytes)."""
        self._check(self.lib.ahv_gather_results_dev(self.h, _vp(d_roots), _vp(d_verdict), int(n_sessions), _vp(d_all)),
                    "ahv_gather_results_dev")

    def barrier(self):
        """All ranks of the communicator have reached this point (and this rank's stream has drained)."""
        self.comm_all_gather(np.zeros(1, np.uint64))

    # -- host-buffer entry points (numpy in / numpy out) ---------------------
    def sha256_batch(self, arena, offsets=None, fixed_len: int = 0, n: Optional[int] = None) -> np.ndarray:
        arena = _arr(arena, np.uint8)
        offsets = _arr(offsets, np.uint64)
        if offsets is not None:
            n = len(offsets) - 1
        elif n is None:
            n = arena.size // fixed_len if fixed_len else 0
        out = np.empty((n, 32), dtype=np.uint8)
        with self._lock:
            self._check(self.lib.ahv_sha256_batch(self.h, _ptr(arena), _ptr(offsets), n, int(fixed_len), _ptr(out)),
                        "ahv_sha256_batch")
        return out

    def sha256_one(self, msg: bytes) -> bytes:
        """SHA-256 of one message through the lowest-latency entry (ahv_sha256_one)."""
        out = C.create_string_buffer(32)
        with self._lock:
            self._check(self.lib.ahv_sha256_one(self.h, msg, len(msg), out), "ahv_sha256_one")
        return out.raw

    def merkle_roots(self, digests, session_rowptr):
        digests = _arr(digests, np.uint8)
        rp = _arr(session_rowptr, np.uint64)
        S = len(rp) - 1
        roots = np.zeros((S, 32), dtype=np.uint8)
        empty = np.zeros(S, dtype=np.uint8)
        with self._lock:
            self._check(self.lib.ahv_merkle_roots(self.h, _ptr(digests), _ptr(rp), S, _ptr(roots), _ptr(empty)),
                        "ahv_merkle_roots")
        return roots, empty

    def commit(self, arena, offsets=None, session_rowptr=None, fixed_len: int = 0, uniform_leaves: int = 0,
               n_sessions: Optional[int] = None, want_digests: bool = True):
        arena = _arr(arena, np.uint8)
        offsets = _arr(offsets, np.uint64)
        rp = _arr(session_rowptr, np.uint64)
        S = (len(rp) - 1) if rp is not None else int(n_sessions)
        n = int(rp[-1]) if rp is not None else S * int(uniform_leaves)
        digs = np.empty((n, 32), dtype=np.uint8) if want_digests else None
        roots = np.zeros((S, 32), dtype=np.uint8)
        empty = np.zeros(S, dtype=np.uint8)
        with self._lock:
            self._check(self.lib.ahv_commit(self.h, _ptr(arena), _ptr(offsets), int(fixed_len), _ptr(rp), S,
                                            int(uniform_leaves), _ptr(digs), _ptr(roots), _ptr(empty)), "ahv_commit")
        return digs, roots, empty

    def verify_chains(self, arena, offsets, parent_digests, parent_kind, session_rowptr=None, fixed_len: int = 0,
                      uniform_leaves: int = 0, n_sessions: Optional[int] = None, expected_roots=None,
                      want_digests: bool = True, want_roots: bool = False):
        arena = _arr(arena, np.uint8)
        offsets = _arr(offsets, np.uint64)
        par = _arr(parent_digests, np.uint8)
        kind = _arr(parent_kind, np.uint8)
        rp = _arr(session_rowptr, np.uint64)
        exp = _arr(expected_roots, np.uint8)
        S = (len(rp) - 1) if rp is not None else int(n_sessions)
        n = int(rp[-1]) if rp is not None else S * int(uniform_leaves)
        verdict = np.ones(S, dtype=np.uint8)
        first_bad = np.full(S, -1, dtype=np.int64)
        digs = np.empty((n, 32), dtype=np.uint8) if want_digests else None
        roots = np.zeros((S, 32), dtype=np.uint8) if (want_roots or exp is not None) else None
        match = np.zeros(S, dtype=np.uint8) if exp is not None else None
        with self._lock:
            self._check(self.lib.ahv_verify_chains(
                self.h, _ptr(arena), _ptr(offsets), int(fixed_len), _ptr(par), _ptr(kind), _ptr(rp), S,
                int(uniform_leaves), _ptr(exp), _ptr(verdict), _ptr(first_bad), _ptr(digs), _ptr(roots),
                _ptr(match)), "ahv_verify_chains")
        return {"verdict": verdict, "first_bad": first_bad, "digests": digs, "roots": roots, "root_match": match}

    def audit_sessions(self, arena, offsets, parent_digests, parent_kind, session_rowptr=None, fixed_len: int = 0,
                       uniform_leaves: int = 0, n_sessions: Optional[int] = None, want_digests: bool = True, out=None,
                       parent_pos=None, expected_roots=None):
        """Commit + verify on one upload (ahv_audit_sessions).  ``out`` may supply preallocated
        (e.g. pinned) result arrays: dict with digests/roots/empty/verdict/first_bad/root_match.
        ``expected_roots`` [S,32]: the stored commitments ``root_match`` is checked against; without them
        ``root_match`` only says that the verify pass rebuilt the roots the commit pass of this call produced."""
        arena = _arr(arena, np.uint8)
        offsets = _arr(offsets, np.uint64)
        par = _arr(parent_digests, np.uint8)
        kind = _arr(parent_kind, np.uint8)
        ppos = _arr(parent_pos, np.uint32)
        rp = _arr(session_rowptr, np.uint64)
        exp = _arr(expected_roots, np.uint8)
        S = (len(rp) - 1) if rp is not None else int(n_sessions)
        n = int(rp[-1]) if rp is not None else S * int(uniform_leaves)
        if exp is not None and exp.size != 32 * S:
            raise ValueError(f"expected_roots holds {exp.size} bytes, need {32 * S}")
        o = dict(out or {})
        o.setdefault("digests", np.empty((n, 32), dtype=np.uint8) if want_digests else None)
        o.setdefault("roots", np.zeros((S, 32), dtype=np.uint8))
        o.setdefault("empty", np.zeros(S, dtype=np.uint8))
        o.setdefault("verdict", np.ones(S, dtype=np.uint8))
        o.setdefault("first_bad", np.full(S, -1, dtype=np.int64))
        o.setdefault("root_match", np.zeros(S, dtype=np.uint8))
        with self._lock:
            self._check(self.lib.ahv_audit_sessions(
                self.h, _ptr(arena), _ptr(offsets), int(fixed_len), _ptr(par), _ptr(kind), _ptr(ppos), _ptr(rp), S,
                int(uniform_leaves), _ptr(exp), _ptr(o["digests"]), _ptr(o["roots"]), _ptr(o["empty"]), _ptr(o["verdict"]),
                _ptr(o["first_bad"]), _ptr(o["root_match"])), "ahv_audit_sessions")
        return o

    def chain_capture(self, arena: np.ndarray, offsets, hexpos, session_rowptr):
        """Patches ``arena`` IN PLACE; returns (digests, parent_digests, parent_kind)."""
        assert arena.dtype == np.uint8 and arena.flags.c_contiguous and arena.flags.writeable
        offsets = _arr(offsets, np.uint64)
        hexpos = _arr(hexpos, np.uint32)
        rp = _arr(session_rowptr, np.uint64)
        S, n = len(rp) - 1, len(offsets) - 1
        digs = np.zeros((n, 32), dtype=np.uint8)
        par = np.zeros((n, 32), dtype=np.uint8)
        kind = np.zeros(n, dtype=np.uint8)
        with self._lock:
            self._check(self.lib.ahv_chain_capture(self.h, _ptr(arena), _ptr(offsets), _ptr(hexpos), _ptr(rp), S,
                                                   _ptr(digs), _ptr(par), _ptr(kind)), "ahv_chain_capture")
        return digs, par, kind

    def sigma_eff_rings(self, rowptr, bonded, active, expiry, now: int, sigma, omega, consensus, current_ring,
                        exact: Optional[bool] = None):
        """ahv_sigma_eff_rings.  exact=True: every row summed in the reference's order (bit-identical);
        exact=False: rows longer than "sigma_long_cut" records take the fixed-shape warp sum (within a few ulp);
        None: whatever the context's "sigma_exact" option says (default: the fast form)."""
        rp = _arr(rowptr, np.uint32)
        A = len(rp) - 1
        bonded = _arr(bonded, np.float64); active = _arr(active, np.uint8); expiry = _arr(expiry, np.int64)
        sigma = _arr(sigma, np.float64); omega = _arr(omega, np.float64)
        consensus = _arr(consensus, np.uint8); current_ring = _arr(current_ring, np.uint8)
        o_s = np.empty(A, np.float64); o_r = np.empty(A, np.uint8)
        o_d = np.empty(A, np.uint8); o_n = np.empty(A, np.uint8)
        with self._lock:
            before = self._options.get("sigma_exact", 0)
            if exact is not None:
                self.set_option("sigma_exact", 1 if exact else 0)
            try:
                self._check(self.lib.ahv_sigma_eff_rings(
                    self.h, _ptr(rp), _ptr(bonded), _ptr(active), _ptr(expiry), int(now), _ptr(sigma), _ptr(omega),
                    _ptr(consensus), _ptr(current_ring), A, _ptr(o_s), _ptr(o_r), _ptr(o_d), _ptr(o_n)),
                    "ahv_sigma_eff_rings")
            finally:
                if exact is not None:
                    self.set_option("sigma_exact", before)
        return o_s, o_r, o_d, o_n

    def exposure_sums(self, rowptr, bonded, active, expiry, now: int, compensated: Optional[bool] = None) -> np.ndarray:
        """Per-row live sums.  compensated=None picks what Python's sum() does under THIS interpreter
        (Neumaier compensation since CPython 3.12), i.e. what the reference would return here."""
        import sys
        if compensated is None:
            compensated = sys.version_info >= (3, 12)
        rp = _arr(rowptr, np.uint32)
        R = len(rp) - 1
        bonded = _arr(bonded, np.float64); active = _arr(active, np.uint8); expiry = _arr(expiry, np.int64)
        out = np.zeros(R, np.float64)
        with self._lock:
            self._check(self.lib.ahv_exposure_sums(self.h, _ptr(rp), _ptr(bonded), _ptr(active), _ptr(expiry), int(now),
                                                   R, 1 if compensated else 0, _ptr(out)), "ahv_exposure_sums")
        return out

    def ring_check(self, agent_ring, action_flags, sigma, consensus, witness):
        ar = _arr(agent_ring, np.uint8); fl = _arr(action_flags, np.uint8); sg = _arr(sigma, np.float64)
        co = _arr(consensus, np.uint8); wi = _arr(witness, np.uint8)
        n = len(ar)
        al = np.empty(n, np.uint8); re = np.empty(n, np.uint8); rq = np.empty(n, np.uint8)
        with self._lock:
            self._check(self.lib.ahv_ring_check(self.h, _ptr(ar), _ptr(fl), _ptr(sg), _ptr(co), _ptr(wi), n,
                                                _ptr(al), _ptr(re), _ptr(rq)), "ahv_ring_check")
        return al, re, rq

    # -- device-pointer entry points (raw ints; async on the ctx stream) -------
    def commit_dev(self, d_arena, arena_bytes, d_offsets, fixed_len, d_rowptr, S, n, uniform, max_leaves,
                   d_digests, d_roots, d_empty):
        self._check(self.lib.ahv_commit_dev(self.h, _vp(d_arena), arena_bytes, _vp(d_offsets), fixed_len,
                                            _vp(d_rowptr), S, n, uniform, max_leaves, _vp(d_digests), _vp(d_roots),
                                            _vp(d_empty)), "ahv_commit_dev")

    def verify_chains_dev(self, d_arena, arena_bytes, d_offsets, fixed_len, d_par, d_kind, d_rowptr, S, n, uniform,
                          max_leaves, d_expected, d_verdict, d_first_bad, d_digests, d_roots, d_match):
        self._check(self.lib.ahv_verify_chains_dev(
            self.h, _vp(d_arena), arena_bytes, _vp(d_offsets), fixed_len, _vp(d_par), _vp(d_kind), _vp(d_rowptr), S,
            n, uniform, max_leaves, _vp(d_expected), _vp(d_verdict), _vp(d_first_bad), _vp(d_digests), _vp(d_roots),
            _vp(d_match)), "ahv_verify_chains_dev")

    def sha256_batch_dev(self, d_arena, arena_bytes, d_offsets, n, fixed_len, d_out):
        self._check(self.lib.ahv_sha256_batch_dev(self.h, _vp(d_arena), arena_bytes, _vp(d_offsets), n, fixed_len,
                                                  _vp(d_out)), "ahv_sha256_batch_dev")

    def merkle_roots_dev(self, d_digests, d_rowptr, S, n, max_leaves, d_roots, d_empty):
        self._check(self.lib.ahv_merkle_roots_dev(self.h, _vp(d_digests), _vp(d_rowptr), S, n, max_leaves,
                                                  _vp(d_roots), _vp(d_empty)), "ahv_merkle_roots_dev")

    def merkle_roots_noalloc_dev(self, d_digests, d_rowptr, S, n, max_leaves, d_roots, d_empty):
        """Same as merkle_roots_dev, but never allocates: scratch is a fixed 8 MiB array in the library."""
        self._check(self.lib.ahv_merkle_roots_noalloc_dev(self.h, _vp(d_digests), _vp(d_rowptr), S, n, max_leaves,
                                                          _vp(d_roots), _vp(d_empty)), "ahv_merkle_roots_noalloc_dev")

    def chain_capture_dev(self, d_arena, arena_bytes, d_offsets, d_hexpos, d_rowptr, S, n, d_digests, d_par, d_kind):
        self._check(self.lib.ahv_chain_capture_dev(self.h, _vp(d_arena), arena_bytes, _vp(d_offsets), _vp(d_hexpos),
                                                   _vp(d_rowptr), S, n, _vp(d_digests), _vp(d_par), _vp(d_kind)),
                    "ahv_chain_capture_dev")

    def slash_cascade(self, rowptr, voucher, active, expiry, now: int, session_agent_rowptr, event_rowptr, event_agent,
                      event_sigma, event_omega, scores, event_depth=None):
        """ahv_slash_cascade.  Inputs are copied; returns dict(scores, touched, active, slash_log, clip_log,
        slash_count, clip_count, status) - session s's slash records start at
        ``event_rowptr[s] + rowptr[a_lo]`` and its clip records at ``rowptr[a_lo]``."""
        rp = _arr(rowptr, np.uint32)
        A = len(rp) - 1
        E = int(rp[-1])
        vo = _arr(voucher, np.uint32)
        ac = np.array(active, dtype=np.uint8, copy=True)
        ex = _arr(expiry, np.int64)
        srp = _arr(session_agent_rowptr, np.uint64); erp = _arr(event_rowptr, np.uint64)
        S = len(srp) - 1
        NE = int(erp[-1])
        ea = _arr(event_agent, np.uint32); es = _arr(event_sigma, np.float64); eo = _arr(event_omega, np.float64)
        ed = _arr(event_depth, np.uint8)
        sc = np.array(scores, dtype=np.float64, copy=True)
        assert ed is None or len(ed) == NE
        assert len(vo) == E and len(ac) == E and len(ex) == E and len(sc) == A and len(ea) == NE
        o = dict(scores=sc, touched=np.zeros(A, np.uint8), active=ac, slash_log=np.zeros(NE + E, dtype=SLASH_DT),
                 clip_log=np.zeros(E, dtype=CLIP_DT), slash_count=np.zeros(S, np.uint32),
                 clip_count=np.zeros(S, np.uint32), status=np.zeros(S, np.uint8))
        b = SlashBatch()
        b.n_agents, b.now, b.n_sessions = A, int(now), S
        for name, arr in (("rowptr", rp), ("voucher", vo), ("active", ac), ("expiry", ex), ("session_agent_rowptr", srp),
                          ("event_rowptr", erp), ("event_agent", ea), ("event_sigma", es), ("event_omega", eo),
                          ("scores", sc), ("touched", o["touched"]), ("slash_log", o["slash_log"]),
                          ("clip_log", o["clip_log"]), ("slash_count", o["slash_count"]),
                          ("clip_count", o["clip_count"]), ("status", o["status"])):
            setattr(b, name, arr.ctypes.data if arr.size else None)
        b.event_depth = ed.ctypes.data if ed is not None and ed.size else None
        with self._lock:
            self._check(self.lib.ahv_slash_cascade(self.h, C.byref(b)), "ahv_slash_cascade")
        return o

    def slash_cascade_dev(self, batch: "SlashBatch"):
        self._check(self.lib.ahv_slash_cascade_dev(self.h, C.byref(batch)), "ahv_slash_cascade_dev")

    def extract_parents_dev(self, d_arena, arena_bytes, d_offsets, fixed_len, d_parent_pos, n, d_par, d_kind):
        self._check(self.lib.ahv_extract_parents_dev(self.h, _vp(d_arena), arena_bytes, _vp(d_offsets), int(fixed_len),
                                                     _vp(d_parent_pos), n, _vp(d_par), _vp(d_kind)),
                    "ahv_extract_parents_dev")

    def extract_parents(self, arena, offsets, parent_pos, fixed_len: int = 0):
        """(parent_digests [n,32], parent_kind [n]) read out of the pre-images on the device."""
        arena = _arr(arena, np.uint8); offsets = _arr(offsets, np.uint64); ppos = _arr(parent_pos, np.uint32)
        n = len(ppos)
        par = np.zeros((n, 32), dtype=np.uint8); kind = np.zeros(n, dtype=np.uint8)
        if n == 0:
            return par, kind
        bufs = [self.dev_alloc(max(arena.nbytes, 1) + 16), self.dev_alloc(8 * (n + 1)), self.dev_alloc(4 * n),
                self.dev_alloc(32 * n), self.dev_alloc(n)]
        try:
            self.h2d(bufs[0], arena); self.h2d(bufs[2], ppos)
            if offsets is not None:
                self.h2d(bufs[1], offsets)
            self.extract_parents_dev(bufs[0], arena.nbytes, bufs[1] if offsets is not None else None, fixed_len,
                                     bufs[2], n, bufs[3], bufs[4])
            self.d2h(par, bufs[3]); self.d2h(kind, bufs[4])
            self.sync()
        finally:
            for b in bufs:
                self.dev_free(b)
        return par, kind

    def sigma_eff_rings_dev(self, d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_cons, d_ring, A,
                            d_os, d_or, d_od, d_on):
        self._check(self.lib.ahv_sigma_eff_rings_dev(
            self.h, _vp(d_rowptr), _vp(d_bonded), _vp(d_active), _vp(d_expiry), int(now), _vp(d_sigma), _vp(d_omega),
            _vp(d_cons), _vp(d_ring), A, _vp(d_os), _vp(d_or), _vp(d_od), _vp(d_on)), "ahv_sigma_eff_rings_dev")


_default_ctx: Optional[Context] = None
_default_lock = threading.Lock()


def get_context(device: Optional[int] = None) -> Context:
    """Process-wide default context (device from AHV_DEVICE / LOCAL_RANK / 0).  Asking for a different
    device than the one the default context already lives on is an error: make a ``Context(device)``."""
    global _default_ctx
    with _default_lock:
        if _default_ctx is None:
            if device is None:
                device = int(os.environ.get("AHV_DEVICE", os.environ.get("LOCAL_RANK", "0")))
            _default_ctx = Context(device)
        elif device is not None and int(device) != _default_ctx.device:
            raise ValueError(f"the default context is on device {_default_ctx.device}, not {device}; "
                             f"create Context({device}) for another GPU")
        return _default_ctx
