"""Host packer: canonical pre-images of audit deltas in a contiguous byte arena.

The bytes must equal what the reference hashes at ``audit/delta.py:43-63``:
``json.dumps({...7 keys...}, sort_keys=True).encode()`` - keys sorted at both nesting
levels, separators ``", "`` / ``": "``, ``ensure_ascii`` escaping, ``None -> null``,
``timestamp.isoformat()``, ``VFSChange.agent_did`` not serialised.  Because the key set is
fixed, the emitter below writes the sorted layout directly and only calls the JSON encoder
for the values (same CPython encoder, so escaping is identical by construction).
"""

from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Iterable, Optional, Sequence

import numpy as np

from .. import _native as nat

_enc_str = json.encoder.encode_basestring_ascii     # what json.dumps uses when ensure_ascii=True
_HEX = frozenset("0123456789abcdef")


def _enc(v) -> str:
    if type(v) is str:
        return _enc_str(v)
    if v is None:
        return "null"
    return json.dumps(v, sort_keys=True)              # ints, floats, bools, nested containers


PARENT_POS_NONE = 0xFFFFFFFF     # parent_pos of a delta whose parent_hash is None


def canonical_preimage_pos(delta) -> tuple[bytes, int]:
    """Pre-image of ``SemanticDelta.compute_hash`` (audit/delta.py:43-62) as bytes, and the position in it
    of the first character of the ``parent_hash`` value (past the opening quote of a string;
    ``PARENT_POS_NONE`` for None): the ``parent_pos`` entry ``ahv_audit_sessions`` reads the parent from."""
    changes = ", ".join(
        '{"content_hash": %s, "operation": %s, "path": %s, "previous_hash": %s}'
        % (_enc(c.content_hash), _enc(c.operation), _enc(c.path), _enc(c.previous_hash))
        for c in delta.changes)
    head = ('{"agent_did": %s, "changes": [%s], "delta_id": %s, "parent_hash": '
            % (_enc(delta.agent_did), changes, _enc(delta.delta_id)))
    ph = delta.parent_hash
    text = (head + '%s, "session_id": %s, "timestamp": %s, "turn_id": %s}'
            % (_enc(ph), _enc(delta.session_id), _enc(delta.timestamp.isoformat()), _enc(delta.turn_id)))
    # ensure_ascii: every character is one byte, so string positions are byte positions
    pos = PARENT_POS_NONE if ph is None else len(head) + (1 if type(ph) is str else 0)
    return text.encode(), pos


def canonical_preimage(delta) -> bytes:
    """Pre-image of ``SemanticDelta.compute_hash`` (audit/delta.py:43-62) as bytes."""
    return canonical_preimage_pos(delta)[0]


def classify_parent(parent_hash) -> tuple[int, bytes]:
    """Side-array encoding of ``parent_hash``: (kind, 32 raw bytes)."""
    if parent_hash is None:
        return nat.PARENT_NONE, b"\0" * 32
    if type(parent_hash) is str and len(parent_hash) == 64 and _HEX.issuperset(parent_hash):
        return nat.PARENT_HEX, bytes.fromhex(parent_hash)
    return nat.PARENT_OTHER, b"\0" * 32


def is_hexdigest(s) -> bool:
    return type(s) is str and len(s) == 64 and _HEX.issuperset(s)


@dataclass
class PackedSessions:
    arena: np.ndarray            # uint8[total]
    offsets: np.ndarray          # uint64[n+1]
    parents: np.ndarray          # uint8[n,32]
    parent_kind: np.ndarray      # uint8[n]
    session_rowptr: np.ndarray   # uint64[S+1]
    parent_pos: Optional[np.ndarray] = None   # uint32[n]; None when a pre-image was added without its position

    @property
    def n(self) -> int:
        return len(self.offsets) - 1

    @property
    def n_sessions(self) -> int:
        return len(self.session_rowptr) - 1

    def uniform_shape(self) -> tuple[int, int]:
        """(fixed_len, uniform_leaves), each 0 when not uniform - lets callers take the fast paths."""
        lens = np.diff(self.offsets.astype(np.int64))
        fixed = int(lens[0]) if len(lens) and np.all(lens == lens[0]) else 0
        cnt = np.diff(self.session_rowptr.astype(np.int64))
        uni = int(cnt[0]) if len(cnt) and cnt[0] > 0 and np.all(cnt == cnt[0]) else 0
        return fixed, uni


class ArenaBuilder:
    """Accumulates sessions of deltas into one arena (+ offsets, parent side arrays, CSR)."""

    def __init__(self) -> None:
        self._buf = bytearray()
        self._offsets = [0]
        self._parents = bytearray()
        self._kind = bytearray()
        self._rowptr = [0]
        self._pos: Optional[list] = []

    def add_preimage(self, preimage: bytes, parent_hash, parent_pos: Optional[int] = None) -> None:
        self._buf += preimage
        self._offsets.append(len(self._buf))
        k, raw = classify_parent(parent_hash)
        self._kind.append(k)
        self._parents += raw
        if parent_pos is None:
            self._pos = None                       # positions unknown: callers use the side arrays
        elif self._pos is not None:
            self._pos.append(parent_pos)

    def add_delta(self, delta) -> None:
        pre, pos = canonical_preimage_pos(delta)
        self.add_preimage(pre, delta.parent_hash, pos)

    def end_session(self) -> None:
        self._rowptr.append(len(self._kind))

    def add_session(self, deltas: Iterable) -> None:
        for d in deltas:
            self.add_delta(d)
        self.end_session()

    def finish(self) -> PackedSessions:
        if self._rowptr[-1] != len(self._kind):
            self.end_session()
        n = len(self._kind)
        return PackedSessions(
            arena=np.frombuffer(bytes(self._buf), dtype=np.uint8),
            offsets=np.array(self._offsets, dtype=np.uint64),
            parents=np.frombuffer(bytes(self._parents), dtype=np.uint8).reshape(n, 32),
            parent_kind=np.frombuffer(bytes(self._kind), dtype=np.uint8),
            session_rowptr=np.array(self._rowptr, dtype=np.uint64),
            parent_pos=None if self._pos is None else np.array(self._pos, dtype=np.uint32))


def pack_sessions(sessions: Sequence[Sequence]) -> PackedSessions:
    b = ArenaBuilder()
    for s in sessions:
        b.add_session(s)
    return b.finish()


# ---------------------------------------------------------------------------------------------
# Columnar path: the C++ host packer (ahv_pack_deltas) for bulk import / replay, where deltas
# arrive as columns rather than Python objects.  ``columns_from_deltas`` exists for tests and
# small batches; a bulk producer fills the same arrays directly.
# ---------------------------------------------------------------------------------------------
import ctypes as _C
from datetime import datetime as _dt, timezone as _tz

_EPOCH = _dt(1970, 1, 1, tzinfo=_tz.utc)
_NAIVE = -(2 ** 31)


class _DeltaColumns(_C.Structure):
    _fields_ = [("n", _C.c_uint64)] + [(nm, _C.c_void_p) for nm in (
        "delta_id_blob", "delta_id_off", "session_id_blob", "session_id_off", "agent_did_blob", "agent_did_off",
        "turn_id", "timestamp_us", "timestamp_utcoffset_s", "parent_kind", "parent_digest",
        "parent_other_blob", "parent_other_off", "change_rowptr",
        "change_path_blob", "change_path_off", "change_op_blob", "change_op_off",
        "change_content_blob", "change_content_off", "change_content_null",
        "change_prev_blob", "change_prev_off", "change_prev_null")]


def _str_col(values):
    enc = [b"" if v is None else v.encode("utf-8", "surrogatepass") for v in values]
    off = np.zeros(len(enc) + 1, dtype=np.uint64)
    if enc:
        np.cumsum([len(e) for e in enc], out=off[1:])
    blob = np.frombuffer(b"".join(enc) or b"\0", dtype=np.uint8)
    null = np.array([1 if v is None else 0 for v in values], dtype=np.uint8) if values else np.zeros(1, np.uint8)
    return blob, off, null


def columns_from_deltas(deltas) -> dict:
    """Columnar form of SemanticDelta-like objects with string ids (ints/other JSON types in id fields
    are not representable as columns: use canonical_preimage for those)."""
    changes = [c for d in deltas for c in d.changes]
    ts_us, ts_off = [], []
    for d in deltas:
        t = d.timestamp
        if t.tzinfo is None:
            delta = t - _EPOCH.replace(tzinfo=None)
            ts_off.append(_NAIVE)
        else:
            delta = t - _EPOCH
            ts_off.append(int(t.utcoffset().total_seconds()))
        ts_us.append((delta.days * 86400 + delta.seconds) * 1_000_000 + delta.microseconds)
    kinds, digs, others = [], bytearray(), []
    for d in deltas:
        k, raw = classify_parent(d.parent_hash)
        kinds.append(k); digs += raw
        others.append(d.parent_hash if k == nat.PARENT_OTHER else "")
    rowptr = np.zeros(len(deltas) + 1, dtype=np.uint64)
    if deltas:
        np.cumsum([len(d.changes) for d in deltas], out=rowptr[1:])
    return {
        "n": len(deltas),
        "delta_id": _str_col([d.delta_id for d in deltas]), "session_id": _str_col([d.session_id for d in deltas]),
        "agent_did": _str_col([d.agent_did for d in deltas]),
        "turn_id": np.array([d.turn_id for d in deltas], dtype=np.int64),
        "timestamp_us": np.array(ts_us, dtype=np.int64), "timestamp_utcoffset_s": np.array(ts_off, dtype=np.int32),
        "parent_kind": np.array(kinds, dtype=np.uint8),
        "parent_digest": np.frombuffer(bytes(digs) or b"\0" * 32, dtype=np.uint8),
        "parent_other": _str_col(others), "change_rowptr": rowptr,
        "change_path": _str_col([c.path for c in changes]), "change_op": _str_col([c.operation for c in changes]),
        "change_content": _str_col([c.content_hash for c in changes]),
        "change_prev": _str_col([c.previous_hash for c in changes]),
    }


def pack_columns(cols: dict, arena: Optional[np.ndarray] = None, want_parent_pos: bool = False,
                 offsets: Optional[np.ndarray] = None, parent_pos: Optional[np.ndarray] = None):
    """Canonical pre-images of ``cols['n']`` deltas via the C++ packer -> (arena u8[], offsets u64[n+1])
    and, with ``want_parent_pos``, the uint32[n] positions of the parent values as a third element.
    ``arena`` / ``offsets`` / ``parent_pos`` may be caller-owned (e.g. pinned) arrays to write into."""
    lib = nat.load_library()
    keep = []

    def ptr(a):
        a = np.ascontiguousarray(a)
        keep.append(a)
        return a.ctypes.data

    st = _DeltaColumns()
    st.n = cols["n"]
    for nm in ("delta_id", "session_id", "agent_did", "parent_other"):
        blob, off, _ = cols[nm]
        setattr(st, nm + "_blob", ptr(blob)); setattr(st, nm + "_off", ptr(off))
    for nm in ("turn_id", "timestamp_us", "timestamp_utcoffset_s", "parent_kind", "parent_digest", "change_rowptr"):
        setattr(st, nm, ptr(cols[nm]))
    for nm, key in (("change_path", "change_path"), ("change_op", "change_op")):
        blob, off, _ = cols[key]
        setattr(st, nm + "_blob", ptr(blob)); setattr(st, nm + "_off", ptr(off))
    for nm in ("change_content", "change_prev"):
        blob, off, null = cols[nm]
        setattr(st, nm + "_blob", ptr(blob)); setattr(st, nm + "_off", ptr(off)); setattr(st, nm + "_null", ptr(null))
    n = cols["n"]
    if offsets is None:
        offsets = np.zeros(n + 1, dtype=np.uint64)
    else:
        assert offsets.dtype == np.uint64 and offsets.flags.c_contiguous and len(offsets) >= n + 1
        offsets = offsets[: n + 1]
    nbytes = _C.c_uint64()
    if parent_pos is not None:
        assert parent_pos.dtype == np.uint32 and parent_pos.flags.c_contiguous and len(parent_pos) >= n
        want_parent_pos = True
    ppos = (parent_pos if parent_pos is not None else np.zeros(max(n, 1), dtype=np.uint32)) if want_parent_pos else None
    rc = lib.ahv_pack_deltas(_C.byref(st), offsets.ctypes.data, None, 0, _C.byref(nbytes),
                             ppos.ctypes.data if want_parent_pos else None, 0)
    if rc:
        raise nat.NativeError(lib.ahv_last_error().decode())
    if arena is None:
        arena = np.empty(max(int(nbytes.value), 1), dtype=np.uint8)
    rc = lib.ahv_pack_deltas(_C.byref(st), offsets.ctypes.data, arena.ctypes.data, arena.nbytes, _C.byref(nbytes), None,
                             1)      # AHV_PACK_OFFSETS_VALID: the size query above filled offsets
    if rc:
        raise nat.NativeError(lib.ahv_last_error().decode())
    if want_parent_pos:
        return arena[: int(nbytes.value)], offsets, ppos[:n]
    return arena[: int(nbytes.value)], offsets


# ---------------------------------------------------------------------------------------------
# Columns on the device: one pinned staging buffer holding every column array (16-byte aligned
# pieces), mirrored by one device buffer, so that a chunk of columns travels in ONE copy and the
# device packer (ahv_pack_deltas_dev) reads it in place.
# ---------------------------------------------------------------------------------------------
_COLUMN_ARRAYS = (
    ("delta_id_blob", "delta_id", 0), ("delta_id_off", "delta_id", 1), ("session_id_blob", "session_id", 0),
    ("session_id_off", "session_id", 1), ("agent_did_blob", "agent_did", 0), ("agent_did_off", "agent_did", 1),
    ("turn_id", "turn_id", None), ("timestamp_us", "timestamp_us", None), ("timestamp_utcoffset_s", "timestamp_utcoffset_s", None),
    ("parent_kind", "parent_kind", None), ("parent_digest", "parent_digest", None),
    ("parent_other_blob", "parent_other", 0), ("parent_other_off", "parent_other", 1), ("change_rowptr", "change_rowptr", None),
    ("change_path_blob", "change_path", 0), ("change_path_off", "change_path", 1), ("change_op_blob", "change_op", 0),
    ("change_op_off", "change_op", 1), ("change_content_blob", "change_content", 0), ("change_content_off", "change_content", 1),
    ("change_content_null", "change_content", 2), ("change_prev_blob", "change_prev", 0), ("change_prev_off", "change_prev", 1),
    ("change_prev_null", "change_prev", 2))


def column_bytes(cols: dict) -> int:
    """Bytes ``stage_columns`` needs for these columns."""
    total = 0
    for _, key, part in _COLUMN_ARRAYS:
        a = cols[key] if part is None else cols[key][part]
        total += (np.asarray(a).nbytes + 15) & ~15
    return total + 16


def stage_columns(cols: dict, staging: np.ndarray):
    """Copy every column array into ``staging`` (a pinned uint8 buffer).  Returns (layout, used_bytes) where
    layout maps struct field -> byte offset inside the buffer; the same offsets hold on the device copy."""
    layout, pos = {}, 0
    for field, key, part in _COLUMN_ARRAYS:
        a = np.ascontiguousarray(cols[key] if part is None else cols[key][part])
        nb = a.nbytes
        if pos + nb > staging.nbytes:
            raise ValueError(f"column staging buffer of {staging.nbytes} bytes is too small")
        staging[pos: pos + nb] = a.view(np.uint8).reshape(-1)
        layout[field] = pos
        pos += (nb + 15) & ~15
    return layout, pos


def device_columns_struct(n: int, layout: dict, d_base: int) -> "_DeltaColumns":
    st = _DeltaColumns()
    st.n = n
    for field, off in layout.items():
        setattr(st, field, d_base + off)
    return st
