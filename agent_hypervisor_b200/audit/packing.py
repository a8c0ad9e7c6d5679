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


def canonical_preimage(delta) -> bytes:
    """Pre-image of ``SemanticDelta.compute_hash`` (audit/delta.py:43-62) as bytes."""
    changes = ", ".join(
        '{"content_hash": %s, "operation": %s, "path": %s, "previous_hash": %s}'
        % (_enc(c.content_hash), _enc(c.operation), _enc(c.path), _enc(c.previous_hash))
        for c in delta.changes)
    text = ('{"agent_did": %s, "changes": [%s], "delta_id": %s, "parent_hash": %s, '
            '"session_id": %s, "timestamp": %s, "turn_id": %s}'
            % (_enc(delta.agent_did), changes, _enc(delta.delta_id), _enc(delta.parent_hash),
               _enc(delta.session_id), _enc(delta.timestamp.isoformat()), _enc(delta.turn_id)))
    return text.encode()


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

    def add_preimage(self, preimage: bytes, parent_hash) -> None:
        self._buf += preimage
        self._offsets.append(len(self._buf))
        k, raw = classify_parent(parent_hash)
        self._kind.append(k)
        self._parents += raw

    def add_delta(self, delta) -> None:
        self.add_preimage(canonical_preimage(delta), delta.parent_hash)

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
            session_rowptr=np.array(self._rowptr, dtype=np.uint64))


def pack_sessions(sessions: Sequence[Sequence]) -> PackedSessions:
    b = ArenaBuilder()
    for s in sessions:
        b.add_session(s)
    return b.finish()
