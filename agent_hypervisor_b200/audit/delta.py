"""Delta Audit Engine with the reference's API (``audit/delta.py:17-160``), hashing on the GPU.

Same classes, fields, signatures, return values and quirks as the reference:

* ``SemanticDelta.compute_hash`` sets and returns ``delta_hash`` (:41-64);
* ``DeltaEngine.capture`` chains ``parent_hash`` to the previous ``delta_hash`` (:85-115);
* ``compute_merkle_root`` hashes hex-concatenated pairs, duplicates an odd last node, returns
  the single leaf for one delta and ``None`` for none (:117-134);
* ``verify_chain`` re-hashes every delta *and writes the fresh hash back* before comparing, so it
  checks only the parent links, stops at the first bad one, and cannot see a tampered last delta
  (:136-152).

All SHA-256 work goes through ``libahv_b200.so``; there is no hashlib path.
"""

from __future__ import annotations

from dataclasses import dataclass
from datetime import datetime, timezone
from typing import Optional

import numpy as np

from .. import _native as nat
import re

from .packing import ArenaBuilder, canonical_preimage

_LOWER_HEX = re.compile(r"[0-9a-f]*")


@dataclass
class VFSChange:
    """audit/delta.py:17-25."""

    path: str
    operation: str  # "add", "modify", "delete", "permission"
    content_hash: Optional[str] = None
    previous_hash: Optional[str] = None
    agent_did: Optional[str] = None


@dataclass
class SemanticDelta:
    """audit/delta.py:28-64."""

    delta_id: str
    turn_id: int
    session_id: str
    agent_did: str
    timestamp: datetime
    changes: list[VFSChange]
    parent_hash: Optional[str]
    delta_hash: str = ""

    def preimage(self) -> bytes:
        return canonical_preimage(self)

    def compute_hash(self) -> str:
        self.delta_hash = nat.get_context().sha256_one(self.preimage()).hex()
        return self.delta_hash


def _merkle_root_of_strings(hashes: list[str]) -> Optional[str]:
    """compute_merkle_root (:117-134) for arbitrary ``delta_hash`` strings.

    Normal case (every hash a 64-char lowercase hexdigest): one device tree over raw digests.
    If a hash was overwritten with some other string, the reference would still hash the string
    concatenation at the first level; that level is then hashed as variable-length messages on the
    device and the rest of the tree proceeds on digests.
    """
    if not hashes:
        return None
    if len(hashes) == 1:
        return hashes[0]
    ctx = nat.get_context()
    joined = "".join(hashes) if all(type(h) is str for h in hashes) else ""
    if len(joined) == 64 * len(hashes) and set(map(len, hashes)) == {64} and _LOWER_HEX.fullmatch(joined):
        digs = np.frombuffer(bytes.fromhex(joined), dtype=np.uint8).reshape(-1, 32)
    else:
        msgs = []
        for i in range(0, len(hashes), 2):
            left = hashes[i]
            right = hashes[i + 1] if i + 1 < len(hashes) else left
            msgs.append(f"{left}{right}".encode())
        offs = np.zeros(len(msgs) + 1, dtype=np.uint64)
        np.cumsum([len(m) for m in msgs], out=offs[1:])
        digs = ctx.sha256_batch(np.frombuffer(b"".join(msgs), dtype=np.uint8), offs)
        if len(digs) == 1:
            return digs[0].tobytes().hex()
    roots, _ = ctx.merkle_roots(digs, np.array([0, len(digs)], dtype=np.uint64))
    return roots[0].tobytes().hex()


class DeltaEngine:
    """Per-session hash-chained audit log (audit/delta.py:67-160)."""

    def __init__(self, session_id: str) -> None:
        self.session_id = session_id
        self._deltas: list[SemanticDelta] = []
        self._turn_counter = 0

    def capture(self, agent_did: str, changes: list[VFSChange], delta_id: Optional[str] = None) -> SemanticDelta:
        self._turn_counter += 1
        parent = self._deltas[-1].delta_hash if self._deltas else None
        delta = SemanticDelta(
            delta_id=delta_id or f"delta:{self._turn_counter}",
            turn_id=self._turn_counter,
            session_id=self.session_id,
            agent_did=agent_did,
            timestamp=datetime.now(timezone.utc),
            changes=changes,
            parent_hash=parent,
        )
        delta.compute_hash()
        self._deltas.append(delta)
        return delta

    def compute_merkle_root(self) -> Optional[str]:
        return _merkle_root_of_strings([d.delta_hash for d in self._deltas])

    def verify_chain(self) -> bool:
        if not self._deltas:
            return True
        b = ArenaBuilder()
        b.add_session(self._deltas)
        p = b.finish()
        res = nat.get_context().verify_chains(p.arena, p.offsets, p.parents, p.parent_kind, p.session_rowptr)
        ok = bool(res["verdict"][0])
        # state parity with the reference: compute_hash() refreshed delta_hash for every delta it
        # reached, i.e. all of them on success and indices <= first_bad on failure (:140, :63)
        last = len(self._deltas) - 1 if ok else int(res["first_bad"][0])
        digs = res["digests"]
        for i in range(last + 1):
            self._deltas[i].delta_hash = digs[i].tobytes().hex()
        return ok

    @property
    def deltas(self) -> list[SemanticDelta]:
        return list(self._deltas)

    @property
    def turn_count(self) -> int:
        return self._turn_counter
