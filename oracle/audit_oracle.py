"""hashlib restatement of the Delta Audit Engine hot path (TEST INFRASTRUCTURE).

Every function cites the reference lines it follows
(``/root/reference/src/hypervisor/audit/delta.py`` unless noted).  The product
package never imports this module; see ``oracle/__init__.py``.

Parity: pinned by ``tests/test_oracle_golden.py`` against vectors generated
from the imported reference (``tests/golden/make_golden.py``).
"""

from __future__ import annotations

import hashlib
import json
from typing import Iterable, Optional, Sequence

import numpy as np

PARENT_NONE = 0   # parent_hash is None
PARENT_HEX = 1    # parent_hash is a 64-char lowercase hex string (raw 32 B in `parents`)
PARENT_OTHER = 2  # any other string: can never equal a hexdigest


# --------------------------------------------------------------------------
# canonical serialisation  (delta.py:43-62)
# --------------------------------------------------------------------------
def canonical_preimage(
    delta_id: str,
    turn_id: int,
    session_id: str,
    agent_did: str,
    timestamp_iso: str,
    changes: Sequence[dict],
    parent_hash: Optional[str],
) -> bytes:
    """delta.py:43-62 - json.dumps(..., sort_keys=True).encode().

    ``changes`` items are dicts with exactly path/operation/content_hash/
    previous_hash (VFSChange.agent_did is not serialised, delta.py:51-57).
    """
    payload = json.dumps(
        {
            "delta_id": delta_id,
            "turn_id": turn_id,
            "session_id": session_id,
            "agent_did": agent_did,
            "timestamp": timestamp_iso,
            "changes": [
                {
                    "path": c["path"],
                    "operation": c["operation"],
                    "content_hash": c.get("content_hash"),
                    "previous_hash": c.get("previous_hash"),
                }
                for c in changes
            ],
            "parent_hash": parent_hash,
        },
        sort_keys=True,
    )
    return payload.encode()


# --------------------------------------------------------------------------
# leaf hashing  (delta.py:63)
# --------------------------------------------------------------------------
def leaf_digests(arena, offsets) -> np.ndarray:
    """delta.py:63 over a packed arena: digest[i] = sha256(arena[o[i]:o[i+1]])."""
    mv = memoryview(np.ascontiguousarray(arena, dtype=np.uint8)).cast("B")
    offs = np.asarray(offsets, dtype=np.uint64)
    n = len(offs) - 1
    out = np.empty((n, 32), dtype=np.uint8)
    sha = hashlib.sha256
    for i in range(n):
        out[i] = np.frombuffer(sha(mv[int(offs[i]):int(offs[i + 1])]).digest(), dtype=np.uint8)
    return out


# --------------------------------------------------------------------------
# Merkle root  (delta.py:117-134)
# --------------------------------------------------------------------------
def merkle_root_hex(hashes: Iterable[str]) -> Optional[str]:
    """delta.py:117-134 verbatim on hex strings."""
    hashes = list(hashes)
    if not hashes:                      # :119-120
        return None
    while len(hashes) > 1:              # :125
        next_level = []
        for i in range(0, len(hashes), 2):                          # :127
            left = hashes[i]
            right = hashes[i + 1] if i + 1 < len(hashes) else left  # :129
            next_level.append(hashlib.sha256(f"{left}{right}".encode()).hexdigest())  # :130
        hashes = next_level
    return hashes[0]


def merkle_roots(digests: np.ndarray, session_rowptr) -> tuple[np.ndarray, np.ndarray]:
    """Segmented form: one tree per session.  Returns (roots[S,32], empty[S])."""
    rp = np.asarray(session_rowptr, dtype=np.uint64)
    S = len(rp) - 1
    roots = np.zeros((S, 32), dtype=np.uint8)
    empty = np.zeros(S, dtype=np.uint8)
    d = np.ascontiguousarray(digests, dtype=np.uint8).reshape(-1, 32)
    for s in range(S):
        lo, hi = int(rp[s]), int(rp[s + 1])
        r = merkle_root_hex(d[i].tobytes().hex() for i in range(lo, hi))
        if r is None:
            empty[s] = 1
        else:
            roots[s] = np.frombuffer(bytes.fromhex(r), dtype=np.uint8)
    return roots, empty


# --------------------------------------------------------------------------
# chain verification  (delta.py:136-152)
# --------------------------------------------------------------------------
def verify_chain_session(preimages: Sequence[bytes], parent_hashes: Sequence[Optional[str]]):
    """delta.py:136-152 restated over pre-images.

    ``delta.compute_hash()`` (:140) overwrites ``delta.delta_hash`` before the
    comparison at :141, so that comparison can never fail; what is left is
    ``parent[0] is None`` (:145-147) and ``parent[i] == sha256(preimage[i-1])``
    (:149), the right-hand side being the hash *recomputed in this call*.

    Returns (verdict, first_bad, recomputed) where ``recomputed`` lists the
    hex digests the reference would have written back into
    ``delta.delta_hash`` (indices 0..first_bad inclusive on failure, all on
    success) and first_bad is -1 on success.
    """
    recomputed: list[str] = []
    for i, pre in enumerate(preimages):
        recomputed.append(hashlib.sha256(pre).hexdigest())      # :140 (+ side effect :63)
        if i == 0:
            if parent_hashes[0] is not None:                    # :146
                return False, 0, recomputed
        else:
            if parent_hashes[i] != recomputed[i - 1]:           # :149
                return False, i, recomputed
    return True, -1, recomputed


def classify_parent(parent_hash: Optional[str]) -> tuple[int, bytes]:
    """Host-side encoding of ``parent_hash`` for the side arrays."""
    if parent_hash is None:
        return PARENT_NONE, b"\0" * 32
    if (
        isinstance(parent_hash, str)
        and len(parent_hash) == 64
        and all(ch in "0123456789abcdef" for ch in parent_hash)
    ):
        return PARENT_HEX, bytes.fromhex(parent_hash)
    return PARENT_OTHER, b"\0" * 32


def verify_chains(arena, offsets, parents, parent_kind, session_rowptr):
    """Arena-level effective rule of delta.py:136-152 for many sessions.

    Returns (verdict u8[S], first_bad i64[S], digests u8[n,32]).
    """
    digs = leaf_digests(arena, offsets)
    rp = np.asarray(session_rowptr, dtype=np.uint64)
    par = np.ascontiguousarray(parents, dtype=np.uint8).reshape(-1, 32)
    kind = np.asarray(parent_kind, dtype=np.uint8)
    S = len(rp) - 1
    verdict = np.ones(S, dtype=np.uint8)
    first_bad = np.full(S, -1, dtype=np.int64)
    for s in range(S):
        lo, hi = int(rp[s]), int(rp[s + 1])
        for i in range(lo, hi):
            if i == lo:
                ok = kind[i] == PARENT_NONE
            else:
                ok = kind[i] == PARENT_HEX and bytes(par[i]) == bytes(digs[i - 1])
            if not ok:
                verdict[s] = 0
                first_bad[s] = i - lo
                break
    return verdict, first_bad, digs


# --------------------------------------------------------------------------
# commitment compare  (audit/commitment.py:59-64)
# --------------------------------------------------------------------------
def commitment_verify(committed_root: Optional[str], expected_root: str) -> bool:
    """commitment.py:59-64 - unknown session -> False, else string equality."""
    if not committed_root:
        return False
    return committed_root == expected_root
