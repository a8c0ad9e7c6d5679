"""Test helpers: turn recipe field dicts into packed arenas the way the
reference would serialise them (via the ORACLE's restatement of
delta.py:43-62), chaining parents with hashlib."""

from __future__ import annotations

import hashlib

import numpy as np

from oracle import audit_oracle as ao


def chain_preimages(fields):
    """[(preimage bytes, parent_hash str|None)] with parent = sha256(prev preimage)."""
    out = []
    parent = None
    for f in fields:
        pre = ao.canonical_preimage(
            f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
            f["timestamp"].isoformat(), f["changes"], parent)
        out.append((pre, parent))
        parent = hashlib.sha256(pre).hexdigest()
    return out


def pack_sessions(sessions):
    """sessions: list of list[(preimage, parent_hash)] -> arena, offsets, parents, kind, rowptr."""
    blobs, offs, par, kind, rp = [], [0], [], [], [0]
    for sess in sessions:
        for pre, parent in sess:
            blobs.append(pre)
            offs.append(offs[-1] + len(pre))
            k, raw = ao.classify_parent(parent)
            kind.append(k)
            par.append(raw)
        rp.append(len(kind))
    arena = np.frombuffer(b"".join(blobs), dtype=np.uint8).copy() if blobs else np.zeros(0, np.uint8)
    parents = np.frombuffer(b"".join(par), dtype=np.uint8).reshape(-1, 32).copy() if par else np.zeros((0, 32), np.uint8)
    return (arena, np.array(offs, dtype=np.uint64), parents,
            np.array(kind, dtype=np.uint8), np.array(rp, dtype=np.uint64))


def config2_session(s: int, L: int = 1000):
    """chained, every pre-image exactly 256 B."""
    import recipes
    out = []
    parent = None
    for i in range(L):
        f0 = recipes.config2_delta(s, i, 0)
        base = ao.canonical_preimage(f0["delta_id"], f0["turn_id"], f0["session_id"], f0["agent_did"],
                                     f0["timestamp"].isoformat(), [], parent)
        pad = recipes.CONFIG2_LEN - len(base)
        f = recipes.config2_delta(s, i, pad)
        pre = ao.canonical_preimage(f["delta_id"], f["turn_id"], f["session_id"], f["agent_did"],
                                    f["timestamp"].isoformat(), [], parent)
        assert len(pre) == recipes.CONFIG2_LEN
        out.append((pre, parent))
        parent = hashlib.sha256(pre).hexdigest()
    return out


def hexes(digests: np.ndarray):
    return [bytes(d).hex() for d in np.asarray(digests, dtype=np.uint8).reshape(-1, 32)]
