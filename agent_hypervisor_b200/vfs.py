"""Batched VFS content hashing: ``session/sso.py:214-216`` (``_hash(content) =
sha256(content.encode()).hexdigest()``), the ``content_hash`` / ``previous_hash`` values that
``SessionVFS.write`` (sso.py:49-69) puts into the edits which later become audit deltas.
One GPU launch for many file contents (variable length; the library length-sorts the batch).
The VFS itself (paths, permissions, snapshots) is control plane and stays in the reference.

A message is hashed by one thread, block after block (SHA-256 has no parallelism inside a message): about
2.3 us per 64 bytes, i.e. ~28 MB/s per message, and a launch lasts as long as its longest message (the
length-sorted launch order only makes the long ones start first).  The pass pays off for MANY contents
(10^7 leaves of 64 B - 16 KB: 990 GB/s); one multi-megabyte file is a job for a CPU core with SHA-NI.
"""

from __future__ import annotations

from typing import Optional, Sequence, Union

import numpy as np

from . import _native as nat


def content_hashes(contents: Sequence[Union[str, bytes]], ctx: Optional[nat.Context] = None) -> list[str]:
    """Hex SHA-256 of every content string (UTF-8 encoded like ``str.encode()``)."""
    if not contents:
        return []
    ctx = ctx or nat.get_context()
    blobs = [c.encode() if isinstance(c, str) else bytes(c) for c in contents]
    offsets = np.zeros(len(blobs) + 1, dtype=np.uint64)
    np.cumsum([len(b) for b in blobs], out=offsets[1:])
    arena = np.frombuffer(b"".join(blobs) or b"\0", dtype=np.uint8)
    digs = ctx.sha256_batch(arena, offsets)
    return [d.tobytes().hex() for d in digs]
