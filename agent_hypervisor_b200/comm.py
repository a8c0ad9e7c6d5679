"""Bringing up the library's own NCCL communicator when one process per GPU is started by a launcher
(``torchrun`` sets RANK / WORLD_SIZE / LOCAL_RANK / MASTER_PORT; it is only the process launcher here, nothing
of torch is imported).

Rank 0 makes the 128-byte NCCL id (``ahv_comm_unique_id``) and publishes it through a small file in the
node's temp directory, written atomically; the other ranks of the node poll for it.  The file name carries
the launcher's pid and the rendezvous port, and a file older than the launcher process is ignored, so a
leftover of an earlier crashed run cannot be mistaken for this run's id.  ``ahv_comm_init`` is collective:
when it returns on rank 0 every rank has read the file, and rank 0 removes it.
"""

from __future__ import annotations

import os
import tempfile
import time
from typing import Optional

from ._native import Context, NativeError


def _launcher_start_time() -> float:
    """Wall-clock start of the parent process (the launcher), from /proc; 0.0 if it cannot be read."""
    try:
        with open(f"/proc/{os.getppid()}/stat") as f:
            fields = f.read().rsplit(")", 1)[1].split()
        ticks = int(fields[19])                         # starttime: field 22 of the stat line
        with open("/proc/stat") as f:
            btime = next(int(line.split()[1]) for line in f if line.startswith("btime"))
        return btime + ticks / os.sysconf("SC_CLK_TCK")
    except Exception:
        return 0.0


def id_path(tag: Optional[str] = None) -> str:
    tag = tag or f"{os.getppid()}_{os.environ.get('MASTER_PORT', '0')}_{os.environ.get('TORCHELASTIC_RUN_ID', 'x')}"
    safe = "".join(c if c.isalnum() or c in "_-" else "_" for c in tag)
    return os.path.join(tempfile.gettempdir(), f"ahv_nccl_id_{os.getuid()}_{safe}")


def init_from_env(ctx: Context, rank: Optional[int] = None, world: Optional[int] = None, tag: Optional[str] = None,
                  timeout_s: float = 120.0) -> tuple[int, int]:
    """Create ``ctx``'s communicator from the launcher's environment.  Returns (rank, world)."""
    rank = int(os.environ.get("RANK", "0")) if rank is None else rank
    world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else world
    if world <= 1:
        ctx.comm_init(1, 0)
        return 0, 1
    path = id_path(tag)
    if rank == 0:
        uid = ctx.comm_unique_id()
        tmp = f"{path}.{os.getpid()}.tmp"
        with open(tmp, "wb") as f:
            f.write(uid)
        os.replace(tmp, path)                           # atomic: a reader sees all 128 bytes or no file
    else:
        not_before = _launcher_start_time() - 2.0
        deadline = time.monotonic() + timeout_s
        uid = None
        while uid is None:
            try:
                if os.stat(path).st_mtime >= not_before:
                    with open(path, "rb") as f:
                        data = f.read()
                    if len(data) == 128:
                        uid = data
            except FileNotFoundError:
                pass
            if uid is None:
                if time.monotonic() > deadline:
                    raise NativeError(f"rank {rank}: no NCCL id from rank 0 at {path} after {timeout_s:.0f} s")
                time.sleep(0.005)
    ctx.comm_init(world, rank, uid)
    if rank == 0:
        try:
            os.unlink(path)
        except OSError:
            pass
    return rank, world
