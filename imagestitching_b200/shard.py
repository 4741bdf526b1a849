"""Job-level sharding of independent stitch jobs across GPUs (BASELINE configs[4]).

The path has no exchange step: a stitch job never needs another job's pixels, so ranks only agree
on WHICH jobs each one runs.  One process per GPU; assignment is static round-robin
(job j -> rank j % world), which keeps every rank's share within one job of the others.
No collective touches pixel data; `gather_digests` exists only so a caller (or a test) can check
that every job ran exactly once.
"""
from __future__ import annotations

import hashlib
from typing import Dict, Iterable, List, Sequence


def shard_jobs(n_jobs: int, rank: int, world: int) -> List[int]:
    """Indices of the jobs `rank` runs."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_jobs, world))


def digest(buf) -> str:
    """Content hash of one job's exported buffer (used to compare runs across GPU counts)."""
    return hashlib.sha1(memoryview(buf)).hexdigest()


def gather_digests(local: Dict[int, str], world: int) -> Dict[int, str]:
    """All ranks' {job: digest} merged on every rank (torch.distributed must be initialised if world > 1)."""
    if world == 1:
        return dict(local)
    import torch.distributed as dist
    parts: List[Dict[int, str]] = [None] * world  # type: ignore
    dist.all_gather_object(parts, local)
    out: Dict[int, str] = {}
    for p in parts:
        for k, v in p.items():
            if k in out:
                raise RuntimeError(f"job {k} ran on two ranks")
            out[k] = v
    return out
