"""The gang a job runs on, visible to the workload callbacks.

The reference knows exactly one execution shape: a job owns one GPU (swarm/worker.py:113-128, `device(func,
**kwargs)` in swarm/gpu/device.py:33-50).  Here a job may own an aligned gang of 2 / 4 / 8 ranks leased from
`gpu.device_pool.RankGroup`; every rank of that gang runs the same callback with the same arguments (SPMD) and the
callbacks ask this module whether they have peers:

    gang = job_gang()        # parallel.runtime.Gang (rank inside the gang, world, process group) or None
    if gang is not None: pipe.enable_patch_parallel(gang) / split the batch / scatter the frames

Rank processes execute one job at a time, so the slot is process-global.
"""
from __future__ import annotations

import contextlib
from typing import Optional

_gang = None


def job_gang():
    """Gang leased to the job this process is executing; None for a single-GPU (data-parallel) job."""
    return _gang


def is_leader() -> bool:
    """Exactly one rank of a gang reports the result to the hive."""
    return _gang is None or _gang.rank == 0


@contextlib.contextmanager
def running_on(gang: Optional[object]):
    global _gang
    prev, _gang = _gang, (gang if gang is not None and getattr(gang, "world", 1) > 1 else None)
    try:
        yield
    finally:
        _gang = prev
