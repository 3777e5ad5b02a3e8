"""Topology-aware rank group -- the reference's dead device pool (swarm/gpu/device_pool.py:1-34:
a mutex-protected free list with add/remove/count that nothing imports) reborn as the gang
scheduler of an 8 x B200 NVSwitch box.

A job leases either one rank (data-parallel: one resident pipeline per GPU) or an aligned gang of
2 / 4 / 8 ranks (patch / sequence / tensor parallel: one image or clip across GPUs).  NVSwitch
gives every subset the same bandwidth, but symmetric-memory arenas, multicast objects and process
groups are created per subset, so only the aligned buddy subsets {0-1, 2-3, ...}, {0-3, 4-7},
{0-7} are ever leased: they are pre-created at start-up and never fragment each other
(SURVEY.md section 5.8).  The reference's three functions are kept as the single-device view.
"""
from __future__ import annotations

import threading
from dataclasses import dataclass, field
from typing import List, Optional

from .device import Device

VALID_GANG_SIZES = (1, 2, 4, 8)


@dataclass(frozen=True)
class Gang:
    """An aligned set of ranks leased to one job."""

    ranks: tuple
    lease_id: int = 0

    @property
    def size(self) -> int:
        return len(self.ranks)

    @property
    def leader(self) -> int:
        return self.ranks[0]


@dataclass
class _State:
    world: int
    free: List[bool] = field(default_factory=list)
    next_lease: int = 1


class RankGroup:
    """Buddy allocator over `world` ranks (world must be a power of two <= 8, or any count for
    size-1 leases)."""

    def __init__(self, world: int):
        if world < 1:
            raise ValueError("world must be >= 1")
        self._cv = threading.Condition()
        self._s = _State(world=world, free=[True] * world)
        self._leases = {}
        self._dead = set()
        self._waiting: List[int] = []   # tickets of lease() calls in arrival order: leases are granted first come, first
        self._next_ticket = 0           # served, so a stream of one-rank jobs cannot starve a job that waits for a gang

    @property
    def world(self) -> int:
        return self._s.world

    def aligned_subsets(self, size: int) -> List[tuple]:
        if size not in VALID_GANG_SIZES:
            raise ValueError(f"gang size must be one of {VALID_GANG_SIZES}, got {size}")
        return [tuple(range(b, b + size)) for b in range(0, self._s.world - size + 1, size)]

    def free_count(self) -> int:
        with self._cv:
            return sum(self._s.free)

    def _try_lease(self, size: int) -> Optional[Gang]:
        best = None
        for subset in self.aligned_subsets(size):
            if all(self._s.free[r] for r in subset):
                # prefer the subset whose buddy is already busy: keeps large aligned blocks intact
                buddy_base = (subset[0] // (size * 2)) * size * 2
                buddy = [r for r in range(buddy_base, min(buddy_base + size * 2, self._s.world)) if r not in subset]
                fragmentation = sum(1 for r in buddy if self._s.free[r])
                if best is None or fragmentation < best[0]:
                    best = (fragmentation, subset)
        if best is None:
            return None
        subset = best[1]
        for r in subset:
            self._s.free[r] = False
        gang = Gang(ranks=subset, lease_id=self._s.next_lease)
        self._s.next_lease += 1
        self._leases[gang.lease_id] = gang
        return gang

    def lease(self, size: int = 1, timeout: Optional[float] = None) -> Gang:
        """Blocks until an aligned gang of `size` ranks is free AND every earlier caller has been served (first come, first
        served: while an 8-rank job waits for the box to drain, later one-rank jobs queue up behind it instead of keeping
        ranks busy forever).  Raises Exception("busy") on timeout (the reference's pool raises the same when empty,
        swarm/gpu/device_pool.py:30-32)."""
        if size > self._s.world:
            raise ValueError(f"gang of {size} does not fit a world of {self._s.world}")
        with self._cv:
            ticket = self._next_ticket
            self._next_ticket += 1
            self._waiting.append(ticket)
            try:
                def ready():
                    return self._waiting[0] == ticket and self._peek(size)

                if not ready():
                    if timeout == 0 or not self._cv.wait_for(ready, timeout):
                        raise Exception("busy")
                return self._try_lease(size)
            finally:
                self._waiting.remove(ticket)
                self._cv.notify_all()  # the next caller in line may be servable now

    def _peek(self, size: int) -> bool:
        return any(all(self._s.free[r] for r in s) for s in self.aligned_subsets(size))

    def release(self, gang: Gang) -> None:
        with self._cv:
            if self._leases.pop(gang.lease_id, None) is None:
                raise ValueError("gang is not leased (double release?)")
            for r in gang.ranks:
                self._s.free[r] = r not in self._dead  # (a rank that died while leased does not come back)
            self._cv.notify_all()

    def mark_dead(self, rank: int) -> List[Gang]:
        """Failure handling: a dead rank is withdrawn; returns the leases it was part of so the
        caller can fail those jobs non-fatally."""
        with self._cv:
            hit = [g for g in self._leases.values() if rank in g.ranks]
            self._s.free[rank] = False
            self._dead.add(rank)
            return hit

    def dead_ranks(self) -> List[int]:
        with self._cv:
            return sorted(self._dead)


# ---- single-device view with the reference's function names --------------------------------
mutex = threading.Lock()
available: List[Device] = []


def get_available_gpu_count() -> int:
    with mutex:
        return len(available)


def add_device_to_pool(device: Device) -> None:
    with mutex:
        available.append(device)


def remove_device_from_pool() -> Device:
    with mutex:
        if available:
            return available.pop(0)
        raise Exception("busy")
