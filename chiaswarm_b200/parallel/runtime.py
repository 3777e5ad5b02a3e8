"""Rank runtime: one process per GPU, NCCL for plumbing, symmetric memory for in-kernel
communication (SURVEY.md sections 2.3, 5.8; the reference has no communication layer at all --
its only parallelism is one asyncio task per GPU in a single process, swarm/worker.py:41-43).

* `init()` joins the job from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE /
  MASTER_ADDR / MASTER_PORT), NCCL on GPUs and gloo on CPU (tests).
* `SymmetricArena` is one symmetric-memory allocation per gang (torch's `_symmetric_memory`:
  every rank maps every peer's buffer, NVLink P2P through NVSwitch).  Tensors carved from it sit
  at the same offset on every rank, so a kernel reaches "the same tensor on rank d" as
  `peer_base[d] + offset`.  A small flag area holds monotonically increasing epoch counters
  written with `st.release.sys` and polled with `ld.acquire.sys` from inside kernels.
* `Gang` wraps a process (sub)group of 1 / 2 / 4 / 8 aligned ranks (gpu/device_pool.RankGroup
  hands those out).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

FLAG_SLOTS = 4096  # int32 flags per rank


def env_rank() -> int:
    return int(os.environ.get("RANK", "0"))


def env_world() -> int:
    return int(os.environ.get("WORLD_SIZE", "1"))


def env_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", "0"))


def init(backend: Optional[str] = None) -> "Gang":
    """Joins the torchrun job (no-op world of 1 without it) and returns the world gang."""
    world = env_world()
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(env_local_rank())
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = backend or ("nccl" if use_cuda else "gloo")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", env_local_rank())
        dist.init_process_group(backend, **kw)
    device = torch.device("cuda", env_local_rank()) if use_cuda else torch.device("cpu")
    return Gang(rank=env_rank(), world=world, group=dist.group.WORLD if world > 1 else None, device=device,
                global_ranks=tuple(range(world)))


@dataclass
class Gang:
    rank: int
    world: int
    group: Optional[object]
    device: torch.device
    global_ranks: tuple = ()   # world ranks of the members (identity of the gang: keys engine caches)

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)

    def all_reduce_max(self, value: float) -> float:
        if self.world == 1:
            return value
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def all_gather_cat(self, t: torch.Tensor) -> torch.Tensor:
        """Concatenates every rank's `t` along dim 0 in rank order; leading dims may differ between ranks (the
        batch / frame scatter of gang jobs hands the last rank a shorter slice)."""
        if self.world == 1:
            return t
        t = t.contiguous()
        counts = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        all_counts = [torch.zeros_like(counts) for _ in range(self.world)]
        dist.all_gather(all_counts, counts, group=self.group)
        counts = [int(c.item()) for c in all_counts]
        cap = max(counts)
        if t.shape[0] < cap:
            pad = torch.zeros((cap - t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            t = torch.cat([t, pad])
        parts = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(parts, t, group=self.group)
        return torch.cat([p[:c] for p, c in zip(parts, counts)])

    def shard(self, count: int) -> range:
        """This rank's contiguous share of `count` items (the first ranks take the remainder)."""
        base, extra = divmod(count, self.world)
        lo = self.rank * base + min(self.rank, extra)
        return range(lo, lo + base + (1 if self.rank < extra else 0))

    def subgang(self, ranks: Sequence[int]) -> "Gang":
        """Aligned sub-gang (every rank of the parent must call this with the same list)."""
        ranks = list(ranks)
        if len(ranks) == self.world:
            return self
        if len(ranks) == 1:  # a lone rank needs no process group
            return Gang(rank=0 if ranks[0] == self.rank else -1, world=1, group=None, device=self.device,
                        global_ranks=tuple(ranks))
        grp = dist.new_group(ranks)
        if self.rank not in ranks:
            return Gang(rank=-1, world=len(ranks), group=None, device=self.device, global_ranks=tuple(ranks))
        return Gang(rank=ranks.index(self.rank), world=len(ranks), group=grp, device=self.device,
                    global_ranks=tuple(ranks))


@dataclass
class SymTensor:
    """A tensor living at the same offset of every rank's symmetric buffer."""

    local: torch.Tensor
    offset: int            # bytes from the arena base
    arena: "SymmetricArena"

    def peer_ptr(self, rank: int) -> int:
        return self.arena.peer_bases[rank] + self.offset

    def peer_ptrs(self) -> List[int]:
        return [self.peer_ptr(r) for r in range(self.arena.world)]


class SymmetricArena:
    """Bump allocator over one symmetric-memory buffer + a flag area. All ranks must perform the
    same sequence of `alloc` calls."""

    def __init__(self, gang: Gang, nbytes: int):
        import torch.distributed._symmetric_memory as symm

        self.gang = gang
        self.world = gang.world
        self.rank = gang.rank
        total = nbytes + FLAG_SLOTS * 4
        self.buffer = symm.empty(total, dtype=torch.uint8, device=gang.device)
        self.buffer.zero_()
        self.handle = symm.rendezvous(self.buffer, gang.group)
        self.peer_bases = [int(p) for p in self.handle.buffer_ptrs]
        try:  # NVLS multicast mapping of the same buffer, when the fabric offers one
            self.multicast_ptr = int(self.handle.multicast_ptr or 0)
        except Exception:
            self.multicast_ptr = 0
        self._cursor = FLAG_SLOTS * 4
        self._flag_cursor = 0
        self.nbytes = total
        self.epoch = 0
        torch.cuda.synchronize()
        gang.barrier()

    # ---- tensors
    def alloc(self, *shape, dtype=torch.float16) -> SymTensor:
        numel = 1
        for s in shape:
            numel *= s
        size = numel * torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + 255) // 256 * 256
        if off + size > self.nbytes:
            raise MemoryError(f"symmetric arena exhausted ({off + size} > {self.nbytes})")
        self._cursor = off + size
        local = self.buffer[off:off + size].view(dtype).view(*shape)
        return SymTensor(local=local, offset=off, arena=self)

    # ---- flags: `count` consecutive int32 slots; slot [src] on rank d is written by rank src
    def alloc_flags(self, count: Optional[int] = None) -> SymTensor:
        count = count or self.world
        if self._flag_cursor + count > FLAG_SLOTS:
            raise MemoryError("flag area exhausted")
        off = self._flag_cursor * 4
        self._flag_cursor += count
        local = self.buffer[off:off + count * 4].view(torch.int32)
        return SymTensor(local=local, offset=off, arena=self)

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch

    def barrier(self):
        """Host-side full barrier of the gang (device-visible writes included)."""
        torch.cuda.synchronize()
        self.gang.barrier()

    def contains(self, t: torch.Tensor) -> bool:
        base = self.buffer.data_ptr()
        return base <= t.data_ptr() < base + self.nbytes

    def slices(self, count: int) -> List["ArenaSlice"]:
        """Cuts what is left of the arena (data and flags) into `count` equal, independently recyclable regions:
        one per cached plan, so that evicting a plan returns its peer-visible memory instead of leaking a whole
        arena (every rank must make the same call)."""
        start = (self._cursor + 4095) // 4096 * 4096
        per = (self.nbytes - start) // count // 4096 * 4096
        fper = (FLAG_SLOTS - self._flag_cursor) // count
        out = [ArenaSlice(self, start + i * per, per, self._flag_cursor + i * fper, fper) for i in range(count)]
        self._cursor = self.nbytes
        self._flag_cursor = FLAG_SLOTS
        return out


class ArenaSlice:
    """A recyclable region of a SymmetricArena with the allocator interface of the arena itself."""

    def __init__(self, parent: SymmetricArena, data_off: int, data_bytes: int, flag_off: int, flag_count: int):
        self.parent = parent
        self.gang, self.world, self.rank = parent.gang, parent.world, parent.rank
        self.peer_bases = parent.peer_bases
        self.multicast_ptr = parent.multicast_ptr
        self.buffer = parent.buffer
        self._lo, self._hi = data_off, data_off + data_bytes
        self._flag_lo, self._flag_hi = flag_off, flag_off + flag_count
        self._cursor, self._flag_cursor = self._lo, self._flag_lo
        self.in_use = False

    def reset(self):
        """Forget every allocation and clear the flags (their epochs restart from zero with the next plan); a gang
        barrier makes sure no peer still polls or writes them."""
        torch.cuda.synchronize()
        self.gang.barrier()
        self.parent.buffer[self._flag_lo * 4:self._flag_hi * 4].zero_()
        self._cursor, self._flag_cursor = self._lo, self._flag_lo
        torch.cuda.synchronize()
        self.gang.barrier()

    def alloc(self, *shape, dtype=torch.float16) -> SymTensor:
        numel = 1
        for s in shape:
            numel *= s
        size = numel * torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + 255) // 256 * 256
        if off + size > self._hi:
            raise MemoryError(f"symmetric arena slice exhausted ({off + size - self._lo} > {self._hi - self._lo})")
        self._cursor = off + size
        local = self.buffer[off:off + size].view(dtype).view(*shape)
        return SymTensor(local=local, offset=off, arena=self)

    def alloc_flags(self, count: Optional[int] = None) -> SymTensor:
        count = count or self.world
        if self._flag_cursor + count > self._flag_hi:
            raise MemoryError("flag area of the arena slice exhausted")
        off = self._flag_cursor * 4
        self._flag_cursor += count
        local = self.buffer[off:off + count * 4].view(torch.int32)
        return SymTensor(local=local, offset=off, arena=self)

    def contains(self, t: torch.Tensor) -> bool:
        base = self.buffer.data_ptr()
        return base + self._lo <= t.data_ptr() < base + self._hi
