"""Resident pipeline / weight cache.  The reference reloads every model
from disk for every job (swarm/diffusion/diffusion_func.py:41-46); on a 180 GB B200 the packed
weights of every model a worker serves stay on the device, keyed by everything that changes the
weights, with least-recently-used eviction by bytes."""
from __future__ import annotations

import threading
from collections import OrderedDict
from typing import Callable, Hashable


def _nbytes(obj) -> int:
    """Bytes of the parameters and buffers an entry keeps resident: the entry itself when it is a module, the modules in
    a tuple / list / dict, or the modules held one attribute deep by a pipeline object (unet, vae, text encoders, vocoder,
    ...).  Shared tensors count once."""
    seen, total = set(), 0

    def add(module):
        nonlocal total
        for t in list(module.parameters()) + list(module.buffers()):
            if id(t) not in seen:
                seen.add(id(t))
                total += t.numel() * t.element_size()

    def visit(x, depth):
        if hasattr(x, "parameters") and hasattr(x, "buffers") and callable(x.parameters):
            add(x)
        elif isinstance(x, (tuple, list)):
            for v in x:
                visit(v, depth)
        elif isinstance(x, dict):
            for v in x.values():
                visit(v, depth)
        elif depth > 0 and hasattr(x, "__dict__"):
            for v in vars(x).values():
                visit(v, depth - 1)

    visit(obj, 1)
    return total


def _device_free_fraction():
    """Free / total memory of the current CUDA device; None without one."""
    try:
        import torch

        if not torch.cuda.is_available():
            return None
        free, total = torch.cuda.mem_get_info()
        return free / max(total, 1)
    except Exception:
        return None


class LRUByBytes:
    """`budget_bytes` bounds what the entries' modules occupy; `min_free` (fraction of device memory) bounds what this
    accounting cannot see -- the engines' packed weight copies, captured graphs and work buffers, other processes: after
    an insertion, least-recently-used entries are dropped (and the allocator's cache released) until that much of the
    device is free again."""

    def __init__(self, budget_bytes: int, min_free: float = 0.08, free_fraction: Callable = _device_free_fraction):
        self.budget = budget_bytes
        self.min_free, self._free_fraction = min_free, free_fraction
        self._items: "OrderedDict[Hashable, tuple]" = OrderedDict()
        self._lock = threading.Lock()
        self.hits = 0
        self.misses = 0
        self.evictions = 0

    def get(self, key: Hashable, loader: Callable):
        with self._lock:
            if key in self._items:
                self._items.move_to_end(key)
                self.hits += 1
                return self._items[key][0]
        value = loader()
        size = _nbytes(value)
        with self._lock:
            self.misses += 1
            self._items[key] = (value, size)
            self._items.move_to_end(key)
            while len(self._items) > 1 and sum(s for _, s in self._items.values()) > self.budget:
                self._items.popitem(last=False)
                self.evictions += 1
            self._relieve_pressure()
        return value

    def _relieve_pressure(self):
        """Called with the lock held, after an insertion."""
        frac = self._free_fraction()
        while frac is not None and frac < self.min_free and len(self._items) > 1:
            self._items.popitem(last=False)
            self.evictions += 1
            self._release_device_memory()
            frac = self._free_fraction()

    @staticmethod
    def _release_device_memory():
        import gc

        gc.collect()
        try:
            import torch

            if torch.cuda.is_available():
                torch.cuda.empty_cache()
        except Exception:
            pass

    def clear(self):
        with self._lock:
            self._items.clear()

    def __len__(self):
        return len(self._items)

    def bytes_resident(self) -> int:
        with self._lock:
            return sum(s for _, s in self._items.values())


_cache = None


def pipeline_cache() -> LRUByBytes:
    global _cache
    if _cache is None:
        from ..settings import load_settings

        try:
            gb = float(getattr(load_settings(), "weight_cache_gb", 120))
        except Exception:
            gb = 120.0
        _cache = LRUByBytes(int(gb * (1 << 30)))
    return _cache
