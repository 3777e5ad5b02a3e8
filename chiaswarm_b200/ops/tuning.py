"""Measured tile configurations for sw_igemm.

`scripts/tune_igemm.py` times every distinct implicit-GEMM shape of the supported pipelines on a B200
(CUDA-graph replays of the real plan ops, device-timed) over the kernel's tile width / CTA-pair /
split-K choices and stores the winners in `igemm_tuning.json`.  `ops.native.igemm` consults the table
whenever the caller left all three knobs on "auto"; shapes that are not in the table fall back to the
heuristics inside the kernel launcher (csrc/igemm.cu).  The table only ever selects between code paths
of the same kernel: a missing or stale file costs speed, never correctness."""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import Dict, Optional, Tuple

_PATH = Path(__file__).with_name("igemm_tuning.json")
_table: Optional[Dict[str, Tuple[int, ...]]] = None


def key_of(n, h, w, kh, kw, c1, c2, cout, act, residual) -> str:
    return f"{n},{h},{w},{kh},{kw},{c1},{c2},{cout},{act},{int(bool(residual))}"


def table() -> Dict[str, Tuple[int, ...]]:
    global _table
    if _table is None:
        _table = {}
        if os.environ.get("SW_IGEMM_TUNING", "1") not in ("0", "") and _PATH.is_file():
            try:
                raw = json.loads(_PATH.read_text())
                _table = {k: (int(v[0]), int(v[1]), int(v[2]), int(v[3]) if len(v) > 3 else 0, int(v[4]) if len(v) > 4 else 0)
                          for k, v in raw.get("configs", {}).items()}
            except (ValueError, OSError, TypeError, IndexError):
                _table = {}
    return _table


def lookup(n, h, w, kh, kw, c1, c2, cout, act, residual) -> Optional[Tuple[int, ...]]:
    """(tile_n, cta_pair, splitk_min_kb, persist, cluster_k) or None."""
    return table().get(key_of(n, h, w, kh, kw, c1, c2, cout, act, residual))
