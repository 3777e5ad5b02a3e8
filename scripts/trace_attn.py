"""Where the softmax warpgroups of k_flash_attn2 spend their clocks (probe build, -DSW_TRACE):
  SW_NATIVE_LIB=chiaswarm_b200/_lib/libswarm_sm100_trace.so python scripts/trace_attn.py
Per CTA and query tile, summed over the KV loop: wait for S, S -> registers, row max, wait for the previous
P.V, exponentials + P store, publish.  Medians over CTAs, in SM clocks per KV tile."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from chiaswarm_b200.ops import native as nat

lib = nat.lib()
lib.sw_trace_dump_attn.argtypes = [C.c_void_p, C.c_int]
lib.sw_trace_dump_attn.restype = C.c_int
for b, h, sq, skv, d in [(2, 8, 4096, 4096, 40), (2, 10, 4096, 4096, 64), (2, 20, 1024, 1024, 64), (2, 8, 1024, 1024, 80)]:
    q = torch.randn(b, sq, h, d, device="cuda", dtype=torch.float16)
    k = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
    v = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
    out = torch.empty(b, sq, h * d, device="cuda", dtype=torch.float16)
    for _ in range(3):
        nat.flash_attn(q, k, v, out, batch=b, heads=h, seq_q=sq, seq_kv=skv, head_dim=d)
    host = np.zeros(8192 * 16, dtype=np.uint64)
    assert lib.sw_trace_dump_attn(host.ctypes.data, 8192) == 0
    tr = host.reshape(8192, 16).astype(np.int64)
    gd = int(tr[0, 15])
    gx, gy, gz = gd & 0xFFFFF, (gd >> 20) & 0xFFFFF, gd >> 40
    nct = min(8192, gx * gy * gz)
    tr = tr[:nct]
    nkv = (skv + 127) // 128
    names = ["wait_S", "load_S", "row_max", "wait_PV", "exp+P", "publish"]
    for t in range(2):
        print(f"({b},{h},{sq},{skv},{d}) tile {t}: " + "  ".join(f"{nm} {np.median(tr[:, t * 6 + i]) / nkv:6.0f}" for i, nm in enumerate(names))
              + f"  | per KV tile {np.median(tr[:, t * 6:t * 6 + 6].sum(1)) / nkv:6.0f} clk, loop total {np.median(tr[:, 12]):8.0f} clk, grid {gx}x{gy}x{gz}",
              flush=True)
