"""sw_flash_attn vs torch SDPA on the SD1.5 / SDXL attention shapes (CUDA events, warm)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


shapes = [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160), (2, 8, 64, 64, 160),
          (2, 8, 4096, 77, 40), (2, 8, 1024, 77, 80), (2, 10, 4096, 4096, 64), (2, 20, 1024, 1024, 64),
          (1, 1, 4096, 4096, 64)]
for b, h, sq, skv, d in shapes:
    q = torch.randn(b, sq, h, d, device="cuda", dtype=torch.float16)
    k = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
    v = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
    out = torch.empty(b, sq, h * d, device="cuda", dtype=torch.float16)
    qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
    t_ref = timeit(lambda: F.scaled_dot_product_attention(qt, kt, vt))
    t_our = timeit(lambda: nat.flash_attn(q, k, v, out, batch=b, heads=h, seq_q=sq, seq_kv=skv, head_dim=d))
    gf = 4.0 * b * h * sq * skv * d / 1e9
    print(f"{str((b,h,sq,skv,d)):28s} {gf:7.2f} GF  sdpa {t_ref:7.1f} us ({gf/t_ref*1e3:5.0f} TF)  ours {t_our:7.1f} us ({gf/t_our*1e3:5.0f} TF)")
