"""Temporal attention (sequence = frames, batch = pixels): packed tiles vs one batch entry per tile."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
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

for f, hw, heads, d in [(24, 2880, 5, 64), (24, 720, 10, 64), (24, 180, 20, 64), (16, 4096, 8, 40)]:
    qkv = torch.randn(f, hw, 3, heads, d, device="cuda", dtype=torch.float16)
    out = torch.zeros(f, hw, heads * d, device="cuda", dtype=torch.float16)
    v = qkv.permute(1, 0, 2, 3, 4)
    t = timeit(lambda: nat.flash_attn(v[:, :, 0], v[:, :, 1], v[:, :, 2], out.permute(1, 0, 2), batch=hw, heads=heads,
                                      seq_q=f, seq_kv=f, head_dim=d))
    print(f"frames={f} pixels={hw} heads={heads} d={d}: {t:8.1f} us  (SW_ATTN_PACK={os.environ.get('SW_ATTN_PACK', '1')})")
