"""One launch of the gn kernel at an SD1.5 shape: the target of `ncu --set full` captures (profiles/)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat
stats = torch.zeros(1 << 20, device="cuda", dtype=torch.float32)
for n, hw, c in [(2, 64, 1280), (2, 4096, 320), (2, 1024, 640)]:
    x = torch.randn(n, hw, c, device="cuda", dtype=torch.float16)
    g = torch.randn(c, device="cuda", dtype=torch.float16)
    b = torch.randn(c, device="cuda", dtype=torch.float16)
    out = torch.empty_like(x)
    for _ in range(3):
        nat.groupnorm(x, g, b, out, n=n, hw=hw, groups=32, eps=1e-5, silu=True, stats=stats, scratch_zeroed=True)
torch.cuda.synchronize()
print("ok")
