"""One launch of the geglu kernel at an SD1.5 shape: the target of `ncu --set full` captures (profiles/)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat
m, n, k = 8192, 2560, 320
x = torch.randn(m, k, device="cuda", dtype=torch.float16)
w = torch.randn(n, k, device="cuda", dtype=torch.float16) * k ** -0.5
b = torch.randn(n, device="cuda", dtype=torch.float16)
out = torch.empty(m, n // 2, device="cuda", dtype=torch.float16)
for _ in range(4):
    nat.igemm(x, w, out, n=1, h=1, w=m, bias=b, act=nat.ACT_GEGLU, tile_n=128)
torch.cuda.synchronize()
print("ok")
