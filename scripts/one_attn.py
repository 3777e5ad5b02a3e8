"""One launch of the attn kernel at an SD1.5 shape: the target of `ncu --set full` captures (profiles/)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat
b, h, sq, skv, d = (int(a) for a in sys.argv[1:6])
q = torch.randn(b, sq, h, d, device="cuda", dtype=torch.float16)
k = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
v = torch.randn(b, skv, h, d, device="cuda", dtype=torch.float16)
out = torch.empty(b, sq, h * d, device="cuda", dtype=torch.float16)
for _ in range(4):
    nat.flash_attn(q, k, v, out, batch=b, heads=h, seq_q=sq, seq_kv=skv, head_dim=d)
torch.cuda.synchronize()
print("ok")
