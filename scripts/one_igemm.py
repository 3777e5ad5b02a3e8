"""One launch of the igemm kernel at an SD1.5 shape: the target of `ncu --set full` captures (profiles/)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.ops.packing import pack_conv_weight
n, h, w, cin, cout, k, tile = (int(a) for a in sys.argv[1:8])
x = torch.randn(n, h, w, cin, device="cuda", dtype=torch.float16)
wp = pack_conv_weight(torch.randn(cout, cin, k, k, device="cuda", dtype=torch.float16) * 0.02)
b = torch.randn(cout, device="cuda", dtype=torch.float16)
out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
scratch = torch.zeros((64 << 20) + 16384, device="cuda", dtype=torch.uint8)
junk = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
for _ in range(4):
    junk.zero_()  # evict the weights from L2: in the UNet every layer streams its weights from HBM
    nat.igemm(x, wp, out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=b, tile_n=tile, splitk_scratch=scratch)
torch.cuda.synchronize()
print("ok")
