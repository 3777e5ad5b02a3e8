"""Which library kernels (cuBLAS / SDPA) torch picks for the UNet's linear and attention shapes.
Run under `ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,...` to list
kernel names (nvjet tile / stage / cluster encodings), grids and device times: the comparison
point for sw_igemm / sw_flash_attn tile choices."""
import torch
import torch.nn.functional as F

dev = "cuda"
shapes = [(8192, 320, 320), (8192, 960, 320), (8192, 2560, 320), (8192, 320, 1280),
          (2048, 640, 640), (2048, 1920, 640), (2048, 5120, 640), (2048, 640, 2560),
          (512, 1280, 1280), (512, 3840, 1280), (512, 10240, 1280), (512, 1280, 5120), (128, 1280, 1280),
          (2048, 1280, 5120), (128, 1280, 11520), (512, 1280, 11520)]
for it in range(2):
    for (m, n, k) in shapes:
        x = torch.randn(m, k, device=dev, dtype=torch.float16)
        w = torch.randn(n, k, device=dev, dtype=torch.float16)
        b = torch.randn(n, device=dev, dtype=torch.float16)
        torch.cuda.nvtx.range_push(f"lin_{m}_{n}_{k}")
        F.linear(x, w, b)
        torch.cuda.nvtx.range_pop()
    for (bsz, h, s, skv, d) in [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 4096, 77, 40), (2, 10, 4096, 4096, 64),
                                (2, 20, 1024, 1024, 64)]:
        q = torch.randn(bsz, h, s, d, device=dev, dtype=torch.float16)
        kk = torch.randn(bsz, h, skv, d, device=dev, dtype=torch.float16)
        v = torch.randn(bsz, h, skv, d, device=dev, dtype=torch.float16)
        F.scaled_dot_product_attention(q, kk, v)
torch.cuda.synchronize()
print("ok")
