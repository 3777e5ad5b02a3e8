"""Timing bisection of the fused GEMM + reduce-scatter kernel on ONE GPU (every "peer" is this GPU, epoch 0):
SW_FUSED_DBG=0/1/2/3 switches off the completion wait + flag (1) and the reduce role (2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat

dev = "cuda:0"
W, m_local, ks, n2 = 2, 4096, 1280, 640
x = torch.randn(W * m_local, ks, device=dev, dtype=torch.float16)
w = torch.randn(n2, ks, device=dev, dtype=torch.float16) * ks ** -0.5
staging = torch.zeros(W, m_local, n2, device=dev, dtype=torch.float16)
flags = torch.zeros(4096, dtype=torch.int32, device=dev)
mode = os.environ.get("RS_PROBE_SYMM", "0")
if mode != "0":  # staging (1), flags (2) or both (3) in symmetric memory, as the multi-GPU runs have them
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29599")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device(dev))
    buf = symm.empty(staging.numel() * 2 + 16384, dtype=torch.uint8, device=dev)
    buf.zero_()
    h = symm.rendezvous(buf, dist.group.WORLD)
    if int(mode) & 1:
        staging = buf[16384:].view(torch.float16).view(W, m_local, n2)
    if int(mode) & 2:
        flags = buf[:16384].view(torch.int32)
out = torch.empty(m_local, n2, device=dev, dtype=torch.float16)
part = torch.empty(W * m_local, n2, device=dev, dtype=torch.float16)
pf = nat.make_peers(W, 0, 0, [staging.data_ptr()] * W, [flags.data_ptr()] * W)


def timed(fn, iters=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


# all-gather GEMM, same local-only arrangement (every "peer" is this GPU's gathered buffer, epoch 0)
k_ag, n_ag = 640, 2560
gathered = torch.randn(W * m_local, k_ag, device=dev, dtype=torch.float16)
w_ag = torch.randn(n_ag, k_ag, device=dev, dtype=torch.float16) * k_ag ** -0.5
out_ag = torch.empty(W * m_local, n_ag, device=dev, dtype=torch.float16)
flags_ag = torch.zeros(4096, dtype=torch.int32, device=dev)
pf_ag = nat.make_peers(W, 0, 0, [gathered.data_ptr()] * W, [flags_ag.data_ptr()] * W)
t_ag = timed(lambda: nat.igemm_allgather(w_ag, out_ag, pf_ag, m_local=m_local, k=k_ag))
print("fused_ag_local %.1f us" % t_ag, "| plain gemm %.1f us" % timed(lambda: nat.linear(gathered, w_ag, out=out_ag)))
print("SYMM", mode, "SW_FUSED_DBG", os.environ.get("SW_FUSED_DBG", "0"), "SW_PDL", os.environ.get("SW_PDL", "1"),
      "fused_rs_local %.1f us" % timed(lambda: nat.igemm_reducescatter(x, w, pf, out, m_local=m_local, ldc=n2)),
      "| plain gemm %.1f us" % timed(lambda: nat.linear(x, w, out=part)),
      "| plain gemm tile128 nopair %.1f us" % timed(lambda: nat.linear(x, w, out=part, tile_n=128, cta_pair=1)))
