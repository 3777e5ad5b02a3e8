"""K/V-gather attention over shapes the patch engine produces (short shards, many heads, batch 1) vs SDPA."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from chiaswarm_b200.parallel import runtime
from chiaswarm_b200.parallel.collectives import FusedOps
gang = runtime.init()
dev, W, R = gang.device, gang.world, gang.rank
arena = runtime.SymmetricArena(gang, 1 << 30)
ops = FusedOps(arena)
torch.manual_seed(1)
worst = 0.0
for (b, h, d, s_loc) in [(1, 20, 64, 256), (1, 10, 64, 1024), (1, 20, 64, 128), (2, 20, 64, 256), (1, 8, 80, 256), (1, 8, 40, 1024),
                         (1, 20, 64, 512), (1, 5, 64, 256)]:
    S = s_loc * W
    q = torch.randn(b, S, h, d, device=dev, dtype=torch.float16)
    k = torch.randn(b, S, h, d, device=dev, dtype=torch.float16)
    v = torch.randn(b, S, h, d, device=dev, dtype=torch.float16)
    for t in (q, k, v):
        torch.distributed.broadcast(t, 0)
    # fused QKV layout as the engine has it: [b, s_loc, 3, h, d] in symmetric memory
    qkv = arena.alloc(b * s_loc, 3 * h * d)
    v5 = qkv.local.view(b, s_loc, 3, h, d)
    sl = slice(R * s_loc, (R + 1) * s_loc)
    v5[:, :, 0].copy_(q[:, sl]); v5[:, :, 1].copy_(k[:, sl]); v5[:, :, 2].copy_(v[:, sl])
    out = torch.zeros(b, s_loc, h * d, device=dev, dtype=torch.float16)
    from chiaswarm_b200.ops import native as nat
    c = h * d
    errs = []
    for rep in range(3):
        e = arena.next_epoch()
        ranks = list(range(W))
        kp = nat.make_peers(W, R, e, [qkv.peer_ptr(r) + 2 * c for r in ranks], ops.flags["kv"].peer_ptrs())
        vptrs = [qkv.peer_ptr(r) + 4 * c for r in ranks]
        torch.cuda.synchronize(); gang.barrier()
        nat.peer_signal(kp)
        nat.flash_attn_gather(v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], out, kp, vptrs, batch=b, heads=h, seq_q=s_loc, s_local=s_loc, head_dim=d)
        torch.cuda.synchronize(); gang.barrier()
        ref = F.scaled_dot_product_attention(q[:, sl].float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2))
        ref = ref.transpose(1, 2).reshape(b, s_loc, c)
        errs.append(((out.float() - ref).abs().max() / ref.abs().max()).item())
    err = gang.all_reduce_max(max(errs))
    worst = max(worst, err)
    if R == 0:
        print(f"b={b} heads={h} d={d} s_local={s_loc} world={W}: rel_err {err:.3e}", flush=True)
if R == 0:
    print("GATHER_PROBE", "PASS" if worst < 5e-3 else "FAIL")
# distributed GroupNorm at the local shapes of SDXL strips (hw_local, c1, c2)
worst = 0.0
for (hw, c1, c2) in [(4096, 320, 0), (4096, 320, 320), (4096, 640, 320), (1024, 640, 0), (1024, 640, 640), (1024, 1280, 640),
                     (1024, 320, 0), (256, 1280, 0), (256, 1280, 1280), (256, 640, 0), (1024, 640, 320), (4096, 640, 0)]:
    c = c1 + c2
    xf = torch.randn(1, W * hw, c, device=dev, dtype=torch.float16) * 1.5 + 0.2
    gam = torch.randn(c, device=dev, dtype=torch.float16)
    bet = torch.randn(c, device=dev, dtype=torch.float16)
    for t in (xf, gam, bet):
        torch.distributed.broadcast(t, 0)
    xl = xf[:, R * hw:(R + 1) * hw]
    x1 = xl[..., :c1].contiguous()
    x2 = xl[..., c1:].contiguous() if c2 else None
    out = torch.zeros(1, hw, c, device=dev, dtype=torch.float16)
    errs = []
    for rep in range(3):
        torch.cuda.synchronize(); gang.barrier()
        ops.groupnorm_fused(x1, gam, bet, out, n=1, hw_local=hw, groups=32, eps=1e-5, silu=True, x2=x2)
        torch.cuda.synchronize(); gang.barrier()
        ref = F.silu(F.group_norm(xf.float().permute(0, 2, 1), 32, gam.float(), bet.float(), 1e-5)).permute(0, 2, 1)[:, R * hw:(R + 1) * hw]
        errs.append((out.float() - ref).abs().max().item())
    err = gang.all_reduce_max(max(errs))
    worst = max(worst, err)
    if R == 0:
        print(f"gn hw_local={hw} c1={c1} c2={c2} world={W}: abs_err {err:.3e}", flush=True)
if R == 0:
    print("GN_PROBE", "PASS" if worst < 3e-2 else "FAIL")
torch.distributed.destroy_process_group()
