"""One transformer block of a row-split plan, head-parallel (SW_PATCH_ATTN=tp: all-gather fused into the QKV GEMM,
out-projection GEMM fused with its reduce-scatter) against the same block on one GPU and against the eager fp32
module.  Run under torchrun with 2 or 4 ranks:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/tp_block_check.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SW_PATCH_ATTN"] = "tp"
import torch

from chiaswarm_b200.diffusion.pipelines import _init_module
from chiaswarm_b200.engine.layers import Recorder, TransformerBlock
from chiaswarm_b200.engine.patch_engine import PatchContext, PatchRecorder
from chiaswarm_b200.models.unet2d import BasicTransformerBlock
from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.parallel import runtime

gang = runtime.init()
dev, P, p = gang.device, gang.world, gang.rank
arena = runtime.SymmetricArena(gang, 1 << 30)
results, ok = {}, True
for (c, heads, dh, S, reps) in [(320, 8, 40, 4096, 1), (640, 8, 80, 1024, 1), (640, 10, 64, 4096, 3), (1280, 20, 64, 1024, 3)]:
    if heads % P:
        continue
    torch.manual_seed(5)
    m = _init_module(BasicTransformerBlock(c, heads, dh, 768), 11).to(dev, torch.float16).eval()
    x = (torch.randn(S, c, device=dev) * 0.7).to(torch.float16)
    ctx_txt = torch.randn(1, 77, 768, device=dev, dtype=torch.float16)
    torch.distributed.broadcast(x, 0)
    torch.distributed.broadcast(ctx_txt, 0)
    with torch.no_grad():
        ref = m.float()(x.float()[None], ctx_txt.float())[0]
    m.half()
    blk = TransformerBlock(m)
    for kv in blk.cross_layers():
        kv.kv = torch.empty(1, 77, kv.lin.nout, dtype=torch.float16, device=dev)
        nat.igemm(ctx_txt.view(77, 768), kv.lin.wgt, kv.kv.view(77, -1), n=1, h=1, w=77)
    # single GPU plan
    rec = Recorder(dev)
    y = rec.buf(S, c)
    out1 = blk(rec, y, 1, S)
    y.copy_(x)
    rec.run()
    torch.cuda.synchronize()
    one = out1.clone()
    # head-parallel over the gang: this rank holds rows [p*s, (p+1)*s)
    s = S // P
    pctx = PatchContext(gang, 1, arena)
    prec = PatchRecorder(dev, pctx)
    yl = prec.buf(s, c)
    assert blk.tp_eligible(prec, s), "shape not eligible for the head-parallel path"
    outp = blk(prec, yl, 1, s)
    prec.add(nat.advance_counter, pctx.tick)
    pctx.finalize()
    # stage by stage on the first input: ops 0..3 = LayerNorm, all-gather GEMM, attention, GEMM reduce-scatter
    gathered, qkv_b, attn_b, y1_b, wqkv, wo = blk._tp_debug
    yl.copy_(x[p * s:(p + 1) * s])
    torch.cuda.synchronize()
    gang.barrier()
    stage = {}
    for i, (fn, a, kw) in enumerate(prec.ops[:4]):
        fn(*a, **kw)
        torch.cuda.synchronize()
        gang.barrier()
        if i == 0:
            ln_full = torch.nn.functional.layer_norm(x.float(), (c,), m.norm1.weight.float(), m.norm1.bias.float(), m.norm1.eps)
            stage["ln_rows"] = (gathered.local[p * s:(p + 1) * s].float() - ln_full[p * s:(p + 1) * s]).abs().max().item()
        if i == 1:
            stage["gathered"] = (gathered.local.float() - ln_full).abs().max().item()
            want = gathered.local.float() @ wqkv.float().t()
            stage["qkv"] = ((qkv_b.float() - want).abs().max() / want.abs().max()).item()
        if i == 2:
            hp = heads // P
            q5 = qkv_b.float().view(1, S, 3, hp, dh)
            want_a = torch.nn.functional.scaled_dot_product_attention(q5[:, :, 0].transpose(1, 2), q5[:, :, 1].transpose(1, 2),
                                                                      q5[:, :, 2].transpose(1, 2)).transpose(1, 2).reshape(S, hp * dh)
            stage["attn"] = ((attn_b.float() - want_a).abs().max() / want_a.abs().max()).item()
            part = attn_b.float() @ wo.float().t()
            torch.distributed.all_reduce(part)
            want_y1 = part[p * s:(p + 1) * s] + m.attn1.to_out[0].bias.float() + x.float()[p * s:(p + 1) * s]
        if i == 3:
            stage["y1"] = ((y1_b.float() - want_y1).abs().max() / want_y1.abs().max()).item()
    stage = {k: gang.all_reduce_max(v) for k, v in stage.items()}
    for fn, a, kw in prec.ops[4:]:
        fn(*a, **kw)
    torch.cuda.synchronize()
    gang.barrier()
    errs = []
    for r in range(reps):  # replays with changing inputs: epochs advance, buffers are reused
        xr = (x.float() * (1.0 + 0.25 * r)).to(torch.float16)
        yl.copy_(xr[p * s:(p + 1) * s])
        torch.cuda.synchronize()
        gang.barrier()
        prec.run()
        torch.cuda.synchronize()
        y.copy_(xr)
        rec.run()
        torch.cuda.synchronize()
        mine = out1[p * s:(p + 1) * s].float()
        errs.append(((outp.float() - mine).abs().max() / mine.abs().max()).item())
    e_ref = ((one.float() - ref).abs().max() / ref.abs().max()).item()
    err = gang.all_reduce_max(max(errs))
    results[f"c{c}_h{heads}_d{dh}_S{S}"] = {"tp_vs_single_gpu_plan": err, "single_gpu_plan_vs_fp32_module": e_ref,
                                            "launches_tp": len(prec), "launches_single": len(rec), "stages": stage}
    ok = ok and err < 5e-3
gang.barrier()
if p == 0:
    print(json.dumps(results, indent=1))
    print("TP_BLOCK_CHECK", "PASS" if ok else "FAIL", flush=True)
torch.distributed.destroy_process_group()
