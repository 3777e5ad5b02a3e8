"""sm_100a plan for UNet3DConditionModel (text-to-video; reference call
site swarm/video/tx2vid.py:55).

Frames ride in the batch dimension: activations are channels-last [B*F, H, W, C], so every spatial
layer is literally the 2-D plan (engine/unet_sm100.py).  The temporal layers reinterpret the same
memory without moving it:

  * TemporalConvLayer: Conv3d (3,1,1) == a 3x1 convolution over the tensor viewed as
    [n=B, h=F, w=H*W, c] -- the implicit-GEMM kernel with kh=3, kw=1, vertical padding 1 and
    horizontal padding 0; its GroupNorm runs over (F*H*W) per clip;
  * TransformerTemporalModel: attention over the frame axis for every pixel -- the flash-attention
    kernel with batch = H*W, sequence = F, expressed purely through strides of the fused QKV buffer.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..ops import native as nat
from ..ops.packing import pad_vector
from .layers import FM, GroupNorm, LayerNorm, Linear, Recorder
from .unet_sm100 import UNetPlan


def _frame_split(rec: Recorder) -> bool:
    """Is this plan one rank's share of a clip that is split along the frame axis (engine/patch_engine.py)?"""
    return rec.patch is not None and not rec.patch.rows and rec.patch.P > 1


class TemporalConv:
    def __init__(self, m):
        seqs = [m.conv1, m.conv2, m.conv3, m.conv4]
        self.norms = [GroupNorm(s[0]) for s in seqs]
        self.wgts, self.biases = [], []
        for s in seqs:
            conv = s[-1]
            w = conv.weight.detach().to(torch.float16)[:, :, :, 0, 0]           # [cout, cin, 3]
            self.wgts.append(w.permute(0, 2, 1).reshape(w.shape[0], -1).contiguous())  # tap-major K
            self.biases.append(conv.bias.detach().to(torch.float16).contiguous())

    def __call__(self, rec: Recorder, x: FM, nf: int) -> FM:
        b, hw = x.n // nf, x.h * x.w
        cur = FM(x.t, b, nf, hw, x.c)
        split = _frame_split(rec)
        for i in range(4):
            halo = {}
            if split:
                # frame-split clip: the GroupNorm spans every rank's frames, and the conv's taps f-1 / f+1 of the
                # first / last local frame live on the neighbours (halo frames pushed into the padded buffer)
                g = rec.halo_fm(nf, hw, x.c)
                self.norms[i](rec, cur, True, out=g, dist=True)
                a, halo = rec.patch.conv_input(rec, g)
            else:
                g = self.norms[i](rec, cur, True)
                a = g.t
            out = rec.fm(b, nf, hw, x.c)
            rec.add(nat.igemm, a, self.wgts[i], out.tokens(), n=b, h=nf, w=hw, kh=3, kw=1, pad=1, pad_w=0,
                    bias=self.biases[i], residual=x.tokens() if i == 3 else None, splitk_scratch=rec.splitk, **halo)
            rec.release(g)
            if i > 0:
                rec.release(cur)
            cur = out
        return FM(cur.t.view(x.n, x.h, x.w, x.c), x.n, x.h, x.w, x.c)


class TemporalBlock:
    """LN -> frame self-attention -> LN -> frame self-attention -> LN -> GEGLU FF (double self-attention)."""

    def __init__(self, m):
        self.heads, self.dh = m.attn1.heads, m.attn1.dim_head
        self.c = self.heads * self.dh
        self.lns = [LayerNorm(m.norm1), LayerNorm(m.norm2), LayerNorm(m.norm3)]
        self.qkv, self.out = [], []
        for a in (m.attn1, m.attn2):
            self.qkv.append(Linear(torch.cat([a.to_q.weight, a.to_k.weight, a.to_v.weight], 0)))
            self.out.append(Linear(a.to_out[0].weight, a.to_out[0].bias))
        self.ff1 = Linear(m.ff.net[0].proj.weight, m.ff.net[0].proj.bias, geglu=True)
        self.ff2 = Linear(m.ff.net[2].weight, m.ff.net[2].bias)

    def __call__(self, rec: Recorder, y: torch.Tensor, b: int, nf: int, hw: int) -> torch.Tensor:
        rows = b * nf * hw
        for k in range(2):
            xn = self.lns[k](rec, y)
            if _frame_split(rec):
                # this rank holds nf of the clip's P * nf frames: queries stay local, its K/V rows go into the
                # full-clip K/V buffer of every rank of the strip group (frame-major, so a rank's share is one
                # contiguous block), and attention runs over all frames for the local query frames
                ctx = rec.patch
                c, tot = self.c, ctx.P * nf
                q = rec.buf(rows, c)
                self.qkv[k].call_rows(rec, xn, 0, c, q)
                kvf = ctx.sym_buf(tot * hw, 2 * c)
                mine = kvf.local[ctx.p * rows:(ctx.p + 1) * rows]
                self.qkv[k].call_rows(rec, xn, c, 3 * c, mine)
                rec.release(xn)
                ctx.group_exchange(rec, mine, kvf, ctx.p * rows * 2 * c * 2)
                attn = rec.buf(rows, c)
                q4 = q.view(nf, hw, self.heads, self.dh).permute(1, 0, 2, 3)
                kv5 = kvf.local.view(tot, hw, 2, self.heads, self.dh).permute(1, 0, 2, 3, 4)
                rec.add(nat.flash_attn, q4, kv5[:, :, 0], kv5[:, :, 1], attn.view(nf, hw, c).permute(1, 0, 2), batch=hw,
                        heads=self.heads, seq_q=nf, seq_kv=tot, head_dim=self.dh)
                rec.release(q)
                y2 = self.out[k](rec, attn, residual=y)
                rec.release(attn)
                rec.release(y)
                y = y2
                continue
            qkv = self.qkv[k](rec, xn)
            rec.release(xn)
            attn = rec.buf(rows, self.c)
            q5 = qkv.view(b, nf, hw, 3, self.heads, self.dh)
            a4 = attn.view(b, nf, hw, self.c)
            for bi in range(b):
                v = q5[bi].permute(1, 0, 2, 3, 4)  # [hw, nf, 3, heads, d]: batch = pixel, sequence = frame
                rec.add(nat.flash_attn, v[:, :, 0], v[:, :, 1], v[:, :, 2], a4[bi].permute(1, 0, 2), batch=hw,
                        heads=self.heads, seq_q=nf, seq_kv=nf, head_dim=self.dh)
            rec.release(qkv)
            y2 = self.out[k](rec, attn, residual=y)
            rec.release(attn)
            rec.release(y)
            y = y2
        xn = self.lns[2](rec, y)
        ff = self.ff1(rec, xn)
        rec.release(xn)
        y2 = self.ff2(rec, ff, residual=y)
        rec.release(ff)
        rec.release(y)
        return y2


class TemporalTransformer:
    def __init__(self, m):
        self.norm = GroupNorm(m.norm)
        self.proj_in = Linear(m.proj_in.weight, m.proj_in.bias)
        self.proj_out = Linear(m.proj_out.weight, m.proj_out.bias)
        self.blocks = [TemporalBlock(b) for b in m.transformer_blocks]

    def __call__(self, rec: Recorder, x: FM, nf: int) -> FM:
        b, hw = x.n // nf, x.h * x.w
        g = self.norm(rec, FM(x.t, b, nf, hw, x.c), False, dist=_frame_split(rec))  # statistics over the clip
        y = self.proj_in(rec, g.tokens())
        rec.release(g)
        for blk in self.blocks:
            y = blk(rec, y, b, nf, hw)
        out = rec.fm(x.n, x.h, x.w, x.c)
        self.proj_out(rec, y, residual=x.tokens(), out=out.tokens())
        rec.release(y)
        return out


class UNet3DPlan(UNetPlan):
    def __init__(self, unet, device):
        super().__init__(unet, device)
        self.num_frames = 1
        self.transformer_in = TemporalTransformer(unet.transformer_in)
        mods = list(unet.down_blocks) + [unet.mid_block] + list(unet.up_blocks)
        for d, m in zip(self.down + [self.mid] + self.up, mods):
            d["tconvs"] = [TemporalConv(t) for t in m.temp_convs]
            d["tattns"] = [TemporalTransformer(t) for t in m.temp_attentions] if m.temp_attentions is not None else None

    def _res(self, rec, d, i, x, skip, temb_all, release_in):
        nf = self.num_frames
        r = d["resnets"][i](rec, x, skip, temb_all)
        if release_in:
            rec.release(x)
        if skip is not None:
            rec.release(skip)
        t = d["tconvs"][i](rec, r, nf)
        rec.release(r)
        return t

    def _attn(self, rec, d, i, x):
        if d["attns"] is None:
            return x
        a = d["attns"][i](rec, x)
        rec.release(x)
        t = d["tattns"][i](rec, a, self.num_frames)
        rec.release(a)
        return t

    def record_encoder(self, rec: Recorder, x_in: FM, temb_all, cond_feat: Optional[FM] = None):
        c = self.conv_in(rec, x_in)
        x = self.transformer_in(rec, c, self.num_frames)
        rec.release(c)
        skips = [x]
        for d in self.down:
            for i in range(len(d["resnets"])):
                x = self._attn(rec, d, i, self._res(rec, d, i, x, None, temb_all, release_in=False))
                skips.append(x)
            if d["down"] is not None:
                x = d["down"](rec, x)
                skips.append(x)
        x = self._res(rec, self.mid, 0, x, None, temb_all, release_in=False)
        x = self._attn(rec, self.mid, 0, x)
        x = self._res(rec, self.mid, 1, x, None, temb_all, release_in=True)
        return x, skips

    def record_decoder(self, rec: Recorder, x: FM, skips: List[FM], temb_all, out: Optional[FM] = None) -> FM:
        skips = list(skips)
        for d in self.up:
            for i in range(len(d["resnets"])):
                x = self._attn(rec, d, i, self._res(rec, d, i, x, skips.pop(), temb_all, release_in=True))
            if d["up"] is not None:
                x2 = d["up"](rec, x)
                rec.release(x)
                x = x2
        g = self.norm_out(rec, x, True)
        rec.release(x)
        out = self.conv_out(rec, g, out=out)
        rec.release(g)
        return out
