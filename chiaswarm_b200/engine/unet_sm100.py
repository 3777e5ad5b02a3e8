"""sm_100a execution plans for UNet2DConditionModel / ControlNetModel / AutoencoderKL decoder.

Each `build_*` walks the torch module (the specification) and records a static plan on a
Recorder (engine/layers.py).  Per denoising step the UNet plan is:

  timestep embedding (reads t_table[*step]) -> time MLP -> ONE fused GEMM for every resnet's
  time_emb_proj -> conv_in -> down blocks -> mid -> up blocks (skip tensors consumed as second
  GroupNorm / shortcut inputs, never concatenated) -> GN+SiLU -> conv_out

followed (engine/denoiser.py) by the fused CFG + scheduler-update kernel.  Reference call site
being replaced: `eps = unet(x_in, t, text, [residuals])` inside diffusers' denoising loop
(swarm/diffusion/diffusion_func.py:96).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..ops import native as nat
from .layers import FM, Conv, GroupNorm, Linear, Recorder, Resnet, Transformer2D, Upsample, _r8


class UNetPlan:
    def __init__(self, unet, device):
        self.cfg = unet.config
        self.device = torch.device(device)
        self.in_ch = self.cfg["in_channels"]
        self.in_pad = _r8(self.in_ch)
        self.out_ch = self.cfg["out_channels"]
        self.boc = list(self.cfg["block_out_channels"])
        self.temb_in = self.boc[0]

        # ---- time MLP + fused time_emb_proj of every resnet
        te = unet.time_embedding
        self.t1 = Linear(te.linear_1.weight, te.linear_1.bias)
        self.t2 = Linear(te.linear_2.weight, te.linear_2.bias)
        resnets = []
        for blk in list(unet.down_blocks) + [unet.mid_block] + list(getattr(unet, "up_blocks", [])):
            resnets.extend(blk.resnets)
        off = 0
        ws, bs = [], []
        self._slices = {}
        for r in resnets:
            if r.time_emb_proj is not None:
                width = r.time_emb_proj.weight.shape[0]
                self._slices[id(r)] = (off, width)
                ws.append(r.time_emb_proj.weight)
                bs.append(r.time_emb_proj.bias)
                off += width
        self.temb_proj = Linear(torch.cat(ws, 0), torch.cat(bs, 0)) if ws else None

        def res(m):
            return Resnet(m, self._slices.get(id(m)))

        self.conv_in = Conv(unet.conv_in, cin_pad=self.in_pad)
        self.down = []
        for blk in unet.down_blocks:
            self.down.append(dict(
                resnets=[res(r) for r in blk.resnets],
                attns=[Transformer2D(a) for a in blk.attentions] if blk.attentions is not None else None,
                down=Conv(blk.downsamplers[0].conv) if blk.downsamplers is not None else None))
        mb = unet.mid_block
        self.mid = dict(resnets=[res(r) for r in mb.resnets],
                        attns=[Transformer2D(a) for a in mb.attentions] if mb.attentions is not None else None)
        self.up = []
        if hasattr(unet, "up_blocks"):
            for blk in unet.up_blocks:
                self.up.append(dict(
                    resnets=[res(r) for r in blk.resnets],
                    attns=[Transformer2D(a) for a in blk.attentions] if blk.attentions is not None else None,
                    up=Upsample(blk.upsamplers[0]) if blk.upsamplers is not None else None))
            self.norm_out = GroupNorm(unet.conv_norm_out)
            self.conv_out = Conv(unet.conv_out)

    # ------------------------------------------------------------------ cross-attention K/V
    def cross_layers(self):
        out = []
        for blk in self.down + [self.mid] + self.up:
            for a in (blk["attns"] or []):
                out.extend(a.cross_layers())
        return out

    def set_context(self, embeds: torch.Tensor):
        """embeds [n, s_text, d] fp16: computes every cross-attention layer's K/V once per job
        (they do not depend on the timestep)."""
        n, s, d = embeds.shape
        flat = embeds.reshape(n * s, d).contiguous()
        for kv in self.cross_layers():
            buf = kv.buffers.get((n, s))
            if buf is None:
                buf = kv.buffers[(n, s)] = torch.empty(n, s, kv.lin.nout, dtype=torch.float16, device=embeds.device)
            kv.kv = buf
            nat.igemm(flat, kv.lin.wgt, buf.view(n * s, -1), n=1, h=1, w=n * s)

    # ------------------------------------------------------------------ plan
    def record_time_embedding(self, rec: Recorder, n: int, t_table, step_idx, aug_emb=None):
        """Returns the fused per-resnet time projections [n, sum(cout)]."""
        sin = rec.buf(n, self.temb_in)
        rec.add(nat.timestep_embedding, t_table, step_idx, sin, n=n, dim=self.temb_in,
                flip_sin_to_cos=self.cfg.get("flip_sin_to_cos", True), freq_shift=float(self.cfg.get("freq_shift", 0)))
        h = self.t1(rec, sin, act=nat.ACT_SILU)
        # emb = linear_2(h) [+ added-condition embedding]; every consumer applies SiLU first
        if aug_emb is None:
            act_emb = self.t2(rec, h, act=nat.ACT_SILU)
        else:
            emb = self.t2(rec, h, residual=aug_emb)
            act_emb = rec.buf(*emb.shape)
            rec.add(nat.silu, emb, act_emb)
        if self.temb_proj is None:
            return None
        return self.temb_proj(rec, act_emb)

    def record_encoder(self, rec: Recorder, x_in: FM, temb_all, cond_feat: Optional[FM] = None):
        """conv_in + down blocks + mid. Returns (mid output, skip list)."""
        x = self.conv_in(rec, x_in, residual=cond_feat)
        skips = [x]
        for blk in self.down:
            for i, r in enumerate(blk["resnets"]):
                x = r(rec, x, None, temb_all)
                if blk["attns"] is not None:
                    x2 = blk["attns"][i](rec, x)
                    rec.release(x)
                    x = x2
                skips.append(x)
            if blk["down"] is not None:
                x = blk["down"](rec, x)
                skips.append(x)
        x = self.mid["resnets"][0](rec, x, None, temb_all)
        if self.mid["attns"] is not None:
            x2 = self.mid["attns"][0](rec, x)
            rec.release(x)
            x = x2
        x2 = self.mid["resnets"][1](rec, x, None, temb_all)
        rec.release(x)
        return x2, skips

    def record_decoder(self, rec: Recorder, x: FM, skips: List[FM], temb_all, out: Optional[FM] = None) -> FM:
        skips = list(skips)
        for blk in self.up:
            for i, r in enumerate(blk["resnets"]):
                skip = skips.pop()
                x2 = r(rec, x, skip, temb_all)
                rec.release(x)
                rec.release(skip)
                x = x2
                if blk["attns"] is not None:
                    x2 = blk["attns"][i](rec, x)
                    rec.release(x)
                    x = x2
            if blk["up"] is not None:
                x2 = blk["up"](rec, x)
                rec.release(x)
                x = x2
        g = self.norm_out(rec, x, True)
        rec.release(x)
        out = self.conv_out(rec, g, out=out)
        rec.release(g)
        return out


class ControlNetPlan(UNetPlan):
    """Encoder copy + conditioning embedder + zero convs.  The zero-conv outputs are not stored
    separately: each zero conv writes `unet_skip + scale * zero_conv(feature)` straight over the
    UNet's skip tensor (the residual add is its epilogue)."""

    def __init__(self, cn, device):
        super().__init__(cn, device)
        ce = cn.controlnet_cond_embedding
        self.cond_in_pad = _r8(ce.conv_in.weight.shape[1])
        self.cond_convs = [Conv(ce.conv_in, cin_pad=self.cond_in_pad)] + [Conv(b) for b in ce.blocks] + [Conv(ce.conv_out)]
        self.zero = [Linear(c.weight, c.bias) for c in cn.controlnet_down_blocks]
        self.zero_mid = Linear(cn.controlnet_mid_block.weight, cn.controlnet_mid_block.bias)
        # controlnet_conditioning_scale is folded into the zero convs (skip + s*(W f + b) = skip + (sW) f + sb):
        # the unscaled parameters are kept and the buffers the recorded plans point at are rewritten in place when
        # a job asks for another scale, so captured graphs are independent of it
        self._zero_unscaled = [(z, z.wgt.clone(), None if z.bias is None else z.bias.clone())
                               for z in self.zero + [self.zero_mid]]
        self._scale = 1.0

    def set_scale(self, scale: float):
        if scale == self._scale:
            return
        for z, w0, b0 in self._zero_unscaled:
            z.wgt.copy_(w0 * scale)
            if b0 is not None:
                z.bias.copy_(b0 * scale)
        self._scale = scale

    def record_cond_embedding(self, rec: Recorder, cond: FM) -> FM:
        """Step-invariant: run once per job, outside the per-step graph."""
        x = cond
        for i, c in enumerate(self.cond_convs):
            last = i == len(self.cond_convs) - 1
            x = c(rec, x, act=nat.ACT_NONE if last else nat.ACT_SILU)
        return x

    def record_residual_adds(self, rec: Recorder, feats: List[FM], mid: FM, unet_skips: List[FM], unet_mid: FM):
        for z, f, s in zip(self.zero, feats, unet_skips):
            rec.add(nat.igemm, f.tokens(), z.wgt, s.tokens(), n=1, h=1, w=f.rows, bias=z.bias, residual=s.tokens())
        rec.add(nat.igemm, mid.tokens(), self.zero_mid.wgt, unet_mid.tokens(), n=1, h=1, w=mid.rows,
                bias=self.zero_mid.bias, residual=unet_mid.tokens())


SCORE_BYTES = 16 << 20  # largest score block of the VAE's single-head attention (the 128x128x512 map itself is 16 MiB)


def vae_attention(rec: Recorder, x: FM, norm: GroupNorm, qkv_lin: Linear, out_lin: Linear, c: int) -> FM:
    """AutoencoderKL mid-block attention: ONE head of width c = 512, sequence = every pixel.  The head is too wide for
    the flash kernel's TMEM budget (S + O accumulators), so it runs as GEMM -> row softmax -> GEMM on blocks of
    query rows sized so that a score block never exceeds SCORE_BYTES: 1024x1024 images (S = 16384) take 32 blocks
    of 512 rows instead of one 512 MiB S x S matrix."""
    s = x.h * x.w
    g = norm(rec, x, False)
    qkv = qkv_lin(rec, g.tokens())  # [n*s, 3c]
    rec.release(g)
    out = rec.fm(x.n, x.h, x.w, c)
    attn = rec.buf(x.n * s, c)
    qc = s if s * s * 2 <= SCORE_BYTES else max(128, (SCORE_BYTES // (2 * s)) // 128 * 128)
    scores = rec.buf(qc, s)
    vt = rec.buf(c, s)
    for b in range(x.n):
        blk = qkv[b * s:(b + 1) * s]
        # q, k are column blocks of the fused projection (strided operands); V^T via the channels-last -> planar
        # transpose kernel, once per image
        rec.add(nat.nhwc_to_nchw, blk[:, 2 * c:], vt.view(1, c, 1, s), 3 * c)
        for q0 in range(0, s, qc):
            rows = min(qc, s - q0)
            sc = scores[:rows]
            rec.add(nat.igemm, blk[q0:q0 + rows, :c], blk[:, c:2 * c], sc, n=1, h=1, w=rows, out_scale=c ** -0.5)
            rec.add(nat.softmax_rows, sc, sc, 1.0)
            rec.add(nat.igemm, sc, vt, attn[b * s + q0:b * s + q0 + rows], n=1, h=1, w=rows)
    out_lin(rec, attn, residual=x.tokens(), out=out.tokens())
    for t in (attn, scores, vt, qkv):
        rec.release(t)
    return out


class VaeEncoderPlan:
    """AutoencoderKL.encode on sm_100a kernels (img2img / inpaint / pix2pix inputs; the reference reaches it inside
    `pipeline(**kwargs)`, swarm/diffusion/diffusion_func.py:96): conv_in -> down blocks (stride-2 convs as implicit
    GEMM with the (0,1,0,1) padding as TMA out-of-bounds fill) -> mid block -> GN+SiLU -> conv_out -> quant_conv.
    Output: the moments [n, h/8, w/8, 2*latent] channels-last."""

    def __init__(self, vae, device):
        enc = vae.encoder
        self.in_pad = _r8(vae.config["in_channels"])
        self.conv_in = Conv(enc.conv_in, cin_pad=self.in_pad)
        self.downs = []
        for blk in enc.down_blocks:
            self.downs.append(dict(resnets=[Resnet(r) for r in blk.resnets],
                                   down=Conv(blk.downsamplers[0].conv, asym_pad=True) if blk.downsamplers is not None else None))
        self.mid_res = [Resnet(r) for r in enc.mid_block.resnets]
        at = enc.mid_block.attentions[0]
        self.attn_norm = GroupNorm(at.group_norm)
        self.attn_qkv = Linear(torch.cat([at.to_q.weight, at.to_k.weight, at.to_v.weight], 0),
                               torch.cat([at.to_q.bias, at.to_k.bias, at.to_v.bias], 0))
        self.attn_out = Linear(at.to_out[0].weight, at.to_out[0].bias)
        self.attn_c = at.to_q.weight.shape[0]
        self.norm_out = GroupNorm(enc.conv_norm_out)
        self.conv_out = Conv(enc.conv_out)
        self.quant = Linear(vae.quant_conv.weight, vae.quant_conv.bias)
        self.moments = 2 * vae.config["latent_channels"]

    def record(self, rec: Recorder, img: FM) -> FM:
        x = self.conv_in(rec, img)
        for blk in self.downs:
            for r in blk["resnets"]:
                x2 = r(rec, x)
                rec.release(x)
                x = x2
            if blk["down"] is not None:
                x2 = blk["down"](rec, x)
                rec.release(x)
                x = x2
        x2 = self.mid_res[0](rec, x)
        rec.release(x)
        x = vae_attention(rec, x2, self.attn_norm, self.attn_qkv, self.attn_out, self.attn_c)
        rec.release(x2)
        x2 = self.mid_res[1](rec, x)
        rec.release(x)
        g = self.norm_out(rec, x2, True)
        rec.release(x2)
        h = self.conv_out(rec, g)
        rec.release(g)
        out = rec.fm(h.n, h.h, h.w, self.quant.nout)
        self.quant(rec, h.tokens(), out=out.tokens())
        rec.release(h)
        return out


class VaeDecoderPlan:
    """AutoencoderKL.decode on sm_100a kernels, ending in a uint8 HWC image (row W8)."""

    def __init__(self, vae, device):
        self.vae_cfg = vae.config
        self.lat = vae.config["latent_channels"]
        self.lat_pad = _r8(self.lat)
        dec = vae.decoder
        self.post_quant = Linear(vae.post_quant_conv.weight, vae.post_quant_conv.bias)
        self.conv_in = Conv(dec.conv_in, cin_pad=self.lat_pad)
        self.mid_res = [Resnet(r) for r in dec.mid_block.resnets]
        at = dec.mid_block.attentions[0]
        self.attn_norm = GroupNorm(at.group_norm)
        self.attn_qkv = Linear(torch.cat([at.to_q.weight, at.to_k.weight, at.to_v.weight], 0),
                               torch.cat([at.to_q.bias, at.to_k.bias, at.to_v.bias], 0))
        self.attn_out = Linear(at.to_out[0].weight, at.to_out[0].bias)
        self.attn_c = at.to_q.weight.shape[0]
        self.ups = []
        for blk in dec.up_blocks:
            self.ups.append(dict(resnets=[Resnet(r) for r in blk.resnets],
                                 up=Upsample(blk.upsamplers[0]) if blk.upsamplers is not None else None))
        self.norm_out = GroupNorm(dec.conv_norm_out)
        self.conv_out = Conv(dec.conv_out)

    def _attention(self, rec: Recorder, x: FM) -> FM:
        return vae_attention(rec, x, self.attn_norm, self.attn_qkv, self.attn_out, self.attn_c)

    def record(self, rec: Recorder, z: FM):
        """z: latents / scaling_factor (+ post-quant bias), channels-last, padded to 8 channels.
        Returns (image FM [n, H, W, 8], of which 3 channels are real)."""
        zq = rec.fm(z.n, z.h, z.w, self.lat_pad)
        self.post_quant(rec, z.tokens(), out=zq.tokens())
        x = self.conv_in(rec, zq)
        rec.release(zq)
        x2 = self.mid_res[0](rec, x)
        rec.release(x)
        x = self._attention(rec, x2)
        rec.release(x2)
        x2 = self.mid_res[1](rec, x)
        rec.release(x)
        x = x2
        for blk in self.ups:
            for r in blk["resnets"]:
                x2 = r(rec, x)
                rec.release(x)
                x = x2
            if blk["up"] is not None:
                x2 = blk["up"](rec, x)
                rec.release(x)
                x = x2
        g = self.norm_out(rec, x, True)
        rec.release(x)
        img = self.conv_out(rec, g)
        rec.release(g)
        return img
