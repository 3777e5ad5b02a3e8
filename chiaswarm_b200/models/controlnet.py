"""ControlNetModel (UNet encoder + mid copy, conditioning embedder
3->16->32->96->256->320 with three stride-2 stages, 12+1 zero 1x1 convs; loaded by the reference
at swarm/diffusion/diffusion_func.py:29-34).  Names follow the flat diffusers ControlNet repo."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet2d import SD15_CONFIG, DownBlock, MidBlock, TimestepEmbedding, _as_list, timestep_embedding


class ControlNetConditioningEmbedding(nn.Module):
    def __init__(self, out_ch, cond_ch=3, block_out=(16, 32, 96, 256)):
        super().__init__()
        self.conv_in = nn.Conv2d(cond_ch, block_out[0], 3, padding=1)
        blocks = []
        for i in range(len(block_out) - 1):
            blocks.append(nn.Conv2d(block_out[i], block_out[i], 3, padding=1))
            blocks.append(nn.Conv2d(block_out[i], block_out[i + 1], 3, padding=1, stride=2))
        self.blocks = nn.ModuleList(blocks)
        self.conv_out = nn.Conv2d(block_out[-1], out_ch, 3, padding=1)

    def forward(self, cond):
        x = F.silu(self.conv_in(cond))
        for b in self.blocks:
            x = F.silu(b(x))
        return self.conv_out(x)


class ControlNetModel(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(SD15_CONFIG)
        c.update(cfg)
        self.config = c
        boc = list(c["block_out_channels"])
        n = len(boc)
        groups, eps = c["norm_num_groups"], c["norm_eps"]
        temb_dim = boc[0] * 4
        layers = _as_list(c["layers_per_block"], n)
        heads = _as_list(c.get("num_attention_heads") or c["attention_head_dim"], n)
        depth = _as_list(c.get("transformer_layers_per_block", 1), n)
        cross, ulp = c["cross_attention_dim"], c.get("use_linear_projection", False)

        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(boc[0], temb_dim)
        self.controlnet_cond_embedding = ControlNetConditioningEmbedding(
            boc[0], c.get("conditioning_channels", 3), tuple(c.get("conditioning_embedding_out_channels", (16, 32, 96, 256))))
        self.down_blocks = nn.ModuleList()
        self.controlnet_down_blocks = nn.ModuleList([nn.Conv2d(boc[0], boc[0], 1)])
        ch = boc[0]
        for i, typ in enumerate(c["down_block_types"]):
            cin, ch = ch, boc[i]
            self.down_blocks.append(DownBlock(cin, ch, temb_dim, layers[i], groups, eps, "CrossAttn" in typ, heads[i],
                                              ch // heads[i], depth[i], cross, i < n - 1, ulp))
            for _ in range(layers[i]):
                self.controlnet_down_blocks.append(nn.Conv2d(ch, ch, 1))
            if i < n - 1:
                self.controlnet_down_blocks.append(nn.Conv2d(ch, ch, 1))
        self.mid_block = MidBlock(boc[-1], temb_dim, groups, eps, heads[-1], boc[-1] // heads[-1], depth[-1], cross, ulp)
        self.controlnet_mid_block = nn.Conv2d(boc[-1], boc[-1], 1)

    def forward(self, sample, timestep, encoder_hidden_states, controlnet_cond, conditioning_scale=1.0):
        t = torch.as_tensor(timestep, device=sample.device)
        if t.dim() == 0:
            t = t[None].expand(sample.shape[0])
        emb = self.time_embedding(timestep_embedding(t, self.config["block_out_channels"][0]).to(sample.dtype))
        x = self.conv_in(sample) + self.controlnet_cond_embedding(controlnet_cond)
        skips = [x]
        for blk in self.down_blocks:
            x, outs = blk(x, emb, encoder_hidden_states)
            skips.extend(outs)
        x = self.mid_block(x, emb, encoder_hidden_states)
        down = [conv(s) * conditioning_scale for s, conv in zip(skips, self.controlnet_down_blocks)]
        mid = self.controlnet_mid_block(x) * conditioning_scale
        return down, mid
