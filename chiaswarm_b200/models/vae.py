"""AutoencoderKL (reference: every image/video pipeline decodes -- and img2img / inpaint / pix2pix
encode -- through it inside `pipeline(**kwargs)`, swarm/diffusion/diffusion_func.py:96; structure
in SURVEY.md section 2.4 row K10).  Parameter names follow the diffusers `vae/` checkpoint layout;
the legacy attention names (query/key/value/proj_attn) are remapped by `utils.checkpoints`."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet2d import Downsample2D, ResnetBlock2D, Upsample2D

SD_VAE_CONFIG = dict(in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(128, 256, 512, 512),
                     layers_per_block=2, norm_num_groups=32, scaling_factor=0.18215)


class VaeAttention(nn.Module):
    """Single-head spatial self-attention (d = channels)."""

    def __init__(self, ch, groups):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, ch, eps=1e-6)
        self.to_q = nn.Linear(ch, ch)
        self.to_k = nn.Linear(ch, ch)
        self.to_v = nn.Linear(ch, ch)
        self.to_out = nn.ModuleList([nn.Linear(ch, ch), nn.Identity()])

    def forward(self, x):
        b, c, h, w = x.shape
        y = self.group_norm(x).reshape(b, c, h * w).transpose(1, 2)
        q, k, v = self.to_q(y), self.to_k(y), self.to_v(y)
        o = F.scaled_dot_product_attention(q[:, None], k[:, None], v[:, None])[:, 0]
        o = self.to_out[0](o).transpose(1, 2).reshape(b, c, h, w)
        return x + o


class VaeMidBlock(nn.Module):
    def __init__(self, ch, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ch, ch, 0, groups, 1e-6) for _ in range(2)])
        self.attentions = nn.ModuleList([VaeAttention(ch, groups)])

    def forward(self, x):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class VaeDownBlock(nn.Module):
    def __init__(self, cin, cout, layers, groups, add_down):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, 0, groups, 1e-6) for i in range(layers)])
        self.downsamplers = nn.ModuleList([Downsample2D(cout, padding=0)]) if add_down else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
        return x


class VaeUpBlock(nn.Module):
    def __init__(self, cin, cout, layers, groups, add_up):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, 0, groups, 1e-6) for i in range(layers)])
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class Encoder(nn.Module):
    def __init__(self, cin, latent, boc, layers, groups):
        super().__init__()
        self.conv_in = nn.Conv2d(cin, boc[0], 3, padding=1)
        self.down_blocks = nn.ModuleList()
        ch = boc[0]
        for i, co in enumerate(boc):
            self.down_blocks.append(VaeDownBlock(ch, co, layers, groups, i < len(boc) - 1))
            ch = co
        self.mid_block = VaeMidBlock(ch, groups)
        self.conv_norm_out = nn.GroupNorm(groups, ch, eps=1e-6)
        self.conv_out = nn.Conv2d(ch, 2 * latent, 3, padding=1)

    def forward(self, x):
        x = self.conv_in(x)
        for b in self.down_blocks:
            x = b(x)
        x = self.mid_block(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class Decoder(nn.Module):
    def __init__(self, latent, cout, boc, layers, groups):
        super().__init__()
        rev = list(boc)[::-1]
        self.conv_in = nn.Conv2d(latent, rev[0], 3, padding=1)
        self.mid_block = VaeMidBlock(rev[0], groups)
        self.up_blocks = nn.ModuleList()
        ch = rev[0]
        for i, co in enumerate(rev):
            self.up_blocks.append(VaeUpBlock(ch, co, layers + 1, groups, i < len(rev) - 1))
            ch = co
        self.conv_norm_out = nn.GroupNorm(groups, ch, eps=1e-6)
        self.conv_out = nn.Conv2d(ch, cout, 3, padding=1)

    def forward(self, z):
        x = self.mid_block(self.conv_in(z))
        for b in self.up_blocks:
            x = b(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class AutoencoderKL(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(SD_VAE_CONFIG)
        c.update(cfg)
        self.config = c
        boc, g = c["block_out_channels"], c["norm_num_groups"]
        self.encoder = Encoder(c["in_channels"], c["latent_channels"], boc, c["layers_per_block"], g)
        self.decoder = Decoder(c["latent_channels"], c["out_channels"], boc, c["layers_per_block"], g)
        self.quant_conv = nn.Conv2d(2 * c["latent_channels"], 2 * c["latent_channels"], 1)
        self.post_quant_conv = nn.Conv2d(c["latent_channels"], c["latent_channels"], 1)

    @property
    def scaling_factor(self):
        return self.config.get("scaling_factor", 0.18215)

    def encode_moments(self, x):
        return self.quant_conv(self.encoder(x)).chunk(2, dim=1)

    def encode(self, x, generator=None, sample=True):
        mean, logvar = self.encode_moments(x)
        if not sample:
            return mean
        std = torch.exp(0.5 * logvar.clamp(-30.0, 20.0))
        noise = torch.randn(mean.shape, generator=generator, device=mean.device, dtype=mean.dtype)
        return mean + std * noise

    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))
