"""NormalBae surface-normal annotator implemented in this tree (reference:
`NormalBaeDetector.from_pretrained("lllyasviel/Annotators")` in swarm/controlnet/input_processor.py:31 -- the detector
ControlNet's normalbae conditioning was trained on; Bae et al., "Estimating and Exploiting the Aleatoric Uncertainty in
Surface Normal Estimation").

`NNET` = an EfficientNet-B5 encoder (the `tf_efficientnet_b5_ap` layout of gen-efficientnet: TensorFlow "same" padding,
swish, squeeze-excite, BatchNorm eps 1e-3; widths 48 | 24 40 64 128 176 304 512 | 2048, depths 3 5 5 7 7 9 3) whose
features after stages 0, 1, 2, 4 and after the 2048-wide head convolution feed a decoder: 1x1 on the deepest map, four
"upsample to the skip, concatenate, two 3x3 conv + BN + LeakyReLU" blocks, a normal + concentration prediction at 1/8
resolution and three refinements (1/4, 1/2, full) by per-pixel MLPs on [bilinearly upsampled features ; upsampled previous
prediction].  Each prediction is (nx, ny, nz) normalised to unit length plus kappa = elu + 1.01.  Parameter names match
`scannet.pt` (`encoder.original_model.*`, `decoder.*`; the file wraps them in {"model": ...} under `module.`), loading is
strict.  The conditioning image is (n + 1) / 2 of the full-resolution normals."""
from __future__ import annotations

import math
from pathlib import Path
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..utils import checkpoints

BN_EPS = 1e-3
# (expand ratio, kernel, stride, output channels, repeats) of EfficientNet-B0, scaled below to B5 (width 1.6, depth 2.2)
_B0 = [(1, 3, 1, 16, 1), (6, 3, 2, 24, 2), (6, 5, 2, 40, 2), (6, 3, 2, 80, 3), (6, 5, 1, 112, 3), (6, 5, 2, 192, 4), (6, 3, 1, 320, 1)]


def _round_channels(c, mult=1.6, divisor=8):
    c *= mult
    new = max(divisor, int(c + divisor / 2) // divisor * divisor)
    return new + divisor if new < 0.9 * c else new


class Conv2dSame(nn.Conv2d):
    """TensorFlow 'SAME' padding: output = ceil(input / stride), the odd pixel of padding goes bottom / right."""

    def forward(self, x):
        ih, iw = x.shape[-2:]
        kh, kw = self.weight.shape[-2:]
        sh, sw = self.stride
        ph = max((math.ceil(ih / sh) - 1) * sh + (kh - 1) * self.dilation[0] + 1 - ih, 0)
        pw = max((math.ceil(iw / sw) - 1) * sw + (kw - 1) * self.dilation[1] + 1 - iw, 0)
        if ph > 0 or pw > 0:
            x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
        return F.conv2d(x, self.weight, self.bias, self.stride, (0, 0), self.dilation, self.groups)


def _bn(c):
    return nn.BatchNorm2d(c, eps=BN_EPS)


class SqueezeExcite(nn.Module):
    def __init__(self, channels, reduced):
        super().__init__()
        self.conv_reduce = nn.Conv2d(channels, reduced, 1, bias=True)
        self.conv_expand = nn.Conv2d(reduced, channels, 1, bias=True)

    def forward(self, x):
        s = x.mean((2, 3), keepdim=True)
        return x * torch.sigmoid(self.conv_expand(F.silu(self.conv_reduce(s))))


class DepthwiseSeparableConv(nn.Module):
    """First-stage block (expand ratio 1): depthwise, SE, pointwise; residual when shapes allow."""

    def __init__(self, cin, cout, kernel, stride):
        super().__init__()
        self.has_residual = stride == 1 and cin == cout
        self.conv_dw = Conv2dSame(cin, cin, kernel, stride, groups=cin, bias=False)
        self.bn1 = _bn(cin)
        self.se = SqueezeExcite(cin, max(1, int(cin * 0.25)))
        self.conv_pw = Conv2dSame(cin, cout, 1, bias=False)
        self.bn2 = _bn(cout)

    def forward(self, x):
        y = self.se(F.silu(self.bn1(self.conv_dw(x))))
        y = self.bn2(self.conv_pw(y))
        return x + y if self.has_residual else y


class InvertedResidual(nn.Module):
    def __init__(self, cin, cout, kernel, stride, expand):
        super().__init__()
        mid = cin * expand
        self.has_residual = stride == 1 and cin == cout
        self.conv_pw = Conv2dSame(cin, mid, 1, bias=False)
        self.bn1 = _bn(mid)
        self.conv_dw = Conv2dSame(mid, mid, kernel, stride, groups=mid, bias=False)
        self.bn2 = _bn(mid)
        self.se = SqueezeExcite(mid, max(1, int(cin * 0.25)))  # reduced from the block's INPUT width
        self.conv_pwl = Conv2dSame(mid, cout, 1, bias=False)
        self.bn3 = _bn(cout)

    def forward(self, x):
        y = F.silu(self.bn1(self.conv_pw(x)))
        y = self.se(F.silu(self.bn2(self.conv_dw(y))))
        y = self.bn3(self.conv_pwl(y))
        return x + y if self.has_residual else y


class EfficientNetB5(nn.Module):
    """gen-efficientnet's GenEfficientNet without pooling / classifier; `features()` returns what the decoder reads."""

    def __init__(self):
        super().__init__()
        stem = _round_channels(32)
        self.conv_stem = Conv2dSame(3, stem, 3, 2, bias=False)
        self.bn1 = _bn(stem)
        stages, cin = [], stem
        for expand, kernel, stride, c, repeats in _B0:
            cout, n = _round_channels(c), int(math.ceil(repeats * 2.2))
            blocks = []
            for i in range(n):
                s = stride if i == 0 else 1
                blocks.append(DepthwiseSeparableConv(cin, cout, kernel, s) if expand == 1
                              else InvertedResidual(cin, cout, kernel, s, expand))
                cin = cout
            stages.append(nn.Sequential(*blocks))
        self.blocks = nn.Sequential(*stages)
        self.conv_head = Conv2dSame(cin, _round_channels(1280), 1, bias=False)
        self.bn2 = _bn(_round_channels(1280))

    def features(self, x):
        x = F.silu(self.bn1(self.conv_stem(x)))
        outs = []
        for stage in self.blocks:
            x = stage(x)
            outs.append(x)
        # stages 0, 1, 2, 4 (1/2, 1/4, 1/8, 1/16 of the input) and the head convolution BEFORE its norm (1/32)
        return outs[0], outs[1], outs[2], outs[4], self.conv_head(x)


class _Encoder(nn.Module):
    def __init__(self):
        super().__init__()
        self.original_model = EfficientNetB5()

    def forward(self, x):
        return self.original_model.features(x)


class UpSampleBN(nn.Module):
    def __init__(self, skip_input, output_features):
        super().__init__()
        self._net = nn.Sequential(nn.Conv2d(skip_input, output_features, 3, 1, 1), nn.BatchNorm2d(output_features), nn.LeakyReLU(),
                                  nn.Conv2d(output_features, output_features, 3, 1, 1), nn.BatchNorm2d(output_features),
                                  nn.LeakyReLU())

    def forward(self, x, skip):
        up = F.interpolate(x, size=skip.shape[2:], mode="bilinear", align_corners=True)
        return self._net(torch.cat([up, skip], dim=1))


def _pixel_mlp(cin):
    return nn.Sequential(nn.Conv1d(cin, 128, 1), nn.ReLU(), nn.Conv1d(128, 128, 1), nn.ReLU(), nn.Conv1d(128, 128, 1), nn.ReLU(),
                         nn.Conv1d(128, 4, 1))


def norm_normalize(out, min_kappa=0.01):
    n, kappa = out[:, :3], out[:, 3:4]
    n = n / (torch.sqrt((n * n).sum(1, keepdim=True)) + 1e-10)
    return torch.cat([n, F.elu(kappa) + 1.0 + min_kappa], dim=1)


class _Decoder(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv2 = nn.Conv2d(2048, 2048, 1)
        self.up1, self.up2 = UpSampleBN(2048 + 176, 1024), UpSampleBN(1024 + 64, 512)
        self.up3, self.up4 = UpSampleBN(512 + 40, 256), UpSampleBN(256 + 24, 128)
        self.out_conv_res8 = nn.Conv2d(512, 4, 3, 1, 1)
        self.out_conv_res4, self.out_conv_res2, self.out_conv_res1 = _pixel_mlp(512 + 4), _pixel_mlp(256 + 4), _pixel_mlp(128 + 4)

    @staticmethod
    def _refine(mlp, feat, prev):
        """Test-time path of the original: every pixel is refined (no uncertainty-guided sampling)."""
        feat = F.interpolate(feat, scale_factor=2, mode="bilinear", align_corners=True)
        prev = F.interpolate(prev, scale_factor=2, mode="bilinear", align_corners=True)
        x = torch.cat([feat, prev], dim=1)
        b, c, h, w = x.shape
        return norm_normalize(mlp(x.view(b, c, -1))).view(b, 4, h, w)

    def forward(self, feats):
        b0, b1, b2, b3, b4 = feats
        d0 = self.conv2(b4)
        d1 = self.up1(d0, b3)
        d2 = self.up2(d1, b2)
        d3 = self.up3(d2, b1)
        d4 = self.up4(d3, b0)
        res8 = norm_normalize(self.out_conv_res8(d2))
        res4 = self._refine(self.out_conv_res4, d2, res8)
        res2 = self._refine(self.out_conv_res2, d3, res4)
        res1 = self._refine(self.out_conv_res1, d4, res2)
        return [res8, res4, res2, res1]


class NNET(nn.Module):
    def __init__(self):
        super().__init__()
        self.encoder, self.decoder = _Encoder(), _Decoder()

    def forward(self, x):
        """x: [n, 3, h, w], ImageNet-normalised RGB, h and w multiples of 32.  Returns the four predictions
        [n, 4, ., .] (unit normal + kappa) at 1/8, 1/4, 1/2 and full resolution."""
        return self.decoder(self.encoder(x))

    @torch.no_grad()
    def normals(self, image: Image.Image, device="cpu") -> Image.Image:
        rgb = image.convert("RGB")
        w, h = rgb.size
        w64, h64 = max(64, int(round(w / 64.0)) * 64), max(64, int(round(h / 64.0)) * 64)
        arr = np.asarray(rgb.resize((w64, h64), Image.Resampling.BILINEAR), dtype=np.float32) / 255.0
        x = torch.from_numpy(arr).permute(2, 0, 1)[None].to(device)
        mean = torch.tensor([0.485, 0.456, 0.406], device=x.device).view(1, 3, 1, 1)
        std = torch.tensor([0.229, 0.224, 0.225], device=x.device).view(1, 3, 1, 1)
        n = self((x - mean) / std)[-1][0, :3].float()
        out = (((n + 1) * 0.5).clamp(0, 1) * 255.0).permute(1, 2, 0).to(torch.uint8).cpu().numpy()
        img = Image.fromarray(out)
        return img if (w64, h64) == (w, h) else img.resize((w, h), Image.Resampling.BILINEAR)


_cache = {}


def load_normalbae(device="cpu") -> Optional[NNET]:
    """The network with its trained weights when `lllyasviel/Annotators/scannet.pt` is available locally, else None."""
    if "nnet" in _cache:
        return _cache["nnet"]
    net = None
    root = checkpoints.locate_checkpoint("lllyasviel/Annotators")
    if root is not None and (Path(root) / "scannet.pt").is_file():
        sd = checkpoints.load_state_dict(Path(root) / "scannet.pt")
        sd = sd.get("model", sd)
        net = NNET()
        net.load_state_dict({k[len("module."):] if k.startswith("module.") else k: v for k, v in sd.items()}, strict=True)
        net = net.to(device).eval()
    _cache["nnet"] = net
    return net
