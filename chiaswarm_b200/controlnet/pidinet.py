"""PiDiNet soft-edge annotator implemented in this tree (reference: `PidiNetDetector.from_pretrained("lllyasviel/Annotators")`
in swarm/controlnet/input_processor.py:51 -- the detector behind the `softedge` pre-processor).

`PiDiNet` is the "table5_pidinet" network (width 60, configuration carv4, dilation module 24, spatial attention): sixteen
3x3 depthwise PIXEL-DIFFERENCE convolutions in four stages (60, 120, 240, 240 channels; 2x2 max-pool + 1x1 shortcut at
each stage change), each followed by a 1x1 convolution and a residual; per stage a compact dilation module (1x1, then four
dilated 3x3 convolutions summed), a spatial attention gate and a 1x1 reduction to one map; the four maps, resized to the
input, are fused by a 1x1 classifier.  A pixel-difference convolution is an ordinary convolution with a rewritten kernel:

    cv  the kernel as stored
    cd  central difference:  conv(x, w) - conv(x, sum(w))           (every tap minus the centre pixel)
    ad  angular difference:  w - w rotated one step clockwise around the centre
    rd  radial difference:   the eight outer taps of a 5x5 kernel get +w, the eight inner ring taps -w

carv4 cycles cd, ad, rd, cv over the sixteen layers.  Parameter names match `table5_pidinet.pth` (`init_block`,
`block{s}_{i}.conv1 / conv2 / shortcut`, `dilations`, `attentions`, `conv_reduces`, `classifier`; the file prefixes them
with `module.`), loading is strict."""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..utils import checkpoints

CARV4 = ["cd", "ad", "rd", "cv"] * 4
_CLOCKWISE = [3, 0, 1, 6, 4, 2, 7, 8, 5]
_RD_OUTER, _RD_INNER = [0, 2, 4, 10, 14, 20, 22, 24], [6, 7, 8, 11, 13, 16, 17, 18]


class PDConv2d(nn.Module):
    """3x3 pixel-difference convolution (bias-free), evaluated as a plain convolution with a rewritten kernel."""

    def __init__(self, op, cin, cout, groups=1):
        super().__init__()
        self.op, self.groups = op, groups
        self.weight = nn.Parameter(torch.empty(cout, cin // groups, 3, 3))
        nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

    def forward(self, x):
        w = self.weight
        if self.op == "cv":
            return F.conv2d(x, w, padding=1, groups=self.groups)
        if self.op == "cd":
            centre = w.sum(dim=[2, 3], keepdim=True)
            return F.conv2d(x, w, padding=1, groups=self.groups) - F.conv2d(x, centre, padding=0, groups=self.groups)
        flat = w.view(w.shape[0], w.shape[1], -1)
        if self.op == "ad":
            return F.conv2d(x, (flat - flat[:, :, _CLOCKWISE]).view(w.shape), padding=1, groups=self.groups)
        assert self.op == "rd", self.op
        buf = torch.zeros(w.shape[0], w.shape[1], 25, device=w.device, dtype=w.dtype)
        buf[:, :, _RD_OUTER] = flat[:, :, 1:]
        buf[:, :, _RD_INNER] = -flat[:, :, 1:]
        return F.conv2d(x, buf.view(w.shape[0], w.shape[1], 5, 5), padding=2, groups=self.groups)


class PDCBlock(nn.Module):
    def __init__(self, op, cin, cout, stride=1):
        super().__init__()
        self.stride = stride
        if stride > 1:
            self.pool = nn.MaxPool2d(kernel_size=2, stride=2)
            self.shortcut = nn.Conv2d(cin, cout, kernel_size=1, padding=0)
        self.conv1 = PDConv2d(op, cin, cin, groups=cin)
        self.relu2 = nn.ReLU()
        self.conv2 = nn.Conv2d(cin, cout, kernel_size=1, padding=0, bias=False)

    def forward(self, x):
        if self.stride > 1:
            x = self.pool(x)
        y = self.conv2(self.relu2(self.conv1(x)))
        return y + (self.shortcut(x) if self.stride > 1 else x)


class CSAM(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.relu1 = nn.ReLU()
        self.conv1 = nn.Conv2d(channels, 4, kernel_size=1, padding=0)
        self.conv2 = nn.Conv2d(4, 1, kernel_size=3, padding=1, bias=False)
        self.sigmoid = nn.Sigmoid()

    def forward(self, x):
        return x * self.sigmoid(self.conv2(self.conv1(self.relu1(x))))


class CDCM(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.relu1 = nn.ReLU()
        self.conv1 = nn.Conv2d(cin, cout, kernel_size=1, padding=0)
        for i, d in enumerate((5, 7, 9, 11), start=1):
            setattr(self, f"conv2_{i}", nn.Conv2d(cout, cout, kernel_size=3, dilation=d, padding=d, bias=False))

    def forward(self, x):
        x = self.conv1(self.relu1(x))
        return self.conv2_1(x) + self.conv2_2(x) + self.conv2_3(x) + self.conv2_4(x)


class MapReduce(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, 1, kernel_size=1, padding=0)

    def forward(self, x):
        return self.conv(x)


class PiDiNet(nn.Module):
    def __init__(self, width=60, ops=CARV4, dil=24):
        super().__init__()
        planes = [width, width * 2, width * 4, width * 4]
        self.init_block = PDConv2d(ops[0], 3, width)
        k, cin = 1, width
        for stage, cout in enumerate(planes, start=1):
            count = 3 if stage == 1 else 4
            for i in range(1, count + 1):
                first = stage > 1 and i == 1
                setattr(self, f"block{stage}_{i}", PDCBlock(ops[k], cin if first else cout, cout, stride=2 if first else 1))
                k += 1
            cin = cout
        self.stages = [[f"block{s}_{i}" for i in range(1, (3 if s == 1 else 4) + 1)] for s in range(1, 5)]
        self.attentions = nn.ModuleList([CSAM(dil) for _ in planes])
        self.dilations = nn.ModuleList([CDCM(c, dil) for c in planes])
        self.conv_reduces = nn.ModuleList([MapReduce(dil) for _ in planes])
        self.classifier = nn.Conv2d(4, 1, kernel_size=1)

    def forward(self, x):
        """x: [n, 3, h, w] (BGR, 0..1).  Returns the five edge maps [n, 1, h, w] in 0..1; the last is the fused one."""
        h, w = x.shape[-2:]
        y = self.init_block(x)
        maps = []
        for i, names in enumerate(self.stages):
            for name in names:
                y = getattr(self, name)(y)
            e = self.conv_reduces[i](self.attentions[i](self.dilations[i](y)))
            maps.append(F.interpolate(e, (h, w), mode="bilinear", align_corners=False))
        maps.append(self.classifier(torch.cat(maps, dim=1)))
        return [torch.sigmoid(m) for m in maps]

    @torch.no_grad()
    def edges(self, image: Image.Image, device="cpu") -> Image.Image:
        arr = np.asarray(image.convert("RGB"), dtype=np.float32)[:, :, ::-1].copy() / 255.0
        x = torch.from_numpy(arr).permute(2, 0, 1)[None].to(device)
        edge = self(x)[-1][0, 0].float().clamp(0, 1)
        out = (edge * 255).to(torch.uint8).cpu().numpy()
        return Image.fromarray(np.stack([out] * 3, axis=2))


_cache = {}


def load_pidinet(device="cpu") -> Optional[PiDiNet]:
    """The detector with its trained weights when `lllyasviel/Annotators/table5_pidinet.pth` is local, else None."""
    if "pidi" in _cache:
        return _cache["pidi"]
    net = None
    root = checkpoints.locate_checkpoint("lllyasviel/Annotators")
    if root is not None and (Path(root) / "table5_pidinet.pth").is_file():
        sd = checkpoints.load_state_dict(Path(root) / "table5_pidinet.pth")
        sd = sd.get("state_dict", sd)
        net = PiDiNet()
        net.load_state_dict({k[len("module."):] if k.startswith("module.") else k: v for k, v in sd.items()}, strict=True)
        net = net.to(device).eval()
    _cache["pidi"] = net
    return net
