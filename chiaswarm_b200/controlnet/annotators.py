"""Learned ControlNet annotators implemented in this tree (reference: `controlnet_aux` detectors loaded in
swarm/controlnet/input_processor.py:100-231).

`HEDNet` is the holistically-nested edge detector that ControlNet's soft-edge / scribble conditioning was trained on
(`lllyasviel/Annotators/ControlNetHED.pth`): five VGG-style stages, each with a 1x1 side output; the edge map is the
sigmoid of the mean of the up-sampled side outputs.  Parameter names match that checkpoint (`norm`,
`block{1..5}.convs.{i}`, `block{n}.projection`), so the file loads with `load_state_dict` unchanged.  It runs on
whatever device the job has (a few hundred MFLOP per image: far off the hot path) and is used whenever the checkpoint
is present locally; without it the classical stand-in in input_processor.py is used.

`LineartNet` is the line-drawing generator behind ControlNet's lineart conditioning (`lllyasviel/Annotators/sk_model.pth`,
the "informative drawings" generator): a 7x7 stem, two stride-2 convolutions, three residual blocks, two transposed
convolutions, a 7x7 head with a sigmoid; instance normalisation without parameters, reflection padding.  Parameter names
match the checkpoint (`model0.1`, `model1.{0,3}`, `model2.{i}.conv_block.{1,5}`, `model3.{0,3}`, `model4.1`); the lines
come out dark on white and are inverted to ControlNet's white-on-black convention."""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..utils import checkpoints


class _Stage(nn.Module):
    def __init__(self, cin, cout, layers):
        super().__init__()
        self.convs = nn.Sequential(*[nn.Conv2d(cin if i == 0 else cout, cout, 3, padding=1) for i in range(layers)])
        self.projection = nn.Conv2d(cout, 1, 1)

    def forward(self, x, down: bool):
        if down:
            x = F.max_pool2d(x, 2, 2)
        for conv in self.convs:
            x = F.relu(conv(x))
        return x, self.projection(x)


class HEDNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.norm = nn.Parameter(torch.zeros(1, 3, 1, 1))
        self.block1 = _Stage(3, 64, 2)
        self.block2 = _Stage(64, 128, 2)
        self.block3 = _Stage(128, 256, 3)
        self.block4 = _Stage(256, 512, 3)
        self.block5 = _Stage(512, 512, 3)

    def forward(self, x):
        """x: [n, 3, h, w] in 0..255.  Returns the five side outputs (logits) at their own resolutions."""
        h = x - self.norm
        sides = []
        for i, blk in enumerate((self.block1, self.block2, self.block3, self.block4, self.block5)):
            h, p = blk(h, down=i > 0)
            sides.append(p)
        return sides

    @torch.no_grad()
    def edges(self, image: Image.Image, device="cpu", scribble: bool = False) -> Image.Image:
        arr = np.asarray(image.convert("RGB"), dtype=np.float32)
        x = torch.from_numpy(arr).permute(2, 0, 1)[None].to(device)
        hh, ww = x.shape[-2:]
        sides = [F.interpolate(s.float(), size=(hh, ww), mode="bilinear", align_corners=False) for s in self(x)]
        edge = torch.sigmoid(torch.stack(sides).mean(0))[0, 0]
        out = (edge.clamp(0, 1) * 255).to(torch.uint8).cpu().numpy()
        if scribble:
            out = scribble_strokes(out)
        return Image.fromarray(np.stack([out] * 3, axis=2))


class _ResidualBlock(nn.Module):
    def __init__(self, ch):
        super().__init__()
        self.conv_block = nn.Sequential(nn.ReflectionPad2d(1), nn.Conv2d(ch, ch, 3), nn.InstanceNorm2d(ch), nn.ReLU(inplace=True),
                                        nn.ReflectionPad2d(1), nn.Conv2d(ch, ch, 3), nn.InstanceNorm2d(ch))

    def forward(self, x):
        return x + self.conv_block(x)


class LineartNet(nn.Module):
    def __init__(self, cin=3, cout=1, blocks=3):
        super().__init__()
        self.model0 = nn.Sequential(nn.ReflectionPad2d(3), nn.Conv2d(cin, 64, 7), nn.InstanceNorm2d(64), nn.ReLU(inplace=True))
        self.model1 = nn.Sequential(nn.Conv2d(64, 128, 3, stride=2, padding=1), nn.InstanceNorm2d(128), nn.ReLU(inplace=True),
                                    nn.Conv2d(128, 256, 3, stride=2, padding=1), nn.InstanceNorm2d(256), nn.ReLU(inplace=True))
        self.model2 = nn.Sequential(*[_ResidualBlock(256) for _ in range(blocks)])
        self.model3 = nn.Sequential(
            nn.ConvTranspose2d(256, 128, 3, stride=2, padding=1, output_padding=1), nn.InstanceNorm2d(128), nn.ReLU(inplace=True),
            nn.ConvTranspose2d(128, 64, 3, stride=2, padding=1, output_padding=1), nn.InstanceNorm2d(64), nn.ReLU(inplace=True))
        self.model4 = nn.Sequential(nn.ReflectionPad2d(3), nn.Conv2d(64, cout, 7), nn.Sigmoid())

    def forward(self, x):
        """x: [n, 3, h, w] in 0..1 (h, w multiples of 4).  Returns [n, 1, h, w] in 0..1: dark lines on white."""
        return self.model4(self.model3(self.model2(self.model1(self.model0(x)))))

    @torch.no_grad()
    def lines(self, image: Image.Image, device="cpu") -> Image.Image:
        rgb = image.convert("RGB")
        w, h = rgb.size
        w4, h4 = max(8, w - w % 4), max(8, h - h % 4)
        arr = np.asarray(rgb if (w4, h4) == (w, h) else rgb.resize((w4, h4), Image.Resampling.BILINEAR), dtype=np.float32)
        x = torch.from_numpy(arr / 255.0).permute(2, 0, 1)[None].to(device)
        line = self(x)[0, 0].float().clamp(0, 1)
        out = (255 - (line * 255).round()).to(torch.uint8).cpu().numpy()
        img = Image.fromarray(np.stack([out] * 3, axis=2))
        return img if (w4, h4) == (w, h) else img.resize((w, h), Image.Resampling.BILINEAR)


def scribble_strokes(edge_map: np.ndarray) -> np.ndarray:
    """Soft edge map (uint8) -> binary strokes, the way the reference's `HEDdetector(..., scribble=True)` does it:
    blur, keep pixels that are the maximum along one of the four line directions through them and above 127, blur again,
    binarise at 4."""
    import cv2

    x = cv2.GaussianBlur(edge_map.astype(np.float32), (0, 0), 3.0)
    lines = [np.array(k, dtype=np.uint8) for k in ([[0, 0, 0], [1, 1, 1], [0, 0, 0]], [[0, 1, 0], [0, 1, 0], [0, 1, 0]],
                                                    [[1, 0, 0], [0, 1, 0], [0, 0, 1]], [[0, 0, 1], [0, 1, 0], [1, 0, 0]])]
    y = np.zeros_like(x)
    for k in lines:
        np.putmask(y, cv2.dilate(x, kernel=k) == x, x)
    z = np.zeros_like(y, dtype=np.uint8)
    z[y > 127] = 255
    z = cv2.GaussianBlur(z.astype(np.float32), (0, 0), 3.0)
    return np.where(z > 4, 255, 0).astype(np.uint8)


_cache = {}


def load_hed(device="cpu") -> Optional[HEDNet]:
    """The detector with its trained weights when `lllyasviel/Annotators` is available locally, else None."""
    if "hed" in _cache:
        return _cache["hed"]
    net = None
    root = checkpoints.locate_checkpoint("lllyasviel/Annotators")
    if root is not None:
        for name in ("ControlNetHED.pth", "ControlNetHED.safetensors"):
            path = Path(root) / name
            if path.is_file():
                net = HEDNet()
                net.load_state_dict(checkpoints.load_state_dict(path))
                net = net.to(device).eval()
                break
    _cache["hed"] = net
    return net


def load_lineart(device="cpu") -> Optional[LineartNet]:
    """The lineart generator with its trained weights when `lllyasviel/Annotators/sk_model.pth` is local, else None."""
    if "lineart" in _cache:
        return _cache["lineart"]
    net = None
    root = checkpoints.locate_checkpoint("lllyasviel/Annotators")
    if root is not None:
        for name in ("sk_model.pth", "sk_model.safetensors"):
            path = Path(root) / name
            if path.is_file():
                net = LineartNet()
                net.load_state_dict(checkpoints.load_state_dict(path))  # strict: a foreign file is an error, not noise
                net = net.to(device).eval()
                break
    _cache["lineart"] = net
    return net
