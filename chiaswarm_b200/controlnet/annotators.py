"""Learned ControlNet annotators implemented in this tree (reference: `controlnet_aux` detectors loaded in
swarm/controlnet/input_processor.py:100-231).

`HEDNet` is the holistically-nested edge detector that ControlNet's soft-edge / scribble conditioning was trained on
(`lllyasviel/Annotators/ControlNetHED.pth`): five VGG-style stages, each with a 1x1 side output; the edge map is the
sigmoid of the mean of the up-sampled side outputs.  Parameter names match that checkpoint (`norm`,
`block{1..5}.convs.{i}`, `block{n}.projection`), so the file loads with `load_state_dict` unchanged.  It runs on
whatever device the job has (a few hundred MFLOP per image: far off the hot path) and is used whenever the checkpoint
is present locally; without it the classical stand-in in input_processor.py is used."""
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
        if scribble:  # thin the soft map to strokes: local maxima above a threshold, then a small dilation
            e = edge[None, None]
            peak = (e >= F.max_pool2d(e, 3, 1, 1) - 1e-6) & (e > 0.25)
            edge = F.max_pool2d(peak.float(), 3, 1, 1)[0, 0]
        out = (edge.clamp(0, 1) * 255).round().to(torch.uint8).cpu().numpy()
        return Image.fromarray(np.stack([out] * 3, axis=2))


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
