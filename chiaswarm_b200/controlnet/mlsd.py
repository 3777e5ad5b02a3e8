"""M-LSD straight-line annotator implemented in this tree (reference: `MLSDdetector.from_pretrained("lllyasviel/ControlNet")`
in swarm/controlnet/input_processor.py:25 -- the detector ControlNet's mlsd conditioning was trained on).

`MLSDLarge` is the "MobileV2_MLSD_Large" network of `annotator/ckpts/mlsd_large_512_fp32.pth`: a truncated MobileNetV2
(4 input channels: RGB + a plane of ones; TFLite-style bottom/right padding on the stride-2 convolutions) whose five
feature levels are merged top-down by blocks of type A (1x1 on both inputs, bilinear 2x on the deeper one, concatenate),
B (3x3 residual + 3x3) and a final C (dilated 3x3, 3x3, 1x1); the last 9 of its 16 output maps are the line map: one
centre-ness logit, four displacements (start x, y and end x, y relative to the centre) and four unused ones.  Parameter
names match the checkpoint (`backbone.features.{i}...`, `block15` .. `block23`), loading is strict.

`decode_lines`: sigmoid + 3x3 non-maximum suppression on the centre map, the 200 best centres, those above the score
threshold whose segment is longer than the distance threshold become segments, scaled by two (the map is half the input
resolution).  `draw_lines`: white 1-pixel segments on black, the conditioning image."""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..utils import checkpoints


class ConvBNReLU(nn.Sequential):
    def __init__(self, cin, cout, kernel=3, stride=1, groups=1):
        padding = 0 if stride == 2 else (kernel - 1) // 2
        super().__init__(nn.Conv2d(cin, cout, kernel, stride, padding, groups=groups, bias=False), nn.BatchNorm2d(cout),
                         nn.ReLU6(inplace=True))
        self.stride = stride

    def forward(self, x):
        if self.stride == 2:  # the weights come from a TFLite model: "same" padding there is bottom / right only
            x = F.pad(x, (0, 1, 0, 1), "constant", 0)
        for m in self:
            x = m(x)
        return x


class InvertedResidual(nn.Module):
    def __init__(self, cin, cout, stride, expand):
        super().__init__()
        hidden = int(round(cin * expand))
        self.use_res = stride == 1 and cin == cout
        layers = [] if expand == 1 else [ConvBNReLU(cin, hidden, kernel=1)]
        layers += [ConvBNReLU(hidden, hidden, stride=stride, groups=hidden), nn.Conv2d(hidden, cout, 1, 1, 0, bias=False),
                   nn.BatchNorm2d(cout)]
        self.conv = nn.Sequential(*layers)

    def forward(self, x):
        return x + self.conv(x) if self.use_res else self.conv(x)


class _Backbone(nn.Module):
    SETTING = [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1)]  # expand, channels, repeats, stride
    TAPS = (1, 3, 6, 10, 13)

    def __init__(self):
        super().__init__()
        feats = [ConvBNReLU(4, 32, stride=2)]
        cin = 32
        for t, c, n, s in self.SETTING:
            for i in range(n):
                feats.append(InvertedResidual(cin, c, s if i == 0 else 1, t))
                cin = c
        self.features = nn.Sequential(*feats)

    def forward(self, x):
        outs = []
        for i, f in enumerate(self.features):
            x = f(x)
            if i in self.TAPS:
                outs.append(x)
        return outs


def _cbr(cin, cout, kernel, **kw):
    return nn.Sequential(nn.Conv2d(cin, cout, kernel, **kw), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))


class BlockTypeA(nn.Module):
    def __init__(self, in_c1, in_c2, out_c1, out_c2, upscale=True):
        super().__init__()
        self.conv1, self.conv2, self.upscale = _cbr(in_c2, out_c2, 1), _cbr(in_c1, out_c1, 1), upscale

    def forward(self, a, b):
        b, a = self.conv1(b), self.conv2(a)
        if self.upscale:
            b = F.interpolate(b, scale_factor=2.0, mode="bilinear", align_corners=True)
        return torch.cat((a, b), dim=1)


class BlockTypeB(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.conv1, self.conv2 = _cbr(cin, cin, 3, padding=1), _cbr(cin, cout, 3, padding=1)

    def forward(self, x):
        return self.conv2(self.conv1(x) + x)


class BlockTypeC(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.conv1, self.conv2 = _cbr(cin, cin, 3, padding=5, dilation=5), _cbr(cin, cin, 3, padding=1)
        self.conv3 = nn.Conv2d(cin, cout, 1)

    def forward(self, x):
        return self.conv3(self.conv2(self.conv1(x)))


class MLSDLarge(nn.Module):
    def __init__(self):
        super().__init__()
        self.backbone = _Backbone()
        self.block15, self.block16 = BlockTypeA(64, 96, 64, 64, upscale=False), BlockTypeB(128, 64)
        self.block17, self.block18 = BlockTypeA(32, 64, 64, 64), BlockTypeB(128, 64)
        self.block19, self.block20 = BlockTypeA(24, 64, 64, 64), BlockTypeB(128, 64)
        self.block21, self.block22 = BlockTypeA(16, 64, 64, 64), BlockTypeB(128, 64)
        self.block23 = BlockTypeC(64, 16)

    def forward(self, x):
        """x: [n, 4, h, w] (RGB and ones, / 127.5 - 1).  Returns the line map [n, 9, h/2, w/2]."""
        c1, c2, c3, c4, c5 = self.backbone(x)
        x = self.block16(self.block15(c4, c5))
        x = self.block18(self.block17(c3, x))
        x = self.block20(self.block19(c2, x))
        x = self.block22(self.block21(c1, x))
        return self.block23(x)[:, 7:]


def decode_lines(tp_map: torch.Tensor, score_thr: float = 0.1, dist_thr: float = 0.1, topk: int = 200) -> np.ndarray:
    """tp_map: [9, h, w] (centre logit, 4 displacements, ...) -> segments [n, 4] = x0, y0, x1, y1 in INPUT pixels."""
    h, w = tp_map.shape[-2:]
    heat = torch.sigmoid(tp_map[0].float())
    keep = F.max_pool2d(heat[None, None], 3, 1, 1)[0, 0] == heat
    scores, idx = torch.topk((heat * keep).reshape(-1), min(topk, h * w))
    ys, xs = torch.div(idx, w, rounding_mode="floor"), idx % w
    disp = tp_map[1:5].float()[:, ys, xs].T                            # [k, 4]: start dx, dy, end dx, dy
    length = torch.sqrt(((disp[:, :2] - disp[:, 2:]) ** 2).sum(-1))
    ok = (scores > score_thr) & (length > dist_thr)
    centre = torch.stack([xs, ys, xs, ys], dim=1).float()
    return (2.0 * (centre + disp)[ok]).cpu().numpy()


def draw_lines(size, segments: np.ndarray) -> Image.Image:
    import cv2

    w, h = size
    canvas = np.zeros((h, w, 3), dtype=np.uint8)
    for x0, y0, x1, y1 in segments:
        cv2.line(canvas, (int(x0), int(y0)), (int(x1), int(y1)), (255, 255, 255), 1)
    return Image.fromarray(canvas)


class MLSDDetector:
    def __init__(self, net: MLSDLarge, device="cpu"):
        self.net, self.device = net.to(device).eval(), torch.device(device)

    @torch.no_grad()
    def __call__(self, image: Image.Image, score_thr: float = 0.1, dist_thr: float = 0.1) -> Image.Image:
        rgb = image.convert("RGB")
        w, h = rgb.size
        w64, h64 = max(64, int(round(w / 64.0)) * 64), max(64, int(round(h / 64.0)) * 64)
        arr = np.asarray(rgb.resize((w64, h64), Image.Resampling.BOX), dtype=np.float32)
        x = torch.from_numpy(np.concatenate([arr, np.ones((h64, w64, 1), np.float32)], axis=-1)).permute(2, 0, 1)[None]
        segments = decode_lines(self.net(x.to(self.device) / 127.5 - 1.0)[0], score_thr, dist_thr)
        if len(segments):
            segments = segments * np.array([w / w64, h / h64, w / w64, h / h64])
        return draw_lines((w, h), segments)


_cache = {}


def load_mlsd(device="cpu") -> Optional[MLSDDetector]:
    """The detector with its trained weights when `lllyasviel/ControlNet/annotator/ckpts/mlsd_large_512_fp32.pth` (or
    `lllyasviel/Annotators/mlsd_large_512_fp32.pth`) is available locally, else None."""
    if "mlsd" in _cache:
        return _cache["mlsd"]
    det = None
    for repo, rel in (("lllyasviel/Annotators", "mlsd_large_512_fp32.pth"),
                      ("lllyasviel/ControlNet", "annotator/ckpts/mlsd_large_512_fp32.pth")):
        root = checkpoints.locate_checkpoint(repo)
        if root is not None and (Path(root) / rel).is_file():
            net = MLSDLarge()
            net.load_state_dict(checkpoints.load_state_dict(Path(root) / rel), strict=True)
            det = MLSDDetector(net, device)
            break
    _cache["mlsd"] = det
    return det
