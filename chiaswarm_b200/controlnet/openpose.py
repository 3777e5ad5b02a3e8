"""OpenPose body annotator implemented in this tree (reference: `OpenposeDetector.from_pretrained("lllyasviel/ControlNet")`
in swarm/controlnet/input_processor.py -- the detector ControlNet's openpose conditioning was trained on).

Three parts, as in the original CMU / pytorch-openpose pipeline the `controlnet_aux` detector wraps:

  `BodyPoseNet`   the two-branch, six-stage convolutional pose machine: ten VGG-19 layers + two CPM layers, then per stage a
                  part-affinity-field branch (38 maps = 19 limbs x (x, y)) and a heat-map branch (18 body parts + background),
                  each later stage reading [PAF, heat maps, backbone features] (185 channels) through 7x7 convolutions.
                  Parameter names are those of `annotator/ckpts/body_pose_model.pth` with the owning sub-module in front
                  (`model0.conv1_1`, `model1_1.conv5_1_CPM_L1`, `model3_2.Mconv4_stage3_L2`, ...): the checkpoint stores them
                  without that prefix, `load_body_pose` puts it back and refuses anything missing or left over.
  `decode_poses`  heat maps + PAFs -> people: smoothed local maxima per part, candidate limbs scored by the line integral of
                  the PAF along the segment (10 samples, 80 % above 0.05, a prior against limbs longer than half the image),
                  greedy one-to-one matching per limb type, limbs merged into people, people with fewer than 4 parts or a
                  mean score under 0.4 dropped.
  `draw_poses`    the 17 coloured limbs (filled ellipses, dimmed to 60 %) and 18 joints on black: the conditioning image.

Runs on whatever device the job has (one 368-pixel-high pass: a few GFLOP, far off the hot path)."""
from __future__ import annotations

import math
from collections import OrderedDict
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..utils import checkpoints

# limb k joins parts LIMBS[k] (1-based part numbers) and its PAF is the channel pair PAF_CHANNELS[k] of the 38-map branch
LIMBS = [(2, 3), (2, 6), (3, 4), (4, 5), (6, 7), (7, 8), (2, 9), (9, 10), (10, 11), (2, 12), (12, 13), (13, 14), (2, 1), (1, 15),
         (15, 17), (1, 16), (16, 18), (3, 17), (6, 18)]
PAF_CHANNELS = [(12, 13), (20, 21), (14, 15), (16, 17), (22, 23), (24, 25), (0, 1), (2, 3), (4, 5), (6, 7), (8, 9), (10, 11),
                (28, 29), (30, 31), (34, 35), (32, 33), (36, 37), (18, 19), (26, 27)]
COLORS = [(255, 0, 0), (255, 85, 0), (255, 170, 0), (255, 255, 0), (170, 255, 0), (85, 255, 0), (0, 255, 0), (0, 255, 85),
          (0, 255, 170), (0, 255, 255), (0, 170, 255), (0, 85, 255), (0, 0, 255), (85, 0, 255), (170, 0, 255), (255, 0, 255),
          (255, 0, 170), (255, 0, 85)]
BOX, STRIDE, PAD_VALUE = 368, 8, 128
PEAK_THRESHOLD, LIMB_THRESHOLD, LIMB_SAMPLES = 0.1, 0.05, 10


def _sequential(spec, no_relu=()):
    layers = OrderedDict()
    for name, v in spec:
        if name.startswith("pool"):
            layers[name] = nn.MaxPool2d(kernel_size=v[0], stride=v[1], padding=v[2])
        else:
            layers[name] = nn.Conv2d(v[0], v[1], kernel_size=v[2], stride=v[3], padding=v[4])
            if name not in no_relu:
                layers["relu_" + name] = nn.ReLU(inplace=True)
    return nn.Sequential(layers)


class BodyPoseNet(nn.Module):
    def __init__(self):
        super().__init__()
        c3 = lambda i, o: [i, o, 3, 1, 1]  # noqa: E731
        self.model0 = _sequential([
            ("conv1_1", c3(3, 64)), ("conv1_2", c3(64, 64)), ("pool1_stage1", [2, 2, 0]),
            ("conv2_1", c3(64, 128)), ("conv2_2", c3(128, 128)), ("pool2_stage1", [2, 2, 0]),
            ("conv3_1", c3(128, 256)), ("conv3_2", c3(256, 256)), ("conv3_3", c3(256, 256)), ("conv3_4", c3(256, 256)),
            ("pool3_stage1", [2, 2, 0]), ("conv4_1", c3(256, 512)), ("conv4_2", c3(512, 512)),
            ("conv4_3_CPM", c3(512, 256)), ("conv4_4_CPM", c3(256, 128))])
        for branch, outs in ((1, 38), (2, 19)):
            last = f"conv5_5_CPM_L{branch}"
            setattr(self, f"model1_{branch}", _sequential(
                [(f"conv5_{i}_CPM_L{branch}", c3(128, 128)) for i in (1, 2, 3)]
                + [(f"conv5_4_CPM_L{branch}", [128, 512, 1, 1, 0]), (last, [512, outs, 1, 1, 0])], no_relu=(last,)))
            for stage in range(2, 7):
                last = f"Mconv7_stage{stage}_L{branch}"
                setattr(self, f"model{stage}_{branch}", _sequential(
                    [(f"Mconv1_stage{stage}_L{branch}", [185, 128, 7, 1, 3])]
                    + [(f"Mconv{i}_stage{stage}_L{branch}", [128, 128, 7, 1, 3]) for i in (2, 3, 4, 5)]
                    + [(f"Mconv6_stage{stage}_L{branch}", [128, 128, 1, 1, 0]), (last, [128, outs, 1, 1, 0])], no_relu=(last,)))

    def forward(self, x):
        """x: [n, 3, h, w], BGR, pixel / 256 - 0.5.  Returns (PAFs [n, 38, h/8, w/8], heat maps [n, 19, h/8, w/8])."""
        feat = self.model0(x)
        paf, heat = self.model1_1(feat), self.model1_2(feat)
        for stage in range(2, 7):
            y = torch.cat([paf, heat, feat], dim=1)
            paf, heat = getattr(self, f"model{stage}_1")(y), getattr(self, f"model{stage}_2")(y)
        return paf, heat

    def load_original(self, sd: dict):
        """`body_pose_model.pth` names its tensors by layer only (`conv1_1.weight`); put the sub-module back in front."""
        own = self.state_dict()
        mapped, missing = {}, []
        for key in own:
            short = key.split(".", 1)[1]
            src = sd.get(short, sd.get(key))
            if src is None:
                missing.append(short)
            else:
                mapped[key] = src
        used = {k.split(".", 1)[1] for k in own} | set(own)
        extra = [k for k in sd if k not in used]
        if missing or extra:
            raise ValueError(f"not an OpenPose body checkpoint: missing {missing[:4]}, unexpected {extra[:4]}")
        self.load_state_dict(mapped, strict=True)
        return self


# ------------------------------------------------------------------------------------------ maps -> people
def _symmetric_pad(x: torch.Tensor, r: int, dim: int) -> torch.Tensor:
    """scipy's default border ("reflect": d c b a | a b c d -- the edge sample repeats), r may exceed the size."""
    n = x.shape[dim]
    idx = torch.arange(-r, n + r, device=x.device)
    idx = idx % (2 * n)
    idx = torch.where(idx >= n, 2 * n - 1 - idx, idx)
    return x.index_select(dim, idx)


def _gaussian_blur(maps: torch.Tensor, sigma: float = 3.0) -> torch.Tensor:
    """[c, h, w] -> smoothed exactly as scipy.ndimage.gaussian_filter does (kernel truncated at 4 sigma, its borders)."""
    r = int(4 * sigma + 0.5)
    k = torch.exp(-0.5 * (torch.arange(-r, r + 1, dtype=maps.dtype, device=maps.device) / sigma) ** 2)
    k = k / k.sum()
    x = maps[:, None]
    x = F.conv2d(_symmetric_pad(x, r, 3), k.view(1, 1, 1, -1))
    x = F.conv2d(_symmetric_pad(x, r, 2), k.view(1, 1, -1, 1))
    return x[:, 0]


def find_peaks(heat: torch.Tensor, threshold: float = PEAK_THRESHOLD) -> List[List[Tuple[int, int, float, int]]]:
    """heat: [18+, h, w] at image resolution.  Per part: [(x, y, score, peak id)], ids running over all parts."""
    parts = heat[:18]
    smooth = _gaussian_blur(parts)
    big = torch.full_like(smooth, -1e9)
    left, right, up, down = big.clone(), big.clone(), big.clone(), big.clone()
    left[:, :, 1:], right[:, :, :-1] = smooth[:, :, :-1], smooth[:, :, 1:]
    up[:, 1:, :], down[:, :-1, :] = smooth[:, :-1, :], smooth[:, 1:, :]
    # (borders compare against zero in the original, which pads the shifted maps with zeros)
    left[:, :, 0] = right[:, :, -1] = 0
    up[:, 0, :] = down[:, -1, :] = 0
    peak = (smooth >= left) & (smooth >= right) & (smooth >= up) & (smooth >= down) & (smooth > threshold)
    out, next_id = [], 0
    raw = parts.cpu().numpy()
    for p, mask in enumerate(peak.cpu().numpy()):
        ys, xs = np.nonzero(mask)  # row-major, as the original enumerates them
        found = []
        for x, y in zip(xs, ys):
            found.append((int(x), int(y), float(raw[p, y, x]), next_id))
            next_id += 1
        out.append(found)
    return out


def decode_poses(heat: torch.Tensor, paf: torch.Tensor):
    """heat [19, h, w], paf [38, h, w] at image resolution -> (candidate [n, 4] = x, y, score, id; subset [people, 20] =
    candidate id per part (-1 = missing), total score, part count)."""
    h = heat.shape[1]
    peaks = find_peaks(heat)
    pafs = paf.cpu().numpy()
    connections, special = [], []
    for k, ((a, b), (cx, cy)) in enumerate(zip(LIMBS, PAF_CHANNELS)):
        cand_a, cand_b = peaks[a - 1], peaks[b - 1]
        if not cand_a or not cand_b:
            special.append(k)
            connections.append(np.zeros((0, 5)))
            continue
        scored = []
        for i, pa in enumerate(cand_a):
            for j, pb in enumerate(cand_b):
                vec = np.array([pb[0] - pa[0], pb[1] - pa[1]], dtype=np.float64)
                norm = max(0.001, math.sqrt(float(vec @ vec)))
                vec /= norm
                xs = np.linspace(pa[0], pb[0], LIMB_SAMPLES)
                ys = np.linspace(pa[1], pb[1], LIMB_SAMPLES)
                xi, yi = np.round(xs).astype(int), np.round(ys).astype(int)
                along = pafs[cx, yi, xi] * vec[0] + pafs[cy, yi, xi] * vec[1]
                prior = along.mean() + min(0.5 * h / norm - 1, 0)
                if np.count_nonzero(along > LIMB_THRESHOLD) > 0.8 * LIMB_SAMPLES and prior > 0:
                    scored.append((i, j, prior, prior + pa[2] + pb[2]))
        scored.sort(key=lambda c: c[2], reverse=True)
        rows, used_a, used_b = [], set(), set()
        for i, j, s, _ in scored:
            if i not in used_a and j not in used_b:
                rows.append([cand_a[i][3], cand_b[j][3], s, i, j])
                used_a.add(i)
                used_b.add(j)
                if len(rows) >= min(len(cand_a), len(cand_b)):
                    break
        connections.append(np.array(rows, dtype=np.float64).reshape(-1, 5))

    candidate = np.array([p for part in peaks for p in part], dtype=np.float64).reshape(-1, 4)
    subset = -1 * np.ones((0, 20))
    for k, (a, b) in enumerate(LIMBS):
        if k in special:
            continue
        ia, ib = a - 1, b - 1
        for row in connections[k]:
            pa, pb, score = int(row[0]), int(row[1]), row[2]
            found = [j for j in range(len(subset)) if subset[j][ia] == pa or subset[j][ib] == pb][:2]
            if len(found) == 1:
                j = found[0]
                if subset[j][ib] != pb:
                    subset[j][ib] = pb
                    subset[j][-1] += 1
                    subset[j][-2] += candidate[pb, 2] + score
            elif len(found) == 2:
                j1, j2 = found
                both = ((subset[j1] >= 0).astype(int) + (subset[j2] >= 0).astype(int))[:-2]
                if not np.any(both == 2):  # two partial people that share no part: one person
                    subset[j1][:-2] += subset[j2][:-2] + 1
                    subset[j1][-2:] += subset[j2][-2:]
                    subset[j1][-2] += score
                    subset = np.delete(subset, j2, 0)
                else:
                    subset[j1][ib] = pb
                    subset[j1][-1] += 1
                    subset[j1][-2] += candidate[pb, 2] + score
            elif not found and k < 17:
                row20 = -1 * np.ones(20)
                row20[ia], row20[ib] = pa, pb
                row20[-1] = 2
                row20[-2] = candidate[pa, 2] + candidate[pb, 2] + score
                subset = np.vstack([subset, row20])
    keep = [i for i in range(len(subset)) if subset[i][-1] >= 4 and subset[i][-2] / subset[i][-1] >= 0.4]
    return candidate, subset[keep]


def draw_poses(size: Tuple[int, int], candidate: np.ndarray, subset: np.ndarray, stick: int = 4) -> Image.Image:
    import cv2

    w, h = size
    canvas = np.zeros((h, w, 3), dtype=np.uint8)
    for k in range(17):
        for person in subset:
            idx = person[np.array(LIMBS[k]) - 1]
            if -1 in idx:
                continue
            xs, ys = candidate[idx.astype(int), 0], candidate[idx.astype(int), 1]
            length = math.hypot(xs[0] - xs[1], ys[0] - ys[1])
            angle = math.degrees(math.atan2(ys[0] - ys[1], xs[0] - xs[1]))
            poly = cv2.ellipse2Poly((int(xs.mean()), int(ys.mean())), (int(length / 2), stick), int(angle), 0, 360, 1)
            cv2.fillConvexPoly(canvas, poly, COLORS[k])
    canvas = (canvas * 0.6).astype(np.uint8)
    for p in range(18):
        for person in subset:
            i = int(person[p])
            if i >= 0:
                cv2.circle(canvas, (int(candidate[i, 0]), int(candidate[i, 1])), 4, COLORS[p], thickness=-1)
    return Image.fromarray(canvas)


# ------------------------------------------------------------------------------------------ the detector
class OpenposeBody:
    def __init__(self, net: BodyPoseNet, device="cpu"):
        self.net, self.device = net.to(device).eval(), torch.device(device)

    @torch.no_grad()
    def maps(self, image: Image.Image):
        """One pass at scale 0.5 * 368 / height (the detector's default search): (heat [19, h, w], paf [38, h, w])."""
        rgb = np.asarray(image.convert("RGB"), dtype=np.float32)
        h, w = rgb.shape[:2]
        scale = 0.5 * BOX / h
        x = torch.from_numpy(rgb[:, :, ::-1].copy()).permute(2, 0, 1)[None].to(self.device)  # BGR, as the weights expect
        hs, ws = max(STRIDE, int(round(h * scale))), max(STRIDE, int(round(w * scale)))
        x = F.interpolate(x, size=(hs, ws), mode="bicubic", align_corners=False)
        ph, pw = (-hs) % STRIDE, (-ws) % STRIDE
        x = F.pad(x, (0, pw, 0, ph), value=PAD_VALUE) / 256.0 - 0.5
        paf, heat = self.net(x)

        def back(m):
            m = F.interpolate(m.float(), scale_factor=STRIDE, mode="bicubic", align_corners=False)[:, :, :hs, :ws]
            return F.interpolate(m, size=(h, w), mode="bicubic", align_corners=False)[0]

        return back(heat), back(paf)

    def __call__(self, image: Image.Image) -> Image.Image:
        heat, paf = self.maps(image)
        candidate, subset = decode_poses(heat, paf)
        return draw_poses(image.size, candidate, subset)


_cache = {}


def load_body_pose(device="cpu") -> Optional[OpenposeBody]:
    """The detector with its trained weights when `lllyasviel/ControlNet/annotator/ckpts/body_pose_model.pth` (or
    `lllyasviel/Annotators/body_pose_model.pth`) is available locally, else None."""
    if "body" in _cache:
        return _cache["body"]
    det = None
    for repo, rel in (("lllyasviel/Annotators", "body_pose_model.pth"),
                      ("lllyasviel/ControlNet", "annotator/ckpts/body_pose_model.pth")):
        root = checkpoints.locate_checkpoint(repo)
        if root is not None and (Path(root) / rel).is_file():
            det = OpenposeBody(BodyPoseNet().load_original(checkpoints.load_state_dict(Path(root) / rel)), device)
            break
    _cache["body"] = det
    return det
