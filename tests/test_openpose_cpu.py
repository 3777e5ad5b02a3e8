"""In-tree OpenPose body annotator (controlnet/openpose.py): checkpoint layout of `body_pose_model.pth`, the heat-map /
part-affinity decoding on synthetic maps with a known answer, the rendering, and the pre-processor wiring."""
import math

import numpy as np
import pytest
import torch
from PIL import Image

from chiaswarm_b200.controlnet import openpose as op

# a standing figure in a 160 x 220 frame: part number (1-based, OpenPose order) -> (x, y)
FIGURE = {1: (80, 30), 2: (80, 55), 3: (60, 57), 4: (50, 90), 5: (46, 120), 6: (100, 57), 7: (110, 90), 8: (114, 120),
          9: (68, 115), 10: (66, 160), 11: (64, 200), 12: (92, 115), 13: (94, 160), 14: (96, 200), 15: (74, 24), 16: (86, 24),
          17: (66, 30), 18: (94, 30)}


def _maps(figures, size=(160, 220)):
    w, h = size
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    heat = np.zeros((19, h, w), np.float32)
    paf = np.zeros((38, h, w), np.float32)
    for fig in figures:
        for p, (x, y) in fig.items():
            if x < 0:
                continue  # (a part this figure does not have)
            heat[p - 1] = np.maximum(heat[p - 1], np.exp(-((xx - x) ** 2 + (yy - y) ** 2) / (2 * 4.0 ** 2)))
        for (a, b), (cx, cy) in zip(op.LIMBS, op.PAF_CHANNELS):
            (xa, ya), (xb, yb) = fig[a], fig[b]
            if xa < 0 or xb < 0:
                continue
            vx, vy = xb - xa, yb - ya
            n = math.hypot(vx, vy)
            vx, vy = vx / n, vy / n
            t = (xx - xa) * vx + (yy - ya) * vy                 # position along the limb
            d = np.abs((xx - xa) * vy - (yy - ya) * vx)         # distance from its axis
            on = (t >= 0) & (t <= n) & (d <= 3)
            paf[cx][on], paf[cy][on] = vx, vy
    return torch.from_numpy(heat), torch.from_numpy(paf)


def test_network_has_the_layout_of_the_original_checkpoint(tmp_path, monkeypatch):
    net = op.BodyPoseNet()
    sd = net.state_dict()
    expect = {"model0.conv1_1.weight": (64, 3, 3, 3), "model0.conv4_4_CPM.weight": (128, 256, 3, 3),
              "model1_1.conv5_5_CPM_L1.weight": (38, 512, 1, 1), "model1_2.conv5_5_CPM_L2.bias": (19,),
              "model2_1.Mconv1_stage2_L1.weight": (128, 185, 7, 7), "model6_2.Mconv7_stage6_L2.weight": (19, 128, 1, 1)}
    for k, shape in expect.items():
        assert tuple(sd[k].shape) == shape, k
    assert len(sd) == 2 * (12 + 2 * 5 + 2 * 5 * 7)
    # the original file names tensors by layer only; anything missing or foreign is refused
    short = {k.split(".", 1)[1]: v for k, v in sd.items()}
    assert "conv1_1.weight" in short and "Mconv7_stage6_L1.bias" in short
    other = op.BodyPoseNet().load_original(short)
    assert all(torch.equal(a, b) for a, b in zip(other.state_dict().values(), sd.values()))
    with pytest.raises(ValueError, match="not an OpenPose body checkpoint"):
        op.BodyPoseNet().load_original({k: v for k, v in short.items() if k != "conv3_4.bias"})
    with pytest.raises(ValueError, match="not an OpenPose body checkpoint"):
        op.BodyPoseNet().load_original(dict(short, stray=torch.zeros(1)))
    # picked up from the local model root by the pre-processor
    from chiaswarm_b200.controlnet import input_processor

    root = tmp_path / "models" / "lllyasviel" / "ControlNet" / "annotator" / "ckpts"
    root.mkdir(parents=True)
    torch.save(short, root / "body_pose_model.pth")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    op._cache.clear()
    img = Image.new("RGB", (96, 128), (120, 130, 140))
    out = input_processor.preprocess_image(img, {"preprocess": True, "type": "openpose"})
    assert op._cache["body"] is not None and out.size == img.size and out.mode == "RGB"
    op._cache.clear()


def test_decoding_recovers_a_known_figure():
    heat, paf = _maps([FIGURE])
    candidate, subset = op.decode_poses(heat, paf)
    assert len(subset) == 1 and subset[0][-1] == 18
    for p, (x, y) in FIGURE.items():
        cx, cy = candidate[int(subset[0][p - 1]), :2]
        assert abs(cx - x) <= 1 and abs(cy - y) <= 1, (p, cx, cy)


def test_two_people_are_not_cross_linked_and_weak_detections_are_dropped():
    second = {p: (x + 150, y + 6) for p, (x, y) in FIGURE.items()}
    heat, paf = _maps([FIGURE, second], size=(320, 230))
    heat[0, 100, 300] = 0.9  # a stray "nose" with no limbs: must not become a person
    candidate, subset = op.decode_poses(heat, paf)
    assert len(subset) == 2 and all(row[-1] == 18 for row in subset)
    for row in subset:
        xs = candidate[row[:18].astype(int), 0]
        assert xs.max() - xs.min() < 80  # every part of one person comes from the same figure
    # only an arm (3 parts): below the 4-part minimum
    arm = {p: FIGURE[p] for p in (2, 3, 4, 5)}
    partial = {p: arm.get(p, (-50, -50)) for p in FIGURE}
    heat, paf = _maps([partial])
    _, subset = op.decode_poses(heat, paf)
    assert len(subset) == 1 and subset[0][-1] == 4
    heat[1] = 0  # without the neck only shoulder-elbow-wrist remain: 3 parts, dropped
    _, subset = op.decode_poses(heat, paf)
    assert len(subset) == 0


def test_rendering_uses_the_training_colours():
    heat, paf = _maps([FIGURE])
    candidate, subset = op.decode_poses(heat, paf)
    img = op.draw_poses((160, 220), candidate, subset)
    arr = np.asarray(img)
    assert img.size == (160, 220) and arr[0, 0].tolist() == [0, 0, 0]
    for p, (x, y) in FIGURE.items():
        assert tuple(arr[y, x]) == op.COLORS[p - 1], p                      # joints at full brightness, drawn last
    # the middle of the right upper arm (limb 3: parts 3 -> 4) carries that limb's colour at 60 %
    (xa, ya), (xb, yb) = FIGURE[3], FIGURE[4]
    mid = arr[(ya + yb) // 2, (xa + xb) // 2]
    assert mid.tolist() == [int(c * 0.6) for c in op.COLORS[2]]
    empty = op.draw_poses((64, 48), np.zeros((0, 4)), np.zeros((0, 20)))
    assert not np.asarray(empty).any()


def test_heat_map_smoothing_is_scipy_gaussian_filter():
    """The original smooths every heat map with scipy.ndimage.gaussian_filter(sigma=3) before looking for peaks."""
    from scipy.ndimage import gaussian_filter

    rng = np.random.default_rng(0)
    for shape in ((2, 40, 50), (1, 5, 7), (1, 30, 9)):  # maps smaller than the 25-tap kernel included
        m = rng.random(shape).astype(np.float32)
        want = np.stack([gaussian_filter(x, sigma=3) for x in m])
        assert np.abs(op._gaussian_blur(torch.from_numpy(m)).numpy() - want).max() < 1e-6
