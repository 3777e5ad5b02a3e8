"""In-tree M-LSD line annotator (controlnet/mlsd.py): layout of `mlsd_large_512_fp32.pth`, line decoding on a synthetic
map with a known answer, rendering, pre-processor wiring."""
import numpy as np
import pytest
import torch
from PIL import Image

from chiaswarm_b200.controlnet import mlsd


def test_network_has_the_layout_of_the_original_checkpoint(tmp_path, monkeypatch):
    net = mlsd.MLSDLarge().eval()
    sd = net.state_dict()
    expect = {"backbone.features.0.0.weight": (32, 4, 3, 3),              # RGB + a plane of ones
              "backbone.features.1.conv.0.0.weight": (32, 1, 3, 3),        # expand ratio 1: depthwise first
              "backbone.features.1.conv.1.weight": (16, 32, 1, 1),
              "backbone.features.2.conv.0.0.weight": (96, 16, 1, 1), "backbone.features.2.conv.1.0.weight": (96, 1, 3, 3),
              "backbone.features.13.conv.2.weight": (96, 576, 1, 1),
              "block15.conv1.0.weight": (64, 96, 1, 1), "block15.conv2.0.weight": (64, 64, 1, 1),
              "block16.conv1.0.weight": (128, 128, 3, 3), "block16.conv2.0.weight": (64, 128, 3, 3),
              "block21.conv2.0.weight": (64, 16, 1, 1), "block23.conv1.0.weight": (64, 64, 3, 3),
              "block23.conv3.weight": (16, 64, 1, 1), "block23.conv3.bias": (16,)}
    for k, shape in expect.items():
        assert tuple(sd[k].shape) == shape, k
    assert "backbone.features.14.conv.0.0.weight" not in sd               # the backbone stops at 96 channels
    assert 1.4e6 < sum(v.numel() for k, v in sd.items() if "num_batches" not in k) < 1.7e6
    with torch.no_grad():
        assert net(torch.zeros(1, 4, 128, 192)).shape == (1, 9, 64, 96)   # the line map is half the input resolution
    # strict loading from the local model root; older files carry no num_batches_tracked entries
    from chiaswarm_b200.controlnet import input_processor

    root = tmp_path / "models" / "lllyasviel" / "ControlNet" / "annotator" / "ckpts"
    root.mkdir(parents=True)
    torch.save({k: v for k, v in sd.items() if "num_batches" not in k}, root / "mlsd_large_512_fp32.pth")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    mlsd._cache.clear()
    img = Image.new("RGB", (100, 70), (90, 90, 90))
    out = input_processor.preprocess_image(img, {"preprocess": True, "type": "mlsd"})
    assert mlsd._cache["mlsd"] is not None and out.size == img.size and out.mode == "RGB"
    mlsd._cache.clear()
    torch.save({"backbone.features.0.0.weight": torch.zeros(32, 4, 3, 3)}, root / "mlsd_large_512_fp32.pth")
    with pytest.raises(RuntimeError):
        mlsd.load_mlsd()
    mlsd._cache.clear()


def test_decoding_recovers_known_segments():
    tp = torch.full((9, 64, 96), -8.0)
    tp[1:5] = 0
    # a centre at (x 30, y 20) whose segment runs from (20, 15) to (45, 25), and one at (70, 50): (60, 50) -> (80, 50)
    for (x, y), (x0, y0, x1, y1), logit in (((30, 20), (20, 15, 45, 25), 4.0), ((70, 50), (60, 50, 80, 50), 2.0)):
        tp[0, y, x] = logit
        tp[0, y, x + 1] = logit - 1.0                                   # a weaker neighbour: suppressed by the 3x3 NMS
        tp[1:5, y, x] = torch.tensor([x0 - x, y0 - y, x1 - x, y1 - y], dtype=torch.float32)
    tp[0, 5, 5] = 3.0                                                    # a centre with a zero-length segment: dropped
    tp[0, 40, 10] = -3.0                                                 # below the score threshold (sigmoid = 0.047)
    tp[1:5, 40, 10] = torch.tensor([-5.0, 0.0, 5.0, 0.0])
    segs = mlsd.decode_lines(tp)
    assert segs.shape == (2, 4)
    assert segs[0].tolist() == [40.0, 30.0, 90.0, 50.0] and segs[1].tolist() == [120.0, 100.0, 160.0, 100.0]  # x2: input px
    img = mlsd.draw_lines((192, 128), segs)
    arr = np.asarray(img)
    assert arr[100, 140].tolist() == [255, 255, 255] and arr[40, 65].tolist() == [255, 255, 255] and arr[0, 0].tolist() == [0, 0, 0]
    assert not np.asarray(mlsd.draw_lines((32, 32), np.zeros((0, 4)))).any()
