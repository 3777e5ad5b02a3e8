"""In-tree NormalBae annotator (controlnet/normalbae.py): the EfficientNet-B5 encoder's layout and TensorFlow padding, the
decoder's refinement path, the layout of `scannet.pt`, pre-processor wiring."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F
from PIL import Image

from chiaswarm_b200.controlnet import normalbae as nb


def test_network_has_the_layout_of_the_original_checkpoint():
    with torch.device("meta"):
        net = nb.NNET()
    sd = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    expect = {
        "encoder.original_model.conv_stem.weight": (48, 3, 3, 3), "encoder.original_model.bn1.running_var": (48,),
        "encoder.original_model.blocks.0.0.conv_dw.weight": (48, 1, 3, 3),         # stage 0: depthwise-separable, no expansion
        "encoder.original_model.blocks.0.0.se.conv_reduce.weight": (12, 48, 1, 1),
        "encoder.original_model.blocks.0.0.conv_pw.weight": (24, 48, 1, 1),
        "encoder.original_model.blocks.1.0.conv_pw.weight": (144, 24, 1, 1),       # expand 6
        "encoder.original_model.blocks.1.0.se.conv_reduce.weight": (6, 144, 1, 1),  # SE width from the block INPUT (24 / 4)
        "encoder.original_model.blocks.2.0.conv_dw.weight": (240, 1, 5, 5),
        "encoder.original_model.blocks.4.6.conv_pwl.weight": (176, 1056, 1, 1),
        "encoder.original_model.blocks.5.8.bn3.weight": (304,),
        "encoder.original_model.blocks.6.2.conv_pwl.weight": (512, 3072, 1, 1),
        "encoder.original_model.conv_head.weight": (2048, 512, 1, 1), "encoder.original_model.bn2.bias": (2048,),
        "decoder.conv2.weight": (2048, 2048, 1, 1), "decoder.up1._net.0.weight": (1024, 2224, 3, 3),
        "decoder.up2._net.3.weight": (512, 512, 3, 3), "decoder.up4._net.0.weight": (128, 280, 3, 3),
        "decoder.out_conv_res8.weight": (4, 512, 3, 3), "decoder.out_conv_res4.0.weight": (128, 516, 1),
        "decoder.out_conv_res2.0.weight": (128, 260, 1), "decoder.out_conv_res1.6.weight": (4, 128, 1),
    }
    for k, shape in expect.items():
        assert sd.get(k) == shape, (k, sd.get(k))
    assert [len(s) for s in net.encoder.original_model.blocks] == [3, 5, 5, 7, 7, 9, 3]
    count = lambda prefix: sum(math.prod(v) for k, v in sd.items() if k.startswith(prefix) and "num_batches" not in k)  # noqa: E731
    assert 28.2e6 < count("encoder") < 28.8e6       # EfficientNet-B5 without its classifier
    for absent in ("encoder.original_model.classifier.weight", "encoder.original_model.blocks.0.0.conv_pwl.weight",
                   "encoder.original_model.blocks.7.0.conv_pw.weight", "decoder.up5._net.0.weight"):
        assert absent not in sd


@pytest.mark.parametrize("size,kernel,stride", [(7, 3, 2), (8, 3, 2), (9, 5, 2), (10, 5, 1), (5, 1, 1)])
def test_same_padding_is_tensorflows(size, kernel, stride):
    """Output ceil(n / s); total padding (ceil(n/s) - 1) s + k - n, the odd pixel after the data."""
    conv = nb.Conv2dSame(1, 1, kernel, stride, bias=False)
    torch.nn.init.ones_(conv.weight)
    x = torch.arange(size * size, dtype=torch.float32).view(1, 1, size, size)
    y = conv(x)
    out = math.ceil(size / stride)
    assert y.shape[-2:] == (out, out)
    total = max((out - 1) * stride + kernel - size, 0)
    before, after = total // 2, total - total // 2
    want = F.conv2d(F.pad(x, [before, after, before, after]), conv.weight, stride=stride)
    assert torch.equal(y, want) and after - before in (0, 1)


def test_predictions_are_unit_normals_at_four_scales():
    torch.manual_seed(0)
    net = nb.NNET().eval()
    with torch.no_grad():
        outs = net(torch.randn(1, 3, 64, 96))
    assert [tuple(o.shape) for o in outs] == [(1, 4, 8, 12), (1, 4, 16, 24), (1, 4, 32, 48), (1, 4, 64, 96)]
    for o in outs:
        assert torch.allclose((o[:, :3] ** 2).sum(1).sqrt(), torch.ones_like(o[:, 0]), atol=1e-4)
        assert float(o[:, 3].min()) > 0.01 - 1e-6                                   # kappa = elu + 1 + 0.01 > 0.01
    raw = torch.tensor([[[3.0]], [[0.0]], [[4.0]], [[-50.0]]])[None]
    n = nb.norm_normalize(raw)[0, :, 0, 0]
    assert torch.allclose(n[:3], torch.tensor([0.6, 0.0, 0.8]), atol=1e-6) and float(n[3]) == pytest.approx(0.01, abs=1e-6)
    img = net.normals(Image.radial_gradient("L").resize((70, 50)).convert("RGB"))
    assert img.size == (70, 50) and img.mode == "RGB"


def test_preprocessor_uses_the_network_when_its_checkpoint_is_local(monkeypatch):
    from chiaswarm_b200.controlnet import input_processor

    class Stub(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.p = torch.nn.Parameter(torch.zeros(1))

        def normals(self, image, device):
            return Image.new("RGB", image.size, (128, 128, 255))

    monkeypatch.setattr(nb, "load_normalbae", lambda device="cpu": Stub())
    img = Image.new("RGB", (40, 24), (9, 9, 9))
    out = input_processor.preprocess_image(img, {"preprocess": True, "type": "normalbae"})
    assert out.getpixel((0, 0)) == (128, 128, 255) and "controlnet_annotator" not in out.info
    monkeypatch.setattr(nb, "load_normalbae", lambda device="cpu": None)
    out = input_processor.preprocess_image(img, {"preprocess": True, "type": "normalbae"})
    assert out.info["controlnet_annotator"].startswith("stand-in for normalbae")


def test_loader_unwraps_the_original_file_layout(tmp_path, monkeypatch):
    """`scannet.pt` = {"model": {"module.<name>": tensor}}; anything else is refused.  (A tiny stand-in class keeps the
    file small: the unwrapping and the strictness are what is tested.)"""
    class Tiny(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.encoder = torch.nn.Linear(2, 2)

    monkeypatch.setattr(nb, "NNET", Tiny)
    root = tmp_path / "models" / "lllyasviel" / "Annotators"
    root.mkdir(parents=True)
    good = Tiny()
    torch.save({"model": {"module." + k: v for k, v in good.state_dict().items()}, "iter": 5}, root / "scannet.pt")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    nb._cache.clear()
    net = nb.load_normalbae()
    assert isinstance(net, Tiny) and torch.equal(net.encoder.weight, good.encoder.weight)
    nb._cache.clear()
    torch.save({"model": {"module.encoder.weight": torch.zeros(2, 2)}}, root / "scannet.pt")
    with pytest.raises(RuntimeError):
        nb.load_normalbae()
    nb._cache.clear()
