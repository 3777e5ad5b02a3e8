"""In-tree PiDiNet soft-edge annotator (controlnet/pidinet.py): the pixel-difference convolutions against their
definitions, the layout of `table5_pidinet.pth`, pre-processor wiring."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F
from PIL import Image

from chiaswarm_b200.controlnet import pidinet

# offsets (dy, dx) of the nine taps of a 3x3 kernel in storage order
TAPS = [(dy, dx) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]


def _shift(x, dy, dx):
    """x[p + (dy, dx)] with zeros outside."""
    h, w = x.shape[-2:]
    pad = 2
    xp = F.pad(x, (pad,) * 4)
    return xp[..., pad + dy:pad + dy + h, pad + dx:pad + dx + w]


@pytest.mark.parametrize("op", ["cv", "cd", "ad", "rd"])
def test_pixel_difference_convolutions_are_what_they_say(op):
    torch.manual_seed(0)
    conv = pidinet.PDConv2d(op, 4, 4, groups=4).double()
    x = torch.randn(2, 4, 7, 9, dtype=torch.float64)
    w = conv.weight.detach()                                   # [4, 1, 3, 3], depthwise
    want = torch.zeros_like(x)
    for t, (dy, dx) in enumerate(TAPS):
        wt = w[:, 0, t // 3, t % 3].view(1, 4, 1, 1)
        if op == "cv":
            want += wt * _shift(x, dy, dx)
        elif op == "cd":                                       # every tap sees the difference to the centre pixel
            want += wt * (_shift(x, dy, dx) - x)
        elif op == "ad":                                       # ... to the next pixel round the ring (centre: to itself)
            dy2, dx2 = TAPS[pidinet._CLOCKWISE.index(t)]
            want += wt * (_shift(x, dy, dx) - _shift(x, dy2, dx2))
        elif op == "rd" and t != 4:                            # outer ring at twice the offset minus the inner ring
            want += wt * (_shift(x, 2 * dy, 2 * dx) - _shift(x, dy, dx))
    if op == "rd":
        # taps are numbered 1..8 in the rewrite (the centre weight is dropped): weights[:, :, 1:] fill the rings in storage
        # order of the 5x5 buffer, so tap k of the 3x3 kernel (k != 0) maps to ring position k - 1
        want = torch.zeros_like(x)
        ring = [(-1, -1), (-1, 0), (-1, 1), (0, -1), (0, 1), (1, -1), (1, 0), (1, 1)]
        flat = w.view(4, 1, 9)
        for k, (dy, dx) in enumerate(ring):
            wk = flat[:, 0, k + 1].view(1, 4, 1, 1)
            want += wk * (_shift(x, 2 * dy, 2 * dx) - _shift(x, dy, dx))
    assert torch.allclose(conv(x), want, atol=1e-10), op
    if op != "cv":  # a constant image has no differences: every difference convolution answers zero away from the border
        flat_img = torch.ones(1, 4, 9, 9, dtype=torch.float64)
        assert conv(flat_img)[..., 2:-2, 2:-2].abs().max() < 1e-12


def test_network_has_the_layout_of_the_original_checkpoint(tmp_path, monkeypatch):
    net = pidinet.PiDiNet().eval()
    sd = net.state_dict()
    expect = {"init_block.weight": (60, 3, 3, 3), "block1_1.conv1.weight": (60, 1, 3, 3), "block1_1.conv2.weight": (60, 60, 1, 1),
              "block2_1.shortcut.weight": (120, 60, 1, 1), "block2_1.shortcut.bias": (120,), "block2_1.conv1.weight": (60, 1, 3, 3),
              "block2_1.conv2.weight": (120, 60, 1, 1), "block3_4.conv1.weight": (240, 1, 3, 3),
              "block4_1.shortcut.weight": (240, 240, 1, 1), "dilations.0.conv1.weight": (24, 60, 1, 1),
              "dilations.3.conv2_4.weight": (24, 24, 3, 3), "attentions.1.conv1.weight": (4, 24, 1, 1),
              "attentions.1.conv2.weight": (1, 4, 3, 3), "conv_reduces.2.conv.weight": (1, 24, 1, 1),
              "classifier.weight": (1, 4, 1, 1), "classifier.bias": (1,)}
    for k, shape in expect.items():
        assert tuple(sd[k].shape) == shape, k
    for absent in ("block1_1.shortcut.weight", "block1_4.conv1.weight", "block1_1.conv2.bias", "init_block.bias",
                   "dilations.0.conv2_1.bias", "attentions.0.conv2.bias"):
        assert absent not in sd, absent
    assert [net.init_block.op] + [getattr(net, n).conv1.op for names in net.stages for n in names] == pidinet.CARV4
    assert 0.6e6 < sum(v.numel() for v in sd.values()) < 0.8e6            # "table5": about 0.7 M parameters
    with torch.no_grad():
        outs = net(torch.rand(1, 3, 40, 56))
    assert len(outs) == 5 and all(o.shape == (1, 1, 40, 56) and 0 <= float(o.min()) and float(o.max()) <= 1 for o in outs)
    # the original file wraps the tensors in {"state_dict": ...} under DataParallel names; picked up by the pre-processor
    from chiaswarm_b200.controlnet import annotators, input_processor

    root = tmp_path / "models" / "lllyasviel" / "Annotators"
    root.mkdir(parents=True)
    torch.save({"state_dict": {"module." + k: v for k, v in sd.items()}, "epoch": 19}, root / "table5_pidinet.pth")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    pidinet._cache.clear()
    annotators._cache.clear()
    img = Image.radial_gradient("L").resize((72, 48)).convert("RGB")
    soft = input_processor.preprocess_image(img, {"preprocess": True, "type": "softedge"})
    assert pidinet._cache["pidi"] is not None and soft.size == img.size and soft.mode == "RGB"
    assert np.asarray(soft)[..., 0].tolist() == np.asarray(soft)[..., 2].tolist()
    pidinet._cache.clear()
    torch.save({"state_dict": {"module.init_block.weight": torch.zeros(60, 3, 3, 3)}}, root / "table5_pidinet.pth")
    with pytest.raises(RuntimeError):
        pidinet.load_pidinet()
    pidinet._cache.clear()
    annotators._cache.clear()
