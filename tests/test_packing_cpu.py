"""Host-side weight packing and the tuning-table plumbing (no GPU needed)."""
import json

import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import packing, tuning


def test_fold_layernorm_is_an_exact_rewrite():
    torch.manual_seed(0)
    rows, c, n = 37, 64, 48
    x = torch.randn(rows, c) * 1.7 + 0.4
    w, b = torch.randn(n, c) * c ** -0.5, torch.randn(n)
    gamma, beta = torch.randn(c) * 0.3 + 1, torch.randn(c) * 0.2
    w2, b2, cs = packing.fold_layernorm(w.half(), b.half(), gamma, beta)
    assert w2.dtype == torch.float16 and cs.dtype == torch.float32 and cs.shape == (n,)
    # what the consumer GEMM epilogue computes from the producer's (sum, sum of squares)
    mean = x.mean(1, keepdim=True)
    rstd = (x.var(1, unbiased=False, keepdim=True) + 1e-5).rsqrt()
    got = rstd * (x @ w2.float().t() - mean * cs[None]) + b2.float()[None]
    ref = F.layer_norm(x, (c,), gamma, beta, 1e-5) @ w.half().float().t() + b.half().float()
    assert (got - ref).abs().max() < 2e-2
    # column sums come from the ROUNDED weights: that is what the tensor core accumulates
    assert torch.equal(cs, w2.float().sum(1))


def test_interleave_geglu_granules():
    w = torch.arange(512 * 8, dtype=torch.float32).reshape(512, 8).half()
    b = torch.arange(512, dtype=torch.float32)
    for half in (64, 128):
        wi, bi = packing.interleave_geglu(w, b, half)
        assert wi.shape == w.shape and bi.shape == b.shape
        # tile t holds `half` value rows followed by their `half` gate rows
        for t in range(256 // half):
            assert torch.equal(wi[2 * half * t:2 * half * t + half], w[half * t:half * (t + 1)])
            assert torch.equal(wi[2 * half * t + half:2 * half * (t + 1)], w[256 + half * t:256 + half * (t + 1)])
            assert torch.equal(bi[2 * half * t + half:2 * half * (t + 1)], b[256 + half * t:256 + half * (t + 1)])


def test_tuning_table_lookup_and_formats(tmp_path, monkeypatch):
    key = tuning.key_of(1, 1, 512, 1, 1, 1280, 0, 1280, 0, True)
    assert key == "1,1,512,1,1,1280,0,1280,0,1"
    path = tmp_path / "t.json"
    path.write_text(json.dumps({"device": "x", "configs": {key: [64, 1, 16], "k4": [128, 2, 8, 2], "k5": [256, 1, 64, 1, 2]}}))
    monkeypatch.setattr(tuning, "_PATH", path)
    monkeypatch.setattr(tuning, "_table", None)
    assert tuning.lookup(1, 1, 512, 1, 1, 1280, 0, 1280, 0, True) == (64, 1, 16, 0, 0)   # older 3-field entries
    assert tuning.table()["k4"] == (128, 2, 8, 2, 0) and tuning.table()["k5"] == (256, 1, 64, 1, 2)
    assert tuning.lookup(2, 8, 8, 3, 3, 1280, 0, 1280, 0, False) is None
    monkeypatch.setattr(tuning, "_table", None)
    monkeypatch.setenv("SW_IGEMM_TUNING", "0")
    assert tuning.table() == {}


def test_shipped_tuning_table_is_well_formed():
    raw = json.loads(tuning._PATH.read_text())
    assert len(raw["configs"]) > 300
    for k, v in raw["configs"].items():
        assert len(k.split(",")) == 10 and 3 <= len(v) <= 5
        assert v[0] in (0, 32, 64, 96, 128, 160, 192, 224, 256) and v[1] in (0, 1, 2)
