"""Numerics of every sm_100a kernel against a plain PyTorch fp32 restatement of the same op
(exact UNet shapes plus odd ones -- head_dim 40, S=45, partial tiles)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope="module")
def nat():
    from chiaswarm_b200.ops import native

    native.lib()
    return native


def _rel_err(got, ref):
    ref = ref.float()
    got = got.float()
    return ((got - ref).abs().max() / ref.abs().max().clamp_min(1e-6)).item()


@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (256, 128, 320), (4096, 320, 320), (8192, 1280, 320),
                                   (154, 640, 768), (77, 320, 1024), (2, 1280, 320), (1000, 2560, 640)])
def test_linear(nat, m, n, k):
    torch.manual_seed(0)
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16) * (k ** -0.5)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    out = nat.linear(x, w, bias=b)
    ref = x.float() @ w.float().t() + b.float()
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("tile_n", [32, 64, 128, 160, 224, 256])
def test_linear_tiles_and_epilogues(nat, tile_n):
    torch.manual_seed(1)
    m, n, k = 512, 640, 1280
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16) * (k ** -0.5)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    res = torch.randn(m, n, device="cuda", dtype=torch.float16)
    out = nat.linear(x, w, bias=b, residual=res, act=nat.ACT_SILU, out_scale=0.5, tile_n=tile_n)
    ref = F.silu(x.float() @ w.float().t() + b.float()) * 0.5 + res.float()
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("m,n,k", [(128, 1280, 11520), (512, 640, 5760), (77, 320, 4096), (2048, 96, 2304)])
def test_linear_split_k(nat, m, n, k):
    """Few output tiles + long K: several CTAs share a tile's K range; fixed-order reduction."""
    torch.manual_seed(12)
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16) * (k ** -0.5)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    res = torch.randn(m, n, device="cuda", dtype=torch.float16)
    scratch = torch.zeros((64 << 20) + 16384, device="cuda", dtype=torch.uint8)
    out = nat.linear(x, w, bias=b, residual=res, splitk_scratch=scratch)
    ref = x.float() @ w.float().t() + b.float() + res.float()
    assert _rel_err(out, ref) < 4e-3
    out2 = nat.linear(x, w, bias=b, residual=res, splitk_scratch=scratch)  # scratch is left reusable
    assert torch.equal(out, out2)


@pytest.mark.parametrize("m,n,k,act,res", [(8192, 2560, 320, "geglu", False), (2048, 640, 640, "none", True), (512, 1280, 1280, "silu", False),
                                           (300, 960, 320, "none", False), (8192, 320, 1280, "none", True)])
def test_linear_persistent_kernel(nat, m, n, k, act, res):
    """persist=2: persistent CTAs with two TMEM accumulators (the producer runs into the next tile while the epilogue
    drains the previous one), asked for directly instead of through the tuning table."""
    from chiaswarm_b200.ops.packing import interleave_geglu

    torch.manual_seed(21)
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16) * (k ** -0.5)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    r = torch.randn(m, n, device="cuda", dtype=torch.float16) if res else None
    y = x.float() @ w.float().t() + b.float()
    if act == "geglu":
        wi, bi = interleave_geglu(w, b, 64)
        out = nat.linear(x, wi, bias=bi, act=nat.ACT_GEGLU, tile_n=128, persist=2)
        ref = y[:, :n // 2] * F.gelu(y[:, n // 2:])
    else:
        out = nat.linear(x, w, bias=b, residual=r, act=nat.ACT_SILU if act == "silu" else nat.ACT_NONE, tile_n=128, cta_pair=1,
                         persist=2)
        ref = (F.silu(y) if act == "silu" else y) + (r.float() if res else 0)
    assert _rel_err(out, ref) < 5e-3


@pytest.mark.parametrize("m,n,k", [(512, 1280, 1280), (128, 1280, 1280), (2048, 640, 2560), (300, 320, 1280)])
def test_linear_cluster_split_k(nat, m, n, k):
    """cluster_k=2: K split over a cluster of two CTAs, the halves of the accumulator exchanged through distributed
    shared memory behind one cluster barrier (no scratch, no global round trip)."""
    torch.manual_seed(22)
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16) * (k ** -0.5)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    r = torch.randn(m, n, device="cuda", dtype=torch.float16)
    out = nat.linear(x, w, bias=b, residual=r, tile_n=128, cta_pair=1, cluster_k=2)
    ref = x.float() @ w.float().t() + b.float() + r.float()
    assert _rel_err(out, ref) < 4e-3
    assert torch.equal(out, nat.linear(x, w, bias=b, residual=r, tile_n=128, cta_pair=1, cluster_k=2))  # deterministic


def test_conv_split_k(nat):
    from chiaswarm_b200.ops.packing import pack_conv_weight

    torch.manual_seed(13)
    n, h, w, cin, cout = 2, 8, 8, 1280, 1280
    x = torch.randn(n, cin, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=torch.float16) * ((9 * cin) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    scratch = torch.zeros((64 << 20) + 16384, device="cuda", dtype=torch.uint8)
    out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
    nat.igemm(x.permute(0, 2, 3, 1).contiguous(), pack_conv_weight(wt), out, n=n, h=h, w=w, kh=3, kw=3, pad=1, bias=b,
              splitk_scratch=scratch)
    ref = F.conv2d(x.float(), wt.float(), b.float(), padding=1).permute(0, 2, 3, 1)
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("half", [64, 128])
def test_linear_geglu(nat, half):
    from chiaswarm_b200.ops.packing import interleave_geglu

    torch.manual_seed(2)
    m, d = 300, 320
    x = torch.randn(m, d, device="cuda", dtype=torch.float16)
    w = torch.randn(8 * d, d, device="cuda", dtype=torch.float16) * (d ** -0.5)
    b = torch.randn(8 * d, device="cuda", dtype=torch.float16)
    wi, bi = interleave_geglu(w, b, half)
    out = nat.linear(x, wi, bias=bi, act=nat.ACT_GEGLU, tile_n=2 * half)
    h = x.float() @ w.float().t() + b.float()
    val, gate = h.chunk(2, dim=-1)
    ref = val * F.gelu(gate)
    assert out.shape == (m, 4 * d)
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("n,h,w,cin,cout", [(2, 64, 64, 320, 320), (2, 32, 32, 640, 640), (2, 16, 16, 1280, 1280),
                                            (2, 8, 8, 1280, 1280), (1, 8, 8, 64, 64), (2, 64, 64, 8, 320),
                                            (2, 64, 64, 320, 8), (1, 96, 96, 64, 128), (3, 40, 72, 64, 64),
                                            (1, 5, 9, 32, 32)])
def test_conv3x3(nat, n, h, w, cin, cout):
    from chiaswarm_b200.ops.packing import pack_conv_weight

    torch.manual_seed(3)
    x = torch.randn(n, cin, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=torch.float16) * ((9 * cin) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
    nat.igemm(x_nhwc, pack_conv_weight(wt), out, n=n, h=h, w=w, kh=3, kw=3, pad=1, bias=b)
    ref = F.conv2d(x.float(), wt.float(), b.float(), padding=1).permute(0, 2, 3, 1)
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("n,h,w,cin,cout,asym", [(2, 64, 64, 320, 320, False), (2, 32, 32, 640, 640, False),
                                                 (2, 16, 16, 1280, 1280, False), (1, 128, 128, 128, 128, True),
                                                 (1, 10, 18, 64, 96, False), (3, 9, 7, 32, 32, True), (2, 96, 96, 320, 320, False)])
def test_conv3x3_stride2_implicit(nat, n, h, w, cin, cout, asym):
    """Stride-2 3x3 conv straight from the NHWC tensor (TMA element strides), UNet (pad 1) and VAE-encoder
    ((0,1,0,1) pad, then valid) flavours, odd sizes included, against F.conv2d(stride=2)."""
    from chiaswarm_b200.ops.packing import pack_conv_weight

    torch.manual_seed(31)
    x = torch.randn(n, cin, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=torch.float16) * ((9 * cin) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    if asym:
        ref = F.conv2d(F.pad(x.float(), (0, 1, 0, 1)), wt.float(), b.float(), stride=2)
    else:
        ref = F.conv2d(x.float(), wt.float(), b.float(), stride=2, padding=1)
    oh, ow = ref.shape[2], ref.shape[3]
    out = torch.empty(n, oh, ow, cout, device="cuda", dtype=torch.float16)
    scratch = torch.zeros((64 << 20) + 16384, dtype=torch.uint8, device="cuda")
    nat.igemm(x.permute(0, 2, 3, 1).contiguous(), pack_conv_weight(wt), out, n=n, h=oh, w=ow, kh=3, kw=3,
              pad=0 if asym else 1, stride=2, in_h=h, in_w=w, bias=b, splitk_scratch=scratch)
    assert _rel_err(out, ref.permute(0, 2, 3, 1)) < 4e-3


@pytest.mark.parametrize("n,h,w,cin,cout", [(2, 16, 16, 1280, 1280), (2, 32, 32, 640, 640), (1, 64, 64, 512, 512),
                                            (1, 5, 9, 32, 64), (1, 128, 128, 256, 256), (6, 20, 36, 640, 640),
                                            (12, 20, 36, 640, 640), (6, 10, 18, 1280, 1280), (3, 40, 24, 320, 320)])
def test_upsample2x_folded_into_conv(nat, n, h, w, cin, cout):
    """conv3x3(nearest2x(x)) as four phase convs (2x2 taps, pre-summed weights) writing every other output pixel,
    against F.interpolate + F.conv2d."""
    from chiaswarm_b200.ops.packing import fold_upsample2x

    torch.manual_seed(32)
    x = torch.randn(n, cin, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=torch.float16) * ((9 * cin) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    out = torch.zeros(n, 2 * h, 2 * w, cout, device="cuda", dtype=torch.float16)
    for (pa, pb), pw in fold_upsample2x(wt).items():
        nat.igemm(x_nhwc, pw, out[:, pa::2, pb::2, :], n=n, h=h, w=w, kh=2, kw=2, pad=1 - pa, pad_w=1 - pb, bias=b,
                  out_sx=2, out_sy=2, ldc=cout)
    ref = F.conv2d(F.interpolate(x.float(), scale_factor=2, mode="nearest"), wt.float(), b.float(), padding=1)
    assert _rel_err(out, ref.permute(0, 2, 3, 1)) < 6e-3
    # the same four phases as the z slices of one launch
    ph = fold_upsample2x(wt)
    stacked = torch.cat([ph[(a, bb)] for a in (0, 1) for bb in (0, 1)], 0).contiguous()
    out1 = torch.zeros_like(out)
    nat.igemm(x_nhwc, stacked, out1, n=n, h=h, w=w, kh=2, kw=2, pad=1, bias=b, out_sx=2, out_sy=2, out_phases=1, ldc=cout)
    assert _rel_err(out1, ref.permute(0, 2, 3, 1)) < 6e-3


@pytest.mark.parametrize("n,h,w,cin,cout,tile_n", [(2, 64, 64, 320, 320, 0), (2, 32, 32, 640, 640, 256), (2, 16, 16, 128, 192, 64),
                                                   (1, 96, 96, 64, 128, 128), (3, 40, 72, 64, 320, 160), (1, 24, 24, 64, 64, 64)])
def test_conv3x3_cta_pair(nat, n, h, w, cin, cout, tile_n):
    """256 x N tiles on tcgen05 cta_group::2 (cluster of two CTAs), incl. an odd tile count."""
    from chiaswarm_b200.ops.packing import pack_conv_weight

    torch.manual_seed(14)
    x = torch.randn(n, cin, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=torch.float16) * ((9 * cin) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    res = torch.randn(n, h, w, cout, device="cuda", dtype=torch.float16)
    out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
    nat.igemm(x.permute(0, 2, 3, 1).contiguous(), pack_conv_weight(wt), out, n=n, h=h, w=w, kh=3, kw=3, pad=1, bias=b,
              residual=res, tile_n=tile_n, cta_pair=2)
    ref = F.conv2d(x.float(), wt.float(), b.float(), padding=1).permute(0, 2, 3, 1) + res.float()
    assert _rel_err(out, ref) < 4e-3


def test_linear_cta_pair_and_geglu(nat):
    from chiaswarm_b200.ops.packing import interleave_geglu

    torch.manual_seed(15)
    m, d = 1300, 640
    x = torch.randn(m, d, device="cuda", dtype=torch.float16)
    w = torch.randn(1920, d, device="cuda", dtype=torch.float16) * (d ** -0.5)
    out = nat.linear(x, w, cta_pair=2)
    assert _rel_err(out, x.float() @ w.float().t()) < 4e-3
    wg = torch.randn(8 * d, d, device="cuda", dtype=torch.float16) * (d ** -0.5)
    bg = torch.randn(8 * d, device="cuda", dtype=torch.float16)
    wi, bi = interleave_geglu(wg, bg)
    out = nat.linear(x, wi, bias=bi, act=nat.ACT_GEGLU, cta_pair=2)
    h = x.float() @ wg.float().t() + bg.float()
    val, gate = h.chunk(2, dim=-1)
    assert _rel_err(out, val * F.gelu(gate)) < 4e-3


def test_conv_dual_input_rowbias_residual(nat):
    from chiaswarm_b200.ops.packing import pack_conv_weight

    torch.manual_seed(4)
    n, h, w, c1, c2, cout = 2, 32, 32, 640, 320, 640
    x1 = torch.randn(n, c1, h, w, device="cuda", dtype=torch.float16)
    x2 = torch.randn(n, c2, h, w, device="cuda", dtype=torch.float16)
    wt = torch.randn(cout, c1 + c2, 3, 3, device="cuda", dtype=torch.float16) * ((9 * (c1 + c2)) ** -0.5)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    rb = torch.randn(n, cout, device="cuda", dtype=torch.float16)
    res = torch.randn(n, h, w, cout, device="cuda", dtype=torch.float16)
    out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
    nat.igemm(x1.permute(0, 2, 3, 1).contiguous(), pack_conv_weight(wt), out, n=n, h=h, w=w, kh=3, kw=3, pad=1,
              a2=x2.permute(0, 2, 3, 1).contiguous(), bias=b, row_bias=rb, residual=res)
    ref = F.conv2d(torch.cat([x1, x2], 1).float(), wt.float(), b.float(), padding=1)
    ref = ref + rb.float()[:, :, None, None]
    ref = ref.permute(0, 2, 3, 1) + res.float()
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("b,heads,sq,skv,d", [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160),
                                              (2, 8, 64, 64, 160), (2, 8, 4096, 77, 40), (2, 8, 256, 77, 160),
                                              (1, 10, 1024, 1024, 64), (3, 5, 45, 45, 64), (2, 4, 300, 200, 64),
                                              (1, 2, 130, 513, 40)])
def test_flash_attn(nat, b, heads, sq, skv, d):
    torch.manual_seed(5)
    q = torch.randn(b, sq, heads, d, device="cuda", dtype=torch.float16)
    k = torch.randn(b, skv, heads, d, device="cuda", dtype=torch.float16)
    v = torch.randn(b, skv, heads, d, device="cuda", dtype=torch.float16)
    out = torch.empty(b, sq, heads * d, device="cuda", dtype=torch.float16)
    nat.flash_attn(q, k, v, out, batch=b, heads=heads, seq_q=sq, seq_kv=skv, head_dim=d)
    qf, kf, vf = (t.float().permute(0, 2, 1, 3) for t in (q, k, v))
    s = (qf @ kf.transpose(-1, -2)) * (d ** -0.5)
    ref = (s.softmax(-1) @ vf).permute(0, 2, 1, 3).reshape(b, sq, heads * d)
    assert _rel_err(out, ref) < 6e-3


@pytest.mark.parametrize("b,heads,s,d", [(2, 12, 77, 64), (1, 20, 77, 64), (3, 8, 128, 40), (2, 4, 45, 80)])
def test_flash_attn_causal(nat, b, heads, s, d):
    """Causal self-attention of the CLIP text encoder (one query tile) against SDPA(is_causal=True)."""
    torch.manual_seed(17)
    qkv = torch.randn(b, s, 3, heads, d, device="cuda", dtype=torch.float16)
    out = torch.zeros(b, s, heads * d, device="cuda", dtype=torch.float16)
    nat.flash_attn(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], out, batch=b, heads=heads, seq_q=s, seq_kv=s, head_dim=d,
                   causal=True)
    q, k, v = (qkv[:, :, i].float().transpose(1, 2) for i in range(3))
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(b, s, heads * d)
    assert _rel_err(out, ref) < 5e-3


def test_flash_attn_fused_qkv_layout(nat):
    """q/k/v as strided views of one [b*s, 3*h*d] projection output (how the UNet calls it)."""
    torch.manual_seed(6)
    b, s, heads, d = 2, 1024, 8, 80
    qkv = torch.randn(b, s, 3, heads, d, device="cuda", dtype=torch.float16)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    out = torch.empty(b, s, heads * d, device="cuda", dtype=torch.float16)
    nat.flash_attn(q, k, v, out, batch=b, heads=heads, seq_q=s, seq_kv=s, head_dim=d)
    ref = F.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2),
                                         v.float().transpose(1, 2)).transpose(1, 2).reshape(b, s, heads * d)
    assert _rel_err(out, ref) < 6e-3


def test_flash_attn_large_score_range(nat):
    """Forces the lazy O-rescale path: the row max keeps growing across KV tiles."""
    torch.manual_seed(7)
    b, heads, s, d = 1, 2, 1024, 64
    q = torch.randn(b, s, heads, d, device="cuda", dtype=torch.float16)
    k = torch.randn(b, s, heads, d, device="cuda", dtype=torch.float16)
    k = k * torch.linspace(0.2, 6.0, s, device="cuda", dtype=torch.float16)[None, :, None, None]
    v = torch.randn(b, s, heads, d, device="cuda", dtype=torch.float16)
    out = torch.empty(b, s, heads * d, device="cuda", dtype=torch.float16)
    nat.flash_attn(q, k, v, out, batch=b, heads=heads, seq_q=s, seq_kv=s, head_dim=d)
    ref = F.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2),
                                         v.float().transpose(1, 2)).transpose(1, 2).reshape(b, s, heads * d)
    assert _rel_err(out, ref) < 8e-3


@pytest.mark.parametrize("n,hw,c1,c2,groups,silu", [(2, 4096, 320, 0, 32, True), (2, 1024, 1280, 640, 32, True),
                                                    (2, 64, 1280, 1280, 32, False), (1, 45, 32, 0, 32, True),
                                                    (2, 16384, 128, 0, 32, True), (3, 100, 64, 32, 8, False)])
def test_groupnorm(nat, n, hw, c1, c2, groups, silu):
    torch.manual_seed(8)
    c = c1 + c2
    x1 = torch.randn(n, hw, c1, device="cuda", dtype=torch.float16) * 2 + 0.5
    x2 = torch.randn(n, hw, c2, device="cuda", dtype=torch.float16) if c2 else None
    g = torch.randn(c, device="cuda", dtype=torch.float16)
    b = torch.randn(c, device="cuda", dtype=torch.float16)
    out = torch.empty(n, hw, c, device="cuda", dtype=torch.float16)
    stats = torch.empty(nat.groupnorm_scratch_floats(n, hw, c, groups), device="cuda", dtype=torch.float32)
    nat.groupnorm(x1, g, b, out, n=n, hw=hw, groups=groups, eps=1e-5, silu=silu, stats=stats, x2=x2)
    out2 = torch.empty_like(out)
    nat.groupnorm(x1, g, b, out2, n=n, hw=hw, groups=groups, eps=1e-5, silu=silu, stats=stats, x2=x2)
    assert torch.equal(out, out2)  # fixed-order reduction: bit-reproducible, scratch reusable
    zeroed = torch.zeros_like(stats)  # engine contract: scratch zeroed once, no per-call memset node
    for _ in range(2):
        out3 = torch.empty_like(out)
        nat.groupnorm(x1, g, b, out3, n=n, hw=hw, groups=groups, eps=1e-5, silu=silu, stats=zeroed, x2=x2, scratch_zeroed=True)
        assert torch.equal(out, out3)
    xc = x1 if x2 is None else torch.cat([x1, x2], -1)
    ref = F.group_norm(xc.float().permute(0, 2, 1), groups, g.float(), b.float(), 1e-5)
    if silu:
        ref = F.silu(ref)
    ref = ref.permute(0, 2, 1)
    assert (out.float() - ref).abs().max().item() < 2e-2


@pytest.mark.parametrize("rows,c", [(8192, 320), (2048, 640), (512, 1280), (77, 768), (3, 2048)])
def test_layernorm(nat, rows, c):
    torch.manual_seed(9)
    x = torch.randn(rows, c, device="cuda", dtype=torch.float16) * 3 + 1
    g = torch.randn(c, device="cuda", dtype=torch.float16)
    b = torch.randn(c, device="cuda", dtype=torch.float16)
    out = torch.empty_like(x)
    nat.layernorm(x, g, b, out, 1e-5)
    ref = F.layer_norm(x.float(), (c,), g.float(), b.float(), 1e-5)
    assert (out.float() - ref).abs().max().item() < 2e-2


def test_upsample_im2col_layout(nat):
    torch.manual_seed(10)
    n, c, h, w = 2, 64, 12, 20
    x = torch.randn(n, c, h, w, device="cuda", dtype=torch.float32)
    nhwc = torch.empty(n, h, w, c, device="cuda", dtype=torch.float16)
    nat.nchw_to_nhwc(x, nhwc, c)
    assert torch.equal(nhwc, x.half().permute(0, 2, 3, 1))
    up = torch.empty(n, 2 * h, 2 * w, c, device="cuda", dtype=torch.float16)
    nat.upsample2x(nhwc, up, n=n, h=h, w=w)
    ref_up = F.interpolate(x.half(), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1)
    assert torch.equal(up, ref_up)
    back = torch.empty(n, c, h, w, device="cuda", dtype=torch.float32)
    nat.nhwc_to_nchw(nhwc, back, c)
    assert torch.equal(back, x.half().float())
    # stride-2 3x3 conv through im2col + GEMM
    from chiaswarm_b200.ops.packing import pack_conv_weight

    wt = torch.randn(32, c, 3, 3, device="cuda", dtype=torch.float16) * 0.05
    oh, ow = h // 2, w // 2
    cols = torch.empty(n, oh, ow, 9 * c, device="cuda", dtype=torch.float16)
    nat.im2col(nhwc, cols, n=n, h=h, w=w, kh=3, kw=3, stride=2, pad_top=1, pad_left=1, oh=oh, ow=ow)
    out = nat.linear(cols.view(-1, 9 * c), pack_conv_weight(wt))
    ref = F.conv2d(x.half().float(), wt.float(), stride=2, padding=1).permute(0, 2, 3, 1).reshape(-1, 32)
    assert _rel_err(out, ref) < 4e-3


def test_timestep_embedding_and_sched(nat):
    t = torch.tensor([999.0, 500.0], device="cuda")
    step = torch.tensor([1], device="cuda", dtype=torch.int32)
    out = torch.empty(2, 320, device="cuda", dtype=torch.float16)
    nat.timestep_embedding(t, step, out, n=2, dim=320)
    half = 160
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, device="cuda").float() / half)
    a = 500.0 * freqs
    ref = torch.cat([a.cos(), a.sin()])
    assert (out[0].float() - ref).abs().max().item() < 2e-3
    assert torch.equal(out[0], out[1])

    rows, c = 4096, 4
    eps = torch.randn(2, rows, 8, device="cuda", dtype=torch.float16)
    x = torch.randn(rows, c, device="cuda")
    x0p = torch.randn(rows, c, device="cuda")
    xin = torch.zeros(2, rows, 8, device="cuda", dtype=torch.float16)
    coef = torch.tensor([[0.0] * 8, [1.1, -0.4, 0.9, 0.3, -0.05, 0.0, 0.7, 0.0]], device="cuda")
    x_ref, x0p_ref = x.clone(), x0p.clone()
    nat.sched_step(eps=eps, n_cond=2, rows=rows, c=c, eps_ld=8, g_text=7.5, g_image=0.0, x=x, x0_prev=x0p,
                   noise=None, x_model_in=xin, xin_ld=8, coef=coef, step_idx=step)
    e = eps[0, :, :4].float() + 7.5 * (eps[1, :, :4].float() - eps[0, :, :4].float())
    x0 = 1.1 * x_ref - 0.4 * e
    xn = 0.9 * x_ref + 0.3 * x0 - 0.05 * x0p_ref
    assert (x - xn).abs().max().item() < 1e-4
    assert (x0p - x0).abs().max().item() < 1e-4
    assert (xin[1, :, :4].float() - xn * 0.7).abs().max().item() < 5e-3
    assert xin[:, :, 4:].abs().max().item() == 0


@pytest.mark.parametrize("b,f,hw,c", [(2, 8, 64, 64), (1, 24, 180, 320), (2, 5, 45, 128)])
def test_temporal_conv_3x1(nat, b, f, hw, c):
    """Conv3d (3,1,1) over frames == 3x1 implicit GEMM on the [b, f, hw, c] view (vertical pad 1, horizontal pad 0)."""
    torch.manual_seed(7)
    x = torch.randn(b, f, hw, c, device="cuda", dtype=torch.float16)
    w3 = torch.randn(c, c, 3, 1, 1, device="cuda", dtype=torch.float16) * (3 * c) ** -0.5
    bias = torch.randn(c, device="cuda", dtype=torch.float16)
    res = torch.randn(b * f * hw, c, device="cuda", dtype=torch.float16)
    wk = w3[:, :, :, 0, 0].permute(0, 2, 1).reshape(c, 3 * c).contiguous()
    out = torch.empty(b * f * hw, c, device="cuda", dtype=torch.float16)
    nat.igemm(x, wk, out, n=b, h=f, w=hw, kh=3, kw=1, pad=1, pad_w=0, bias=bias, residual=res)
    ref = F.conv3d(x.float().permute(0, 3, 1, 2).unsqueeze(-1), w3.float(), bias.float(), padding=(1, 0, 0))
    ref = ref.squeeze(-1).permute(0, 2, 3, 1).reshape(b * f * hw, c) + res.float()
    assert _rel_err(out, ref) < 4e-3


@pytest.mark.parametrize("f,hw,heads,d", [(8, 64, 4, 8), (24, 180, 5, 64), (16, 45, 20, 64), (24, 2883, 5, 64), (5, 7, 2, 40)])
def test_flash_attn_over_frames(nat, f, hw, heads, d):
    """Temporal attention: batch = pixel, sequence = frame, straight from the strides of a fused QKV buffer."""
    torch.manual_seed(8)
    c = heads * d
    qkv = torch.randn(f, hw, 3, heads, d, device="cuda", dtype=torch.float16)
    out = torch.zeros(f, hw, c, device="cuda", dtype=torch.float16)
    v = qkv.permute(1, 0, 2, 3, 4)
    nat.flash_attn(v[:, :, 0], v[:, :, 1], v[:, :, 2], out.permute(1, 0, 2), batch=hw, heads=heads, seq_q=f, seq_kv=f,
                   head_dim=d)
    q, k, vv = (v[:, :, i].float().permute(0, 2, 1, 3) for i in range(3))  # [hw, heads, f, d]
    ref = F.scaled_dot_product_attention(q, k, vv).permute(2, 0, 1, 3).reshape(f, hw, c)
    assert _rel_err(out, ref) < 5e-3


@pytest.mark.parametrize("fq,fkv,hw,heads,d", [(6, 24, 180, 8, 40), (3, 24, 37, 4, 80), (12, 24, 64, 8, 64)])
def test_flash_attn_frames_local_queries(nat, fq, fkv, hw, heads, d):
    """Frame-split clip: queries are this rank's frames, keys / values all frames (packed block-diagonal tiles with
    seq_q != seq_kv)."""
    torch.manual_seed(9)
    c = heads * d
    q = torch.randn(fq, hw, heads, d, device="cuda", dtype=torch.float16)
    kv = torch.randn(fkv, hw, 2, heads, d, device="cuda", dtype=torch.float16)
    out = torch.zeros(fq, hw, c, device="cuda", dtype=torch.float16)
    kv5 = kv.permute(1, 0, 2, 3, 4)
    nat.flash_attn(q.permute(1, 0, 2, 3), kv5[:, :, 0], kv5[:, :, 1], out.permute(1, 0, 2), batch=hw, heads=heads,
                   seq_q=fq, seq_kv=fkv, head_dim=d)
    qq = q.float().permute(1, 2, 0, 3)                     # [hw, heads, fq, d]
    kk, vv = (kv5[:, :, i].float().permute(0, 2, 1, 3) for i in range(2))
    ref = F.scaled_dot_product_attention(qq, kk, vv).permute(2, 0, 1, 3).reshape(fq, hw, c)
    assert _rel_err(out, ref) < 5e-3


@pytest.mark.parametrize("rows,c,nout,geglu", [(512, 1280, 3840, False), (2048, 640, 640, False), (8192, 320, 2560, True),
                                               (300, 320, 960, False), (1024, 1280, 10240, True)])
def test_layernorm_folded_into_gemm(nat, rows, c, nout, geglu):
    """y = x0 @ Wp^T + r (producer, emits row statistics); out = LayerNorm(y) @ W^T + b (consumer, LayerNorm folded
    into the weights + epilogue) against the unfused fp32 computation."""
    from chiaswarm_b200.ops.packing import fold_layernorm, interleave_geglu

    torch.manual_seed(12)
    x0 = torch.randn(rows, c, device="cuda", dtype=torch.float16)
    wp = torch.randn(c, c, device="cuda", dtype=torch.float16) * c ** -0.5
    bp = torch.randn(c, device="cuda", dtype=torch.float16)
    res = torch.randn(rows, c, device="cuda", dtype=torch.float16) * 2 + 0.3
    gamma = torch.randn(c, device="cuda", dtype=torch.float16) * 0.5 + 1
    beta = torch.randn(c, device="cuda", dtype=torch.float16) * 0.2
    w = torch.randn(nout, c, device="cuda", dtype=torch.float16) * c ** -0.5
    b = torch.randn(nout, device="cuda", dtype=torch.float16)
    tiles = (c + 127) // 128
    stats = torch.zeros(tiles, rows, 2, device="cuda", dtype=torch.float32)
    y = torch.empty(rows, c, device="cuda", dtype=torch.float16)
    nat.igemm(x0, wp, y, n=1, h=1, w=rows, bias=bp, residual=res, tile_n=128, row_stats_out=stats)
    y_ref = x0.float() @ wp.float().t() + bp.float() + res.float()
    assert _rel_err(y, y_ref) < 4e-3
    s_ref = torch.stack([y_ref.sum(1), (y_ref * y_ref).sum(1)], 1)
    assert _rel_err(stats.sum(0), s_ref) < 2e-3
    w2, b2, cs = fold_layernorm(w, b, gamma, beta)
    act = nat.ACT_NONE
    if geglu:
        _, cs = interleave_geglu(w2, cs)
        w2, b2 = interleave_geglu(w2, b2)
        act = nat.ACT_GEGLU
    out = torch.empty(rows, nout // 2 if geglu else nout, device="cuda", dtype=torch.float16)
    nat.igemm(y, w2, out, n=1, h=1, w=rows, bias=b2, act=act, tile_n=128 if geglu else 0, row_stats_in=stats,
              stats_tiles=tiles, ln_width=c, ln_eps=1e-5, col_sums=cs.float().contiguous())
    ln = F.layer_norm(y.float(), (c,), gamma.float(), beta.float(), 1e-5)
    ref = ln @ w.float().t() + b.float()
    if geglu:
        val, gate = ref.chunk(2, dim=-1)
        ref = val * F.gelu(gate)
    assert _rel_err(out, ref) < 8e-3
