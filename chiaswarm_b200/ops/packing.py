"""Weight re-packing from the diffusers/torch layouts into what the sm_100a kernels consume.

Done once at model-load time (the reference re-reads checkpoints per job,
swarm/diffusion/diffusion_func.py:41-46; here packed weights stay resident)."""
from __future__ import annotations

import torch


def pack_conv_weight(w: torch.Tensor, cin_pad: int | None = None, cout_pad: int | None = None) -> torch.Tensor:
    """[cout, cin, kh, kw] (torch conv layout) -> [cout_pad, kh*kw*cin_pad], tap-major then channel,
    which is the K order sw_igemm walks (tap, 64-channel slab)."""
    cout, cin, kh, kw = w.shape
    cin_pad = cin if cin_pad is None else cin_pad
    cout_pad = cout if cout_pad is None else cout_pad
    out = torch.zeros(cout_pad, kh, kw, cin_pad, dtype=w.dtype, device=w.device)
    out[:cout, :, :, :cin] = w.permute(0, 2, 3, 1)
    return out.reshape(cout_pad, kh * kw * cin_pad).contiguous()


def pad_vector(b: torch.Tensor | None, n_pad: int) -> torch.Tensor | None:
    if b is None:
        return None
    if b.shape[0] == n_pad:
        return b.contiguous()
    out = torch.zeros(n_pad, dtype=b.dtype, device=b.device)
    out[: b.shape[0]] = b
    return out


def interleave_geglu(w: torch.Tensor, b: torch.Tensor | None, half: int = 64):
    """GEGLU projection [2*inner, d] = [values; gates] -> blocks of `half` value rows followed by their
    `half` gate rows, so one 2*half-wide MMA tile holds both halves and the epilogue can fuse x*gelu(g).
    half = 64 pairs with tile_n = 128, half = 128 with tile_n = 256."""
    two_inner = w.shape[0]
    inner = two_inner // 2
    assert half in (64, 128) and inner % half == 0, "GEGLU inner dim must be a multiple of the tile half"
    val, gate = w[:inner], w[inner:]
    wi = torch.stack([val.reshape(inner // half, half, -1), gate.reshape(inner // half, half, -1)], dim=1)
    wi = wi.reshape(two_inner, -1).contiguous()
    bi = None
    if b is not None:
        bv, bg = b[:inner], b[inner:]
        bi = torch.stack([bv.reshape(inner // half, half), bg.reshape(inner // half, half)], dim=1)
        bi = bi.reshape(two_inner).contiguous()
    return wi, bi


def fold_layernorm(w: torch.Tensor, b: torch.Tensor | None, gamma: torch.Tensor, beta: torch.Tensor):
    """LayerNorm(x) @ w.T + b  ==  rstd * (x @ w2.T - mean * col_sums) + b2  with
    w2 = w * gamma (per input column), b2 = w @ beta + b, col_sums = w2.sum(1): the affine part of the
    LayerNorm moves into the weights, the statistics into the GEMM epilogue (sw_igemm row_stats_in).
    Returns (w2 fp16, b2 fp16, col_sums fp32); col_sums is taken from the ROUNDED w2 so that the epilogue
    subtracts exactly what the tensor core accumulated."""
    w32, g32, be32 = w.detach().float(), gamma.detach().float(), beta.detach().float()
    w2 = (w32 * g32[None, :]).to(torch.float16)
    b2 = w32 @ be32
    if b is not None:
        b2 = b2 + b.detach().float()
    return w2.contiguous(), b2.to(torch.float16).contiguous(), w2.float().sum(1).contiguous()


def pack_kmajor(w: torch.Tensor, taps: int, c1: int, c2: int = 0) -> torch.Tensor:
    """[cout, taps*(c1+c2)] (the K order of sw_igemm: tap, then the channels of input 1, then of input 2) ->
    k-block-major [k-blocks, cout, 64]: k-block = (tap, 64-channel slab), slabs zero-padded to 64 channels.
    Each weight tile the kernel fetches (rows of one N tile, one k-block) is then a single contiguous run of
    memory, so cold weights stream from HBM in whole-page bursts instead of one 128-byte segment per DRAM page
    (sw_igemm_params.w_kmajor).  A round trip through `unpack_kmajor` is exact."""
    cout, k = w.shape
    ctot = c1 + c2
    assert k == taps * ctot, (w.shape, taps, c1, c2)
    nkb1, nkb2 = (c1 + 63) // 64, (c2 + 63) // 64
    out = torch.zeros(taps * (nkb1 + nkb2), cout, 64, dtype=w.dtype, device=w.device)
    wt = w.reshape(cout, taps, ctot)
    for t in range(taps):
        base = t * (nkb1 + nkb2)
        for j in range(nkb1):
            lo, hi = j * 64, min(c1, j * 64 + 64)
            out[base + j, :, : hi - lo] = wt[:, t, lo:hi]
        for j in range(nkb2):
            lo, hi = j * 64, min(c2, j * 64 + 64)
            out[base + nkb1 + j, :, : hi - lo] = wt[:, t, c1 + lo:c1 + hi]
    return out.contiguous()


def unpack_kmajor(wk: torch.Tensor, taps: int, c1: int, c2: int = 0) -> torch.Tensor:
    """Inverse of pack_kmajor."""
    nkb, cout, _ = wk.shape
    nkb1, nkb2 = (c1 + 63) // 64, (c2 + 63) // 64
    assert nkb == taps * (nkb1 + nkb2)
    out = torch.empty(cout, taps, c1 + c2, dtype=wk.dtype, device=wk.device)
    for t in range(taps):
        base = t * (nkb1 + nkb2)
        for j in range(nkb1):
            lo, hi = j * 64, min(c1, j * 64 + 64)
            out[:, t, lo:hi] = wk[base + j, :, : hi - lo]
        for j in range(nkb2):
            lo, hi = j * 64, min(c2, j * 64 + 64)
            out[:, t, c1 + lo:c1 + hi] = wk[base + nkb1 + j, :, : hi - lo]
    return out.reshape(cout, taps * (c1 + c2))


def fold_upsample2x(w: torch.Tensor, cin_pad: int | None = None, cout_pad: int | None = None):
    """conv3x3(pad 1) applied to a nearest-neighbour 2x upsample, rewritten as four 2x2 convolutions on the SMALL map,
    one per output phase (a, b) = (row parity, column parity):

        out[2y+a, 2x+b] = sum_{i,j in {0,1}} P_ab[i, j] . x[y - 1 + a + i, x - 1 + b + j]      (zero outside x)

    because the three upsampled rows 2y+a-1 .. 2y+a+1 come from only two source rows: for a = 0 rows (y-1, y, y) and
    for a = 1 rows (y, y, y+1).  So the row taps fold to a = 0: [w0, w1 + w2], a = 1: [w0 + w1, w2], the same for
    columns.  2.25x fewer multiply-adds than the conv on the upsampled map, and neither the upsample kernel nor its
    4x larger buffer exist (reference call site: diffusers Upsample2D inside `pipeline(**kwargs)`,
    swarm/diffusion/diffusion_func.py:96).  Returns {(a, b): packed [cout_pad, 4 * cin_pad]} (taps summed in fp32)."""
    w32 = w.detach().float()
    rows = {0: [w32[:, :, 0], w32[:, :, 1] + w32[:, :, 2]], 1: [w32[:, :, 0] + w32[:, :, 1], w32[:, :, 2]]}
    out = {}
    for a in (0, 1):
        for b in (0, 1):
            taps = []
            for i in (0, 1):
                r = rows[a][i]                                   # [cout, cin, 3] (columns left)
                cols = [r[:, :, 0], r[:, :, 1] + r[:, :, 2]] if b == 0 else [r[:, :, 0] + r[:, :, 1], r[:, :, 2]]
                taps.append(torch.stack(cols, dim=-1))           # [cout, cin, 2]
            k = torch.stack(taps, dim=2).to(w.dtype)             # [cout, cin, 2, 2]
            out[(a, b)] = pack_conv_weight(k, cin_pad, cout_pad)
    return out
