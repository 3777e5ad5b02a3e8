"""ctypes binding of libswarm_sm100.so (C ABI in include/swarm_b200.h).

The reference has no native layer (every FLOP goes through diffusers ->
torch -> cuBLAS/cuDNN/SDPA, e.g. swarm/diffusion/diffusion_func.py:96).  Here the hot ops are our
own sm_100a kernels; this module is the only place Python touches them.  On a machine with a GPU
a missing library is a hard error -- there is no silent PyTorch fallback on the hot path.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

import torch

from ..build_native import LIB_PATH, build
from . import tuning

_lock = threading.Lock()
_lib = None

ACT_NONE, ACT_SILU, ACT_GELU, ACT_GEGLU = 0, 1, 2, 3

# count of kernel launches issued through this binding (bench.py reports it)
launch_count = 0


class IGemmParams(C.Structure):
    _fields_ = [
        ("a1", C.c_void_p), ("a2", C.c_void_p),
        ("c1", C.c_int32), ("c2", C.c_int32),
        ("n", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
        ("kh", C.c_int32), ("kw", C.c_int32), ("pad", C.c_int32),
        ("wgt", C.c_void_p),
        ("cout", C.c_int32),
        ("bias", C.c_void_p),
        ("row_bias", C.c_void_p),
        ("row_bias_ld", C.c_int32),
        ("residual", C.c_void_p),
        ("ldr", C.c_int32),
        ("out", C.c_void_p),
        ("ldc", C.c_int32),
        ("act", C.c_int32),
        ("out_scale", C.c_float),
        ("bf16", C.c_int32),
        ("tile_n", C.c_int32),
        ("gn_stats", C.c_void_p),
        ("gn_groups", C.c_int32),
        ("a_ld", C.c_int32),
        ("w_ld", C.c_int32),
        ("splitk_scratch", C.c_void_p),
        ("splitk_scratch_bytes", C.c_int64),
        ("splitk_min_kb", C.c_int32),
        ("pad_w1", C.c_int32),
        ("stages", C.c_int32),
        ("cta_pair", C.c_int32),
        ("a_h_total", C.c_int32),
        ("a_y_off", C.c_int32),
        ("row_stats_out", C.c_void_p),
        ("row_stats_in", C.c_void_p),
        ("col_sums", C.c_void_p),
        ("stats_tiles", C.c_int32),
        ("ln_width", C.c_int32),
        ("ln_eps", C.c_float),
        ("cluster_k", C.c_int32),
        ("persist", C.c_int32),
        ("w_kmajor", C.c_int32),
        ("stride", C.c_int32), ("in_h", C.c_int32), ("in_w", C.c_int32),
        ("out_sx", C.c_int32), ("out_sy", C.c_int32), ("out_phases", C.c_int32),
    ]


class AttnParams(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p),
        ("q_stride_b", C.c_int64), ("q_stride_s", C.c_int64), ("q_stride_h", C.c_int64),
        ("k_stride_b", C.c_int64), ("k_stride_s", C.c_int64), ("k_stride_h", C.c_int64),
        ("v_stride_b", C.c_int64), ("v_stride_s", C.c_int64), ("v_stride_h", C.c_int64),
        ("out", C.c_void_p), ("o_stride_b", C.c_int64), ("o_stride_s", C.c_int64),
        ("batch", C.c_int32), ("heads", C.c_int32), ("seq_q", C.c_int32), ("seq_kv", C.c_int32),
        ("head_dim", C.c_int32),
        ("scale", C.c_float),
        ("causal", C.c_int32),
    ]


class SchedParams(C.Structure):
    _fields_ = [
        ("eps", C.c_void_p),
        ("n_cond", C.c_int32), ("rows", C.c_int32), ("c", C.c_int32), ("eps_ld", C.c_int32),
        ("g_text", C.c_float), ("g_image", C.c_float),
        ("x", C.c_void_p), ("x0_prev", C.c_void_p), ("noise", C.c_void_p),
        ("x_model_in", C.c_void_p),
        ("xin_ld", C.c_int32),
        ("coef", C.c_void_p),
        ("step_idx", C.c_void_p),
        ("guidance", C.c_void_p),
    ]


class Peers(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("epoch", C.c_int32),
                ("peer", C.c_void_p * 8), ("flags", C.c_void_p * 8),
                ("epoch_base", C.c_void_p), ("epoch_mul", C.c_int32)]


class Xfer(C.Structure):
    _fields_ = [("dst", C.c_int32), ("src", C.c_void_p), ("dst_ptr", C.c_void_p), ("nbytes", C.c_int64)]


def make_peers(world, rank, epoch, peer_ptrs, flag_ptrs, epoch_base=None, epoch_mul=0) -> Peers:
    """`epoch_base` (int32 device tensor) + `epoch_mul`: effective epoch = epoch + *epoch_base * epoch_mul."""
    p = Peers()
    p.world, p.rank, p.epoch = world, rank, epoch
    p.epoch_base = None if epoch_base is None else epoch_base.data_ptr()
    p.epoch_mul = epoch_mul
    for d in range(world):
        p.peer[d] = peer_ptrs[d] if peer_ptrs is not None else None
        p.flags[d] = flag_ptrs[d]
    return p


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> C.CDLL:
    """Loads (building first if the .so is absent) the native library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            import os

            path = os.environ.get("SW_NATIVE_LIB") or (LIB_PATH if LIB_PATH.exists() else build())
            l = C.CDLL(str(path))
            l.sw_last_error.restype = C.c_char_p
            for name in (
                "sw_igemm", "sw_flash_attn", "sw_groupnorm_nhwc", "sw_layernorm", "sw_upsample2x_nhwc",
                "sw_im2col_nhwc", "sw_nchw_to_nhwc", "sw_nhwc_to_nchw", "sw_timestep_embedding",
                "sw_advance_counter", "sw_add", "sw_silu", "sw_sched_step", "sw_image_to_u8", "sw_softmax_rows",
                "sw_igemm_ag", "sw_igemm_rs", "sw_igemm_allgather", "sw_igemm_reducescatter", "sw_gn_partial_sums", "sw_gn_apply_sums", "sw_rs_reduce", "sw_peer_signal", "sw_peer_wait",
                "sw_flash_attn_gather", "sw_groupnorm_nhwc_dist", "sw_peer_exchange", "sw_groupnorm_nhwc_zeroed",
            ):
                getattr(l, name).restype = C.c_int
            l.sw_igemm.argtypes = [C.POINTER(IGemmParams), C.c_void_p]
            l.sw_flash_attn.argtypes = [C.POINTER(AttnParams), C.c_void_p]
            l.sw_igemm_ag.argtypes = [C.POINTER(IGemmParams), C.POINTER(Peers), C.c_int, C.c_void_p]
            l.sw_igemm_rs.argtypes = [C.POINTER(IGemmParams), C.POINTER(Peers), C.c_int, C.c_void_p, C.c_void_p]
            l.sw_igemm_allgather.argtypes = [C.POINTER(IGemmParams), C.POINTER(Peers), C.c_int, C.c_void_p]
            l.sw_igemm_reducescatter.argtypes = [C.POINTER(IGemmParams), C.POINTER(Peers), C.c_int, C.c_void_p, C.c_int,
                                                 C.c_void_p]
            l.sw_gn_partial_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
            l.sw_gn_apply_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                           C.c_int, C.c_int, C.c_longlong, C.c_float, C.c_int, C.c_void_p]
            l.sw_rs_reduce.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.POINTER(Peers), C.c_void_p]
            l.sw_peer_signal.argtypes = [C.POINTER(Peers), C.c_void_p]
            l.sw_peer_wait.argtypes = [C.POINTER(Peers), C.c_void_p]
            l.sw_peer_exchange.argtypes = [C.POINTER(Peers), C.POINTER(Xfer), C.c_int, C.c_void_p]
            l.sw_flash_attn_gather.argtypes = [C.POINTER(AttnParams), C.POINTER(Peers), C.POINTER(C.c_void_p), C.c_int,
                                               C.c_void_p]
            l.sw_groupnorm_nhwc_dist.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int,
                                                 C.c_void_p, C.POINTER(Peers), C.c_void_p]
            l.sw_groupnorm_nhwc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                            C.c_int, C.c_void_p, C.c_void_p]
            l.sw_groupnorm_nhwc_zeroed.argtypes = l.sw_groupnorm_nhwc.argtypes
            l.sw_groupnorm_kernel_count.restype = C.c_int
            l.sw_groupnorm_kernel_count.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
            l.sw_groupnorm_scratch_floats.restype = C.c_int64
            l.sw_groupnorm_scratch_floats.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
            l.sw_layernorm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                       C.c_float, C.c_void_p]
            l.sw_softmax_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p]
            l.sw_upsample2x_nhwc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_void_p]
            l.sw_im2col_nhwc.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 11 + [C.c_void_p]
            l.sw_nchw_to_nhwc.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]
            l.sw_nhwc_to_nchw.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]
            l.sw_timestep_embedding.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                C.c_float, C.c_int, C.c_float, C.c_void_p]
            l.sw_advance_counter.argtypes = [C.c_void_p, C.c_void_p]
            l.sw_add.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
            l.sw_silu.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
            l.sw_sched_step.argtypes = [C.POINTER(SchedParams), C.c_void_p]
            l.sw_image_to_u8.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
            _lib = l
    return _lib


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().sw_last_error()
        raise RuntimeError(f"{what} failed ({rc}): {msg.decode() if msg else '?'}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _count(n=1):
    global launch_count
    launch_count += n


def _f16(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float16 or not t.is_cuda:
        raise TypeError(f"{name}: expected a CUDA fp16 tensor, got {t.dtype} on {t.device}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: must be contiguous")
    return t


# --------------------------------------------------------------------------------------- ops
def igemm(a1, wgt, out, *, n, h, w, kh=1, kw=1, pad=0, a2=None, bias=None, row_bias=None,
          residual=None, act=ACT_NONE, out_scale=1.0, tile_n=0, splitk_scratch=None, cta_pair=0, splitk_min_kb=0, stages=0, pad_w=None,
          a_h_total=0, a_y_off=0, persist=0, row_stats_out=None, row_stats_in=None, col_sums=None, stats_tiles=0,
          ln_width=0, ln_eps=0.0, cluster_k=0, stride=1, in_h=0, in_w=0, out_sx=1, out_sy=1, ldc=None, out_phases=0):
    """Implicit-GEMM conv / linear (see sw_igemm in swarm_b200.h). `a1`/`a2` are channels-last
    [n,h,w,c] (any view with that memory layout), `wgt` is [cout, kh*kw*(c1+c2)]."""
    for t, nm in ((a1, "a1"), (wgt, "wgt"), (out, "out")):
        if t.dtype != torch.float16 or not t.is_cuda or t.stride(-1) != 1:
            raise TypeError(f"{nm}: expected a CUDA fp16 tensor with contiguous last dim")
    c1 = a1.shape[-1]
    c2 = 0 if a2 is None else _f16(a2, "a2").shape[-1]
    kmajor = wgt.dim() == 3  # k-block-major packing [k-blocks, cout, 64] (ops/packing.pack_kmajor)
    if kmajor:
        cout = wgt.shape[1]
        assert wgt.is_contiguous() and wgt.shape[2] == 64
        assert wgt.shape[0] == kh * kw * ((c1 + 63) // 64 + (c2 + 63) // 64), (wgt.shape, kh, kw, c1, c2)
    else:
        cout = wgt.shape[0] // (4 if out_phases else 1)
        assert wgt.shape[1] == kh * kw * (c1 + c2), (wgt.shape, kh, kw, c1, c2)
    assert a1.numel() == n * (a_h_total or in_h or h) * (in_w or w) * c1
    # strided 2-D views (a column block of a wider matrix) are passed through leading dimensions
    a_ld = 0 if a1.is_contiguous() else a1.stride(-2)
    w_ld = 0 if (kmajor or wgt.is_contiguous()) else wgt.stride(0)
    if a_ld and (a1.dim() != 2 or kh != 1 or a2 is not None):
        raise ValueError("strided a1 is only supported for plain GEMMs")
    if (tile_n == 0 and cta_pair == 0 and splitk_min_kb == 0 and stages == 0 and persist == 0 and cluster_k == 0
            and row_stats_out is None and row_stats_in is None and stride == 1 and out_sx == 1 and out_sy == 1):
        tuned = tuning.lookup(n, h, w, kh, kw, c1, c2, cout, act, residual is not None)
        if tuned is not None:
            tile_n, cta_pair, splitk_min_kb, persist, cluster_k = tuned
    p = IGemmParams()
    p.a1, p.a2 = _ptr(a1), _ptr(a2)
    p.c1, p.c2 = c1, c2
    p.n, p.h, p.w = n, h, w
    p.kh, p.kw, p.pad = kh, kw, pad
    p.wgt = _ptr(wgt)
    p.cout = cout
    p.bias = _ptr(bias)
    p.row_bias = _ptr(row_bias)
    p.row_bias_ld = 0 if row_bias is None else row_bias.stride(-2)
    p.residual = _ptr(residual)
    p.ldr = 0 if residual is None else residual.stride(-2)
    p.out = _ptr(out)
    p.ldc = ldc if ldc is not None else (out.stride(-2) if out.dim() > 1 else out.shape[-1])
    p.act = act
    p.out_scale = out_scale
    p.bf16 = 0
    p.tile_n = tile_n
    p.gn_stats = None
    p.gn_groups = 0
    p.a_ld = a_ld
    p.w_ld = w_ld
    p.cta_pair = cta_pair
    p.splitk_min_kb = splitk_min_kb
    p.stages = stages
    p.pad_w1 = 0 if pad_w is None else pad_w + 1
    p.a_h_total, p.a_y_off = a_h_total, a_y_off
    p.persist = persist
    p.cluster_k = cluster_k
    p.w_kmajor = 1 if kmajor else 0
    p.stride, p.in_h, p.in_w, p.out_sx, p.out_sy, p.out_phases = stride, in_h, in_w, out_sx, out_sy, out_phases
    p.row_stats_out, p.row_stats_in, p.col_sums = _ptr(row_stats_out), _ptr(row_stats_in), _ptr(col_sums)
    p.stats_tiles, p.ln_width, p.ln_eps = stats_tiles, ln_width, ln_eps
    if splitk_scratch is not None:
        p.splitk_scratch = splitk_scratch.data_ptr()
        p.splitk_scratch_bytes = splitk_scratch.numel() * splitk_scratch.element_size()
    _check(lib().sw_igemm(C.byref(p), _stream()), "sw_igemm")
    _count()
    return out


def linear(x, wgt, out=None, **kw):
    """x: [..., K] fp16 contiguous; wgt: [N, K]. Returns [..., N] (N/2 for GEGLU)."""
    k = x.shape[-1]
    m = x.numel() // k
    cout = wgt.shape[1] if wgt.dim() == 3 else wgt.shape[0]
    ncols = cout // 2 if kw.get("act") == ACT_GEGLU else cout
    if out is None:
        out = torch.empty(*x.shape[:-1], ncols, dtype=torch.float16, device=x.device)
    igemm(x, wgt, out, n=1, h=1, w=m, **kw)
    return out


def flash_attn(q, k, v, out, *, batch, heads, seq_q, seq_kv, head_dim, scale=None, causal=False):
    """q/k/v: fp16 tensors viewed as [batch, seq, heads, head_dim] (strides taken from the views);
    out: [batch, seq_q, heads*head_dim]."""
    for t, nm in ((q, "q"), (k, "k"), (v, "v")):
        if t.dtype != torch.float16 or t.stride(-1) != 1 or t.dim() != 4:
            raise ValueError(f"{nm}: need fp16 [b, s, h, d] view with contiguous d")
    p = AttnParams()
    p.q, p.k, p.v = q.data_ptr(), k.data_ptr(), v.data_ptr()
    p.q_stride_b, p.q_stride_s, p.q_stride_h = q.stride(0), q.stride(1), q.stride(2)
    p.k_stride_b, p.k_stride_s, p.k_stride_h = k.stride(0), k.stride(1), k.stride(2)
    p.v_stride_b, p.v_stride_s, p.v_stride_h = v.stride(0), v.stride(1), v.stride(2)
    p.out = out.data_ptr()
    p.o_stride_b, p.o_stride_s = out.stride(0), out.stride(1)
    p.batch, p.heads, p.seq_q, p.seq_kv, p.head_dim = batch, heads, seq_q, seq_kv, head_dim
    p.scale = float(scale if scale is not None else head_dim ** -0.5)
    p.causal = int(bool(causal))
    _check(lib().sw_flash_attn(C.byref(p), _stream()), "sw_flash_attn")
    _count()
    return out


def groupnorm_scratch_floats(n, hw, c, groups) -> int:
    return int(lib().sw_groupnorm_scratch_floats(n, hw, c, groups))


def groupnorm(x, gamma, beta, out, *, n, hw, groups, eps, silu, stats, x2=None, scratch_zeroed=False):
    """`scratch_zeroed`: the caller zeroed `stats` once at allocation (skips the per-call memset node)."""
    c1 = x.shape[-1]
    c2 = 0 if x2 is None else x2.shape[-1]
    need = groupnorm_scratch_floats(n, hw, c1 + c2, groups)
    if stats.numel() < need:
        raise ValueError(f"groupnorm scratch too small: {stats.numel()} < {need} floats")
    fn = lib().sw_groupnorm_nhwc_zeroed if scratch_zeroed else lib().sw_groupnorm_nhwc
    _check(fn(_ptr(x), _ptr(x2), _ptr(out), _ptr(gamma), _ptr(beta), n, hw, c1, c2, groups, eps, int(silu),
              _ptr(stats), _stream()), "sw_groupnorm_nhwc")
    _count(lib().sw_groupnorm_kernel_count(n, hw, c1 + c2, groups))
    return out


def layernorm(x, gamma, beta, out, eps=1e-5):
    c = x.shape[-1]
    _check(lib().sw_layernorm(_ptr(x), _ptr(out), _ptr(gamma), _ptr(beta), x.numel() // c, c, eps, _stream()),
           "sw_layernorm")
    _count()
    return out


def softmax_rows(x, out, scale=1.0):
    cols = x.shape[-1]
    _check(lib().sw_softmax_rows(_ptr(x), _ptr(out), x.numel() // cols, cols, scale, _stream()), "sw_softmax_rows")
    _count()
    return out


def upsample2x(x, out, *, n, h, w):
    _check(lib().sw_upsample2x_nhwc(_ptr(x), _ptr(out), n, h, w, x.shape[-1], _stream()), "sw_upsample2x_nhwc")
    _count()
    return out


def im2col(x, out, *, n, h, w, kh, kw, stride, pad_top, pad_left, oh, ow):
    _check(lib().sw_im2col_nhwc(_ptr(x), _ptr(out), n, h, w, x.shape[-1], kh, kw, stride, pad_top, pad_left,
                                oh, ow, _stream()), "sw_im2col_nhwc")
    _count()
    return out


def nchw_to_nhwc(x, out, c_pad):
    n, c, h, w = x.shape
    _check(lib().sw_nchw_to_nhwc(_ptr(x), _ptr(out), n, c, h, w, c_pad, int(x.dtype == torch.float32), _stream()),
           "sw_nchw_to_nhwc")
    _count()
    return out


def nhwc_to_nchw(x, out, c_pad):
    n, c, h, w = out.shape
    _check(lib().sw_nhwc_to_nchw(_ptr(x), _ptr(out), n, c, h, w, c_pad, int(out.dtype == torch.float32), _stream()),
           "sw_nhwc_to_nchw")
    _count()
    return out


def timestep_embedding(t_table, step_idx, out, *, n, dim, max_period=10000.0, flip_sin_to_cos=True,
                       freq_shift=0.0):
    _check(lib().sw_timestep_embedding(_ptr(t_table), _ptr(step_idx), _ptr(out), n, dim, max_period,
                                       int(flip_sin_to_cos), freq_shift, _stream()), "sw_timestep_embedding")
    _count()
    return out


def advance_counter(counter):
    _check(lib().sw_advance_counter(_ptr(counter), _stream()), "sw_advance_counter")
    _count()


def add(a, b, out):
    _check(lib().sw_add(_ptr(a), _ptr(b), _ptr(out), a.numel(), _stream()), "sw_add")
    _count()
    return out


def silu(x, out):
    _check(lib().sw_silu(_ptr(x), _ptr(out), x.numel(), _stream()), "sw_silu")
    _count()
    return out


def sched_step(*, eps, n_cond, rows, c, eps_ld, g_text, g_image, x, x0_prev, noise, x_model_in, xin_ld,
               coef, step_idx, guidance=None):
    """`guidance`: optional device fp32 [2] = (g_text, g_image) read by the kernel instead of the scalars."""
    p = SchedParams()
    p.eps = _ptr(eps)
    p.n_cond, p.rows, p.c, p.eps_ld = n_cond, rows, c, eps_ld
    p.g_text, p.g_image = g_text, g_image
    p.x, p.x0_prev, p.noise = _ptr(x), _ptr(x0_prev), _ptr(noise)
    p.x_model_in = _ptr(x_model_in)
    p.xin_ld = xin_ld
    p.coef = _ptr(coef)
    p.step_idx = _ptr(step_idx)
    p.guidance = _ptr(guidance)
    _check(lib().sw_sched_step(C.byref(p), _stream()), "sw_sched_step")
    _count()


def image_to_u8(x_nhwc, out, pixels, c_pad):
    _check(lib().sw_image_to_u8(_ptr(x_nhwc), _ptr(out), pixels, c_pad, _stream()), "sw_image_to_u8")
    _count()
    return out


# --------------------------------------------------------------------------- fused collectives
def _gemm_params(wgt, out_ptr, ldc, m_total, k, bias=None, residual=None, act=ACT_NONE):
    p = IGemmParams()
    p.a1 = None
    p.c1, p.c2 = k, 0
    p.n, p.h, p.w = 1, 1, m_total
    p.kh, p.kw, p.pad = 1, 1, 0
    p.wgt = wgt.data_ptr()
    p.cout = wgt.shape[0]
    p.bias = _ptr(bias)
    p.residual = _ptr(residual)
    p.ldr = 0 if residual is None else residual.stride(-2)
    p.out = out_ptr
    p.ldc = ldc
    p.act = act
    p.out_scale = 1.0
    return p


def igemm_ag(wgt, out, peers: Peers, *, m_local, k, bias=None, residual=None, act=ACT_NONE):
    """out[world*m_local, N] = concat_d(X_d) @ wgt^T, A tiles loaded from the owning ranks."""
    p = _gemm_params(wgt, out.data_ptr(), out.stride(-2), m_local * peers.world, k, bias, residual, act)
    _check(lib().sw_igemm_ag(C.byref(p), C.byref(peers), m_local, _stream()), "sw_igemm_ag")
    _count()
    return out


def igemm_allgather(wgt, out, peers: Peers, *, m_local, k, a_ld=0, bias=None, tile_n=0):
    """out[world*m_local, N] = gathered @ wgt^T (+ bias); peers.peer[d] = rank d's gathered buffer, whose rows of
    this rank are pushed to the peers by the kernel's copy role (include/swarm_b200.h: sw_igemm_allgather)."""
    p = _gemm_params(wgt, out.data_ptr(), out.stride(-2), m_local * peers.world, k, bias)
    p.a_ld, p.tile_n = a_ld, tile_n
    _check(lib().sw_igemm_allgather(C.byref(p), C.byref(peers), m_local, _stream()), "sw_igemm_allgather")
    _count()
    return out


def igemm_reducescatter(x, wgt, peers: Peers, out, *, m_local, ldc, bias=None, residual=None, tile_n=0):
    """out[m_local, N] = this rank's rows of sum_r (x_r @ wgt_r^T) (+ bias, + residual): partial tiles TMA-stored
    into the owners' staging slots, reduced by the kernel's own reduce role (sw_igemm_reducescatter)."""
    p = _gemm_params(wgt, None, ldc, m_local * peers.world, x.shape[-1], bias, residual)
    p.a1 = x.data_ptr()
    p.a_ld = x.stride(-2)
    p.tile_n = tile_n
    _check(lib().sw_igemm_reducescatter(C.byref(p), C.byref(peers), m_local, out.data_ptr(), out.stride(-2), _stream()),
           "sw_igemm_reducescatter")
    _count()
    return out


def igemm_rs(x, wgt, peers: Peers, ticket, *, m_local, ldc):
    """Partial GEMM over this rank's K shard, rows pushed to their owners' staging slots."""
    p = _gemm_params(wgt, None, ldc, m_local * peers.world, x.shape[-1])
    p.a1 = x.data_ptr()
    _check(lib().sw_igemm_rs(C.byref(p), C.byref(peers), m_local, ticket.data_ptr(), _stream()), "sw_igemm_rs")
    _count()


def rs_reduce(staging, out, peers: Peers, *, m_local, ncols, bias=None, residual=None):
    _check(lib().sw_rs_reduce(staging.data_ptr(), out.data_ptr(), _ptr(bias), _ptr(residual), m_local, ncols,
                              staging.stride(-2), out.stride(-2), 0 if residual is None else residual.stride(-2),
                              C.byref(peers), _stream()), "sw_rs_reduce")
    _count()
    return out


def gn_partial_sums(x, *, n, hw, groups):
    sums = torch.empty(n, groups, 2, dtype=torch.float32, device=x.device)
    _check(lib().sw_gn_partial_sums(_ptr(x), _ptr(sums), n, hw, x.shape[-1], groups, _stream()), "sw_gn_partial_sums")
    _count(1)
    return sums


def gn_apply_sums(x, sums, gamma, beta, out, *, n, hw, groups, count, eps, silu):
    _check(lib().sw_gn_apply_sums(_ptr(x), _ptr(sums), _ptr(gamma), _ptr(beta), _ptr(out), n, hw, x.shape[-1], groups,
                                  int(count), float(eps), int(silu), _stream()), "sw_gn_apply_sums")
    _count(1)
    return out


def peer_signal(peers: Peers):
    _check(lib().sw_peer_signal(C.byref(peers), _stream()), "sw_peer_signal")
    _count()


def peer_wait(peers: Peers):
    _check(lib().sw_peer_wait(C.byref(peers), _stream()), "sw_peer_wait")
    _count()


def peer_exchange(peers: Peers, transfers):
    """transfers: [(dst index, local source tensor (contiguous), destination address on that rank)]."""
    if not transfers:
        return
    arr = (Xfer * len(transfers))()
    for i, (dst, src, dst_ptr) in enumerate(transfers):
        arr[i].dst = dst
        if src is None:  # flag-only transfer: pairwise barrier
            arr[i].src, arr[i].dst_ptr, arr[i].nbytes = None, None, 0
            continue
        if not src.is_contiguous():
            raise ValueError("peer_exchange: source must be contiguous")
        arr[i].src, arr[i].dst_ptr = src.data_ptr(), dst_ptr
        arr[i].nbytes = src.numel() * src.element_size()
    _check(lib().sw_peer_exchange(C.byref(peers), arr, len(transfers), _stream()), "sw_peer_exchange")
    _count()


def flash_attn_gather(q, k_local, v_local, out, kpeers: Peers, v_ptrs, *, batch, heads, seq_q, s_local, head_dim,
                      scale=None):
    """q: [b, sq, h, d] view; k_local/v_local: this rank's [b, s_local, h, d] views (strides are
    taken from them; every rank uses the same layout)."""
    p = AttnParams()
    p.q = q.data_ptr()
    p.q_stride_b, p.q_stride_s, p.q_stride_h = q.stride(0), q.stride(1), q.stride(2)
    p.k_stride_b, p.k_stride_s, p.k_stride_h = k_local.stride(0), k_local.stride(1), k_local.stride(2)
    p.v_stride_b, p.v_stride_s, p.v_stride_h = v_local.stride(0), v_local.stride(1), v_local.stride(2)
    p.out = out.data_ptr()
    p.o_stride_b, p.o_stride_s = out.stride(0), out.stride(1)
    p.batch, p.heads, p.seq_q, p.seq_kv, p.head_dim = batch, heads, seq_q, s_local * kpeers.world, head_dim
    p.scale = float(scale if scale is not None else head_dim ** -0.5)
    arr = (C.c_void_p * 8)(*[v_ptrs[d] if d < kpeers.world else None for d in range(8)])
    _check(lib().sw_flash_attn_gather(C.byref(p), C.byref(kpeers), arr, s_local, _stream()), "sw_flash_attn_gather")
    _count()
    return out


def groupnorm_dist(x, gamma, beta, out, peers: Peers, *, n, hw, groups, eps, silu, stats, x2=None):
    c1 = x.shape[-1]
    c2 = 0 if x2 is None else x2.shape[-1]
    _check(lib().sw_groupnorm_nhwc_dist(_ptr(x), _ptr(x2), _ptr(out), _ptr(gamma), _ptr(beta), n, hw, c1, c2, groups,
                                        eps, int(silu), _ptr(stats), C.byref(peers), _stream()),
           "sw_groupnorm_nhwc_dist")
    _count(2)
    return out
