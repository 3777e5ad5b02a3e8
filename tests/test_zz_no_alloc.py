"""GPU test of sw_groupnorm_nhwc_static, the GroupNorm entry point for callers that cannot
allocate device memory (a host whose whole memory budget already belongs to its own arena, or a
latency-critical path where cudaMalloc/cudaFree are forbidden because they can stall the device).

The entry point is driven through the C ABI with ctypes, exactly as such a caller would, and its
output is compared with the regular entry point (caller-provided scratch) on two back-to-back
calls.  The shapes cover both launch paths: small feature maps run the single-launch cluster
kernel (which needs no scratch at all), while a 256x256 map and a 128-group norm fall to the
statistics + apply kernels -- those are the calls that really read and write the built-in array,
and their second call runs on the numbers and ticket counters the first one left behind, which
is the situation that buffer is always in.  For them the result is also checked against
torch.nn.functional.group_norm evaluated in fp32.  The file name sorts last on purpose, so a
failure here cannot hide results of the other GPU tests in a stop-at-first-failure run.
"""
import ctypes as C

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

MAX_FLOATS = 256 * 1024


def _lib():
    from chiaswarm_b200.ops import native

    lib = native.lib()
    fn = lib.sw_groupnorm_nhwc_static
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                   C.c_int, C.c_float, C.c_int, C.c_void_p]
    return native, lib, fn


def _ptr(t):
    return None if t is None else t.data_ptr()


@pytest.mark.parametrize("n,hw,c1,c2,groups,silu,kernels", [
    (2, 4096, 320, 0, 32, True, 1), (2, 256, 1280, 640, 32, False, 1),
    # two-kernel path = the calls that use the built-in scratch array
    (1, 65536, 256, 0, 32, True, 2), (2, 4096, 512, 512, 128, False, 2)])
def test_static_scratch_groupnorm_matches_regular_entry_point(n, hw, c1, c2, groups, silu, kernels):
    native, lib, fn = _lib()
    torch.manual_seed(21)
    c = c1 + c2
    assert lib.sw_groupnorm_kernel_count(n, hw, c, groups) == kernels
    assert native.groupnorm_scratch_floats(n, hw, c, groups) <= MAX_FLOATS
    gamma = torch.randn(c, device="cuda", dtype=torch.float16)
    beta = torch.randn(c, device="cuda", dtype=torch.float16)
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.empty(native.groupnorm_scratch_floats(n, hw, c, groups), device="cuda", dtype=torch.float32)
    for call in range(2):  # second call reuses whatever the first left in the built-in buffer
        x1 = torch.randn(n, hw, c1, device="cuda", dtype=torch.float16) * (1.5 + call) + 0.25
        x2 = torch.randn(n, hw, c2, device="cuda", dtype=torch.float16) if c2 else None
        want = torch.empty(n, hw, c, device="cuda", dtype=torch.float16)
        got = torch.full_like(want, float("nan"))
        native.groupnorm(x1, gamma, beta, want, n=n, hw=hw, groups=groups, eps=1e-5, silu=silu, stats=scratch, x2=x2)
        rc = fn(_ptr(x1), _ptr(x2), _ptr(got), _ptr(gamma), _ptr(beta), n, hw, c1, c2, groups, 1e-5, int(silu), stream)
        assert rc == 0, lib.sw_last_error().decode()
        torch.cuda.synchronize()
        assert torch.isfinite(got.float()).all()
        assert (got.float() - want.float()).abs().max().item() < 2e-2
        if kernels == 2:
            xin = (x1 if x2 is None else torch.cat([x1, x2], dim=-1)).float()
            ref = torch.nn.functional.group_norm(xin.transpose(1, 2), groups, gamma.float(), beta.float(), 1e-5)
            if silu:
                ref = torch.nn.functional.silu(ref)
            err = (got.float() - ref.transpose(1, 2)).abs().max().item()
            assert err < 3e-2, err


def test_static_scratch_groupnorm_rejects_oversized_problems():
    native, lib, fn = _lib()
    n, hw, c, groups = 64, 4096, 320, 32  # needs far more than the built-in 1 MiB
    assert native.groupnorm_scratch_floats(n, hw, c, groups) > MAX_FLOATS
    x = torch.zeros(8, device="cuda", dtype=torch.float16)  # never touched: the size check comes first
    rc = fn(_ptr(x), None, _ptr(x), None, None, n, hw, c, 0, groups, 1e-5, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.sw_last_error().decode()
    assert "built-in buffer" in msg and str(MAX_FLOATS) in msg
