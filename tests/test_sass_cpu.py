"""What the compiled library contains, read from its SASS (cuobjdump runs without a GPU): the GEMM and attention kernels
are tcgen05 / TMA kernels and carry no mma.sync, and no TMA LOAD in the whole library is a multicast -- TMA multicast loads
here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU; the library
has none at all, so none can reach a peer mapping.  (The only multicast is `tcgen05.commit` to the two CTAs of a pair.)"""
import re
import shutil
import subprocess

import pytest


@pytest.fixture(scope="module")
def kernels():
    from chiaswarm_b200.ops import native

    if shutil.which("cuobjdump") is None or not native.LIB_PATH.exists():
        pytest.skip("needs cuobjdump and the built library")
    out = subprocess.run(["cuobjdump", "-sass", str(native.LIB_PATH)], capture_output=True, text=True, check=True).stdout
    found, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            found[name] = []
        elif name is not None:
            found[name].append(line)
    assert len(found) > 20
    return found


def _ops(lines, pattern):
    return [m.group(0) for line in lines for m in [re.search(pattern, line)] if m]


def test_gemm_and_attention_kernels_are_tcgen05_tma_kernels(kernels):
    hot = {k: v for k, v in kernels.items() if re.search(r"k_igemm|k_flash_attn", k)}
    assert len(hot) >= 8
    for name, lines in hot.items():
        assert _ops(lines, r"\bUTCHMMA[.\w]*"), f"{name}: no tcgen05.mma"
        assert _ops(lines, r"\bUTMALDG[.\w]*"), f"{name}: no TMA load"
        assert _ops(lines, r"\bLDTM[.\w]*"), f"{name}: no tcgen05.ld (TMEM read-back)"
    for name, lines in kernels.items():
        assert not _ops(lines, r"(?<![A-Z])HMMA[.\w]*"), f"{name}: mma.sync (HMMA) found"
    pair = [k for k, v in kernels.items() if _ops(v, r"\bUTCHMMA\.2CTA")]
    assert pair, "no cta_group::2 kernel"


def test_no_tma_load_is_a_multicast(kernels):
    loads = [op for lines in kernels.values() for op in _ops(lines, r"\b(?:UTMALDG|UBLKCP)[.\w]*")]
    assert len(loads) > 100
    assert not [op for op in loads if "MULTICAST" in op.upper()], "a TMA load with cluster multicast was compiled in"
    multicast = sorted({op for lines in kernels.values() for op in _ops(lines, r"\b[A-Z]+[.\w]*MULTICAST[.\w]*")})
    assert multicast == ["UTCBAR.2CTA.MULTICAST"], multicast   # tcgen05.commit to both CTAs of a pair: an mbarrier arrive


def test_peer_traffic_uses_system_scope(kernels):
    """Kernels that publish or read epoch flags across GPUs do so with system-scope accesses."""
    for needle in ("k_peer_signal", "k_peer_wait", "k_peer_exchange", "k_igemm_lean_ag", "k_igemm_lean_rs", "k_flash_attn2_gather"):
        hits = [k for k in kernels if needle in k]
        assert hits, needle
        for k in hits:
            assert any(re.search(r"\.SYS\b", line) for line in kernels[k]), f"{k}: no .SYS access"


def test_hot_kernels_keep_their_arguments_out_of_the_stack():
    """`cuobjdump -res-usage`: the single-thread TMA / MMA issue loops of the GEMM and attention kernels must not run off a
    stack copy of the argument struct (a `__noinline__` helper taking it by reference once cost 512 bytes of stack and a
    2-3x slower kernel), and nothing spills to local memory."""
    from chiaswarm_b200.ops import native

    if shutil.which("cuobjdump") is None or not native.LIB_PATH.exists():
        pytest.skip("needs cuobjdump and the built library")
    out = subprocess.run(["cuobjdump", "-res-usage", str(native.LIB_PATH)], capture_output=True, text=True, check=True).stdout
    usage, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and name:
            usage[name] = tuple(int(v) for v in m.groups())
            name = None
    hot = {k: v for k, v in usage.items() if re.search(r"k_igemm|k_flash_attn|k_gn_fused|k_layernorm", k)}
    assert len(hot) >= 12
    for k, (reg, stack, shared, local) in hot.items():
        assert stack <= 16 and local == 0, f"{k}: STACK {stack} LOCAL {local}"
        assert reg <= 168, f"{k}: {reg} registers (two CTAs of the attention kernel per SM need <= 168)"
    assert all(v[3] == 0 for v in usage.values()), "a kernel spills to local memory"
