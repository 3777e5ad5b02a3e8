"""Fused compute+collective kernels on >= 2 GPUs (torchrun, one process per GPU): each is checked
against a single-GPU fp32 oracle and the NCCL baseline path."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_fused_collectives_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "scripts", "dist_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=800)
    assert "DIST_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("guidance", [1.0, 7.5])
def test_patch_parallel_engine_two_ranks(guidance):
    """SD1.5-shaped UNet, one image over 2 GPUs: guidance 1.0 -> two strips (halo push, GroupNorm all-reduce,
    K/V-gather attention, K/V push for the short mid-block sequence), 7.5 -> one CFG branch per GPU (epsilon
    exchange only); latents must match the single-GPU engine."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29537", os.path.join(ROOT, "scripts", "patch_check.py"), "--arch", "sd15",
           "--size", "512", "--steps", "3", "--guidance", str(guidance)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=800)
    assert "PATCH_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
