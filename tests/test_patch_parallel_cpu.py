"""Patch-parallel algebra on CPU with real processes (gloo, world_size 2 and 4): the row-sharded
UNet (halo exchange, GroupNorm all-reduce, K/V all-gather) and the row-sharded VAE decoder must reproduce the unsplit
models."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_patch_parallel_unet_matches_unsplit(tmp_path, world):
    port, out = _free_port(), tmp_path / "res.txt"
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_cpu_worker.py"), str(r), str(world),
                               str(port), str(out)], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    logs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(logs)[-4000:]
    err, err2, same, err_vae = out.read_text().split()
    assert float(err) < 1e-4 and float(err2) < 1e-5 and same == "1"
    assert float(err_vae) < 1e-4  # row-split VAE decode == whole decode
