import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.fixture(autouse=True)
def _local_job_inputs(tmp_path, monkeypatch):
    """The suite feeds images as data: / file:// URIs; a worker accepts those only when told to (net.py).  Settings, logs,
    the result spool and models.json live under SDAAS_ROOT: every test gets its own, so nothing is written to ~/.sdaas
    (tests that set SDAAS_ROOT themselves override this)."""
    from chiaswarm_b200 import net

    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path / "sdaas_root"))
    monkeypatch.setenv("SDAAS_LOCAL_INPUT_ROOT", str(tmp_path))
    net.allow_local_inputs(True)
    yield
    net.allow_local_inputs(None)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200); skipped in the CPU tier")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
