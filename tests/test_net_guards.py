"""Input fetching guards (reference: the http-only downloads of swarm/job_arguments.py:156-190 and
swarm/video/pix2pix.py:90-122): local schemes are opt-in, confined to a root, size-checked before any read, and
http bodies are capped while streaming."""
import base64
import io

import pytest
from PIL import Image

from chiaswarm_b200 import net


def _png_bytes(size=(8, 8)):
    buf = io.BytesIO()
    Image.new("RGB", size, (10, 200, 30)).save(buf, format="PNG")
    return buf.getvalue()


def test_local_schemes_are_off_by_default(monkeypatch):
    net.allow_local_inputs(None)
    monkeypatch.delenv("SDAAS_LOCAL_INPUTS", raising=False)
    with pytest.raises(net.InputRefused, match="disabled"):
        net.probe("file:///etc/hostname")
    with pytest.raises(net.InputRefused, match="disabled"):
        net.fetch("data:image/png;base64," + base64.b64encode(_png_bytes()).decode())
    monkeypatch.setenv("SDAAS_LOCAL_INPUTS", "1")
    assert net.local_inputs_allowed()


def test_file_inputs_are_confined_and_stat_checked(tmp_path):
    good = tmp_path / "in.png"
    good.write_bytes(_png_bytes())
    ctype, size = net.probe(good.as_uri())
    assert ctype == "image/png" and size == good.stat().st_size
    assert net.open_image(good.as_uri()).size == (8, 8)
    with pytest.raises(net.InputRefused, match="outside the allowed root"):
        net.probe("file:///etc/hostname")
    with pytest.raises(net.InputRefused, match="not a regular file"):
        (tmp_path / "d").mkdir()
        net.probe((tmp_path / "d").as_uri())
    # a device node would never finish reading: refused by type, never opened
    link = tmp_path / "zero"
    link.symlink_to("/dev/zero")
    with pytest.raises(net.InputRefused):
        net.fetch(link.as_uri())
    big = tmp_path / "big.bin"
    with open(big, "wb") as fh:
        fh.truncate(net.MAX_IMAGE_BYTES + 1)
    with pytest.raises(net.InputRefused, match="too large"):
        net.fetch(big.as_uri())


def test_http_bodies_are_capped_while_streaming(monkeypatch):
    class Resp:
        ok = True
        closed = False

        def __init__(self, total):
            self.total = total

        def iter_content(self, chunk_size):
            sent = 0
            while sent < self.total:
                yield b"\0" * chunk_size
                sent += chunk_size

        def close(self):
            self.closed = True

    import requests

    r = Resp(10 * net.MAX_IMAGE_BYTES)
    monkeypatch.setattr(requests, "get", lambda *a, **k: r)
    with pytest.raises(net.InputRefused, match="too large"):
        net.fetch("http://example.invalid/huge")  # no Content-Length needed: the stream itself is cut
    assert r.closed
    small = Resp(1 << 16)
    monkeypatch.setattr(requests, "get", lambda *a, **k: small)
    assert len(net.fetch("http://example.invalid/small")) == 1 << 16
