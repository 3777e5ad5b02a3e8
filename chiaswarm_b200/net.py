"""Guarded fetching of user-supplied inputs (reference: swarm/job_arguments.py:156-190 for images,
swarm/video/pix2pix.py:90-122 for videos, swarm/toolbox/stitch.py:30-36 for result images).

Supports http(s) through `requests`, plus `file://` and `data:` URIs so that tests, the benchmark
and air-gapped deployments can feed inputs without a network."""
from __future__ import annotations

import base64
import io
import mimetypes
import os
from urllib.parse import unquote, urlparse

MAX_IMAGE_BYTES = 1048576 * 3
MAX_VIDEO_BYTES = 1048576 * 30


def _local(uri: str):
    """Returns (content_type, bytes) for file:// and data: URIs, else None."""
    if uri.startswith("data:"):
        head, _, payload = uri.partition(",")
        ctype = head[5:].split(";")[0] or "application/octet-stream"
        data = base64.b64decode(payload) if ";base64" in head else unquote(payload).encode()
        return ctype, data
    parsed = urlparse(uri)
    if parsed.scheme == "file":
        path = unquote(parsed.path)
        ctype = mimetypes.guess_type(path)[0] or "application/octet-stream"
        with open(path, "rb") as fh:
            return ctype, fh.read()
    return None


def probe(uri: str):
    """(content_type, content_length) without downloading the body (HEAD for http)."""
    local = _local(uri)
    if local is not None:
        return local[0], len(local[1])
    import requests

    head = requests.head(uri, allow_redirects=True, timeout=30)
    return head.headers.get("Content-Type", ""), int(head.headers.get("Content-Length", 0) or 0)


def fetch(uri: str) -> bytes:
    local = _local(uri)
    if local is not None:
        return local[1]
    import requests

    resp = requests.get(uri, allow_redirects=True, stream=True, timeout=60)
    return resp.content


def fetch_to_file(uri: str, path: str) -> bool:
    local = _local(uri)
    if local is not None:
        with open(path, "wb") as fh:
            fh.write(local[1])
        return True
    import requests

    resp = requests.get(uri, allow_redirects=True, stream=True, timeout=120)
    if not resp.ok:
        print(f"Download failed: status code {resp.status_code}\n{resp.text}")
        return False
    with open(path, "wb") as fh:
        for chunk in resp.iter_content(chunk_size=1 << 16):
            if chunk:
                fh.write(chunk)
        fh.flush()
        os.fsync(fh.fileno())
    return True


def open_image(uri: str):
    from PIL import Image, ImageOps

    image = Image.open(io.BytesIO(fetch(uri)))
    image = ImageOps.exif_transpose(image)
    return image.convert("RGB")
