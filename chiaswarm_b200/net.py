"""Guarded fetching of user-supplied inputs (reference: swarm/job_arguments.py:156-190 for images,
swarm/video/pix2pix.py:90-122 for videos, swarm/toolbox/stitch.py:30-36 for result images).

http(s) goes through `requests` with a streaming byte cap (a lying or absent Content-Length cannot make the
worker buffer an unbounded body).  `file://` and `data:` URIs exist for tests, the benchmark and air-gapped
deployments and are OFF by default, because job URIs come from the hive: `SDAAS_LOCAL_INPUTS=1` (or
`allow_local_inputs(True)`) enables them, `file://` then only for regular files under `SDAAS_LOCAL_INPUT_ROOT`
(default: the system temp directory) and only after an `os.stat` size check -- the file is never read to find out
that it is too big, and device nodes such as /dev/zero are refused."""
from __future__ import annotations

import base64
import io
import mimetypes
import os
import stat
import tempfile
from pathlib import Path
from typing import Optional
from urllib.parse import unquote, urlparse

MAX_IMAGE_BYTES = 1048576 * 3
MAX_VIDEO_BYTES = 1048576 * 30

_local_override: Optional[bool] = None


class InputRefused(ValueError):
    """A job input the worker will not read (disabled scheme, outside the allowed root, too large)."""


def allow_local_inputs(flag: Optional[bool]) -> None:
    """Programmatic switch for file:// and data: inputs (None = back to the SDAAS_LOCAL_INPUTS environment)."""
    global _local_override
    _local_override = flag


def local_inputs_allowed() -> bool:
    if _local_override is not None:
        return _local_override
    return os.environ.get("SDAAS_LOCAL_INPUTS", "0") not in ("0", "", "false", "False")


def _local_root() -> Path:
    return Path(os.environ.get("SDAAS_LOCAL_INPUT_ROOT") or tempfile.gettempdir()).resolve()


def _local_path(uri: str, limit: int) -> Optional[Path]:
    """Validated path of a file:// URI (None for other schemes)."""
    parsed = urlparse(uri)
    if parsed.scheme != "file":
        return None
    if not local_inputs_allowed():
        raise InputRefused("file:// inputs are disabled on this worker")
    path = Path(unquote(parsed.path)).resolve()
    root = _local_root()
    if root != path and root not in path.parents:
        raise InputRefused(f"file:// input outside the allowed root {root}")
    st = os.stat(path)
    if not stat.S_ISREG(st.st_mode):
        raise InputRefused("file:// input is not a regular file")
    if st.st_size > limit:
        raise InputRefused(f"Input too large: {st.st_size} bytes (limit {limit})")
    return path


def _data_uri(uri: str, limit: int):
    """(content_type, bytes) of a data: URI (None for other schemes)."""
    if not uri.startswith("data:"):
        return None
    if not local_inputs_allowed():
        raise InputRefused("data: inputs are disabled on this worker")
    head, _, payload = uri.partition(",")
    if len(payload) > limit * 4 // 3 + 8:
        raise InputRefused(f"Input too large (limit {limit} bytes)")
    ctype = head[5:].split(";")[0] or "application/octet-stream"
    data = base64.b64decode(payload) if ";base64" in head else unquote(payload).encode()
    if len(data) > limit:
        raise InputRefused(f"Input too large: {len(data)} bytes (limit {limit})")
    return ctype, data


def probe(uri: str, limit: int = MAX_VIDEO_BYTES):
    """(content_type, content_length) without downloading the body (HEAD for http, stat for files)."""
    d = _data_uri(uri, limit)
    if d is not None:
        return d[0], len(d[1])
    path = _local_path(uri, limit)
    if path is not None:
        return mimetypes.guess_type(str(path))[0] or "application/octet-stream", os.stat(path).st_size
    import requests

    head = requests.head(uri, allow_redirects=True, timeout=30)
    return head.headers.get("Content-Type", ""), int(head.headers.get("Content-Length", 0) or 0)


def _read_capped(resp, limit: int) -> bytes:
    buf = io.BytesIO()
    for chunk in resp.iter_content(chunk_size=1 << 16):
        if not chunk:
            continue
        buf.write(chunk)
        if buf.tell() > limit:
            resp.close()
            raise InputRefused(f"Input too large: more than {limit} bytes")
    return buf.getvalue()


def fetch(uri: str, limit: int = MAX_IMAGE_BYTES) -> bytes:
    d = _data_uri(uri, limit)
    if d is not None:
        return d[1]
    path = _local_path(uri, limit)
    if path is not None:
        with open(path, "rb") as fh:
            return fh.read(limit + 1)[:limit]
    import requests

    resp = requests.get(uri, allow_redirects=True, stream=True, timeout=60)
    return _read_capped(resp, limit)


def fetch_to_file(uri: str, path: str, limit: int = MAX_VIDEO_BYTES) -> bool:
    d = _data_uri(uri, limit)
    src = None if d is not None else _local_path(uri, limit)
    if d is not None or src is not None:
        data = d[1] if d is not None else open(src, "rb").read(limit + 1)[:limit]
        with open(path, "wb") as fh:
            fh.write(data)
        return True
    import requests

    resp = requests.get(uri, allow_redirects=True, stream=True, timeout=120)
    if not resp.ok:
        print(f"Download failed: status code {resp.status_code}\n{resp.text[:500]}")
        return False
    written = 0
    with open(path, "wb") as fh:
        for chunk in resp.iter_content(chunk_size=1 << 16):
            if chunk:
                written += len(chunk)
                if written > limit:
                    resp.close()
                    raise InputRefused(f"Input too large: more than {limit} bytes")
                fh.write(chunk)
        fh.flush()
        os.fsync(fh.fileno())
    return True


def open_image(uri: str):
    from PIL import Image, ImageOps

    image = Image.open(io.BytesIO(fetch(uri, MAX_IMAGE_BYTES)))
    image = ImageOps.exif_transpose(image)
    return image.convert("RGB")
