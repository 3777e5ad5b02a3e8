"""Artifact packaging (reference: swarm/output_processor.py:10-136).

Wire contract kept bit-for-bit: each artifact is {blob (base64), content_type, thumbnail (base64
JPEG <= 100x100), sha256_hash}; 2..9 images are tiled into one grid; text results hash the raw
string rather than the JSON blob (reference quirk Q9, kept because the hive may verify it)."""
from __future__ import annotations

import base64
import hashlib
import io
import json
import threading
from io import BytesIO

from PIL import Image, ImageDraw

# images -> (rows, cols) of the contact grid
_GRID_FOR_COUNT = {2: (1, 2), 3: (2, 2), 4: (2, 2), 5: (2, 3), 6: (2, 3), 7: (3, 3), 8: (3, 3), 9: (3, 3)}
MAX_GRID_IMAGES = 9
THUMB_BOX = (100, 100)


_tls = threading.local()


class DeferredArtifact:
    """An artifact whose encoding (grid, JPEG / PNG, thumbnail, base64, sha256) has not run yet.  The serve loop asks for
    these (`generator.synchronous_do_work_function(..., defer_packaging=True)`) so that a GPU lane starts its next job
    while a result thread encodes the previous one; `resolve()` gives the wire dictionary."""

    def __init__(self, pack):
        self._pack, self._value = pack, None

    def resolve(self) -> dict:
        if self._value is None:
            self._value = self._pack()
            self._pack = None
        return self._value


class BatchArtifacts(list):
    """What the image callback answers when several jobs shared one denoising run (generator.synchronous_do_work_batch):
    one artifacts dict per job, in job order, plus the per-image NSFW flags."""

    def __init__(self, per_job, nsfw=None):
        super().__init__(per_job)
        self.nsfw = list(nsfw) if nsfw is not None else [False] * len(per_job)


def packaging_deferred() -> bool:
    return bool(getattr(_tls, "defer", False))


class deferring_packaging:
    """`with deferring_packaging(flag):` -- OutputProcessors of this thread hand out DeferredArtifacts."""

    def __init__(self, flag: bool = True):
        self.flag = bool(flag)

    def __enter__(self):
        self.prev = packaging_deferred()
        _tls.defer = self.flag
        return self

    def __exit__(self, *exc):
        _tls.defer = self.prev
        return False


def _b64(data: bytes) -> str:
    return base64.b64encode(data).decode("UTF-8")


def image_to_buffer(image, content_type, quality="web_high") -> BytesIO:
    if not content_type.startswith("image"):
        raise ValueError(f"Unsupported content type: {content_type}")
    buf = io.BytesIO()
    if content_type == "image/png":
        image.save(buf, format="PNG")
    elif content_type == "image/jpeg":
        image.save(buf, format="JPEG", quality=quality, optimize=True, progressive=True)
    else:
        raise ValueError(f"Invalid image format: {content_type}")
    buf.seek(0)
    return buf


def image_from_text(text, size=(512, 512), color=0) -> Image.Image:
    image = Image.new(mode="RGB", size=size, color=color)
    ImageDraw.Draw(image).multiline_text((5, 5), str(text))
    return image


def make_thumbnail(buffer) -> BytesIO:
    """`buffer`: encoded image bytes / BytesIO (the reference's form, swarm/output_processor.py:73-80), or the PIL image
    itself -- the packer below hands over the image it just encoded, which saves decoding a progressive JPEG again."""
    if isinstance(buffer, Image.Image):
        image = buffer.convert("RGB")  # (a copy: thumbnail() works in place)
    else:
        if not isinstance(buffer, BytesIO):
            buffer = BytesIO(buffer)
        buffer.seek(0)
        image = Image.open(buffer).convert("RGB")
    image.thumbnail(THUMB_BOX, Image.Resampling.LANCZOS)
    return image_to_buffer(image, "image/jpeg", "web_low")


def _placeholder_thumb(label: str) -> BytesIO:
    return image_to_buffer(image_from_text(label, THUMB_BOX, 1), "image/jpeg", "web_low")


def make_result(buffer: BytesIO, thumb, content_type: str) -> dict:
    thumb = _placeholder_thumb(content_type) if thumb is None else make_thumbnail(thumb)
    payload = buffer.getvalue()
    return {
        "blob": _b64(payload),
        "content_type": content_type,
        "thumbnail": _b64(thumb.getvalue()),
        "sha256_hash": hashlib.sha256(payload).hexdigest(),
    }


def make_text_result(string: str) -> dict:
    blob = json.dumps({"caption": string}).encode("utf-8")
    return {
        "blob": _b64(blob),
        "content_type": "application/json",
        "thumbnail": _b64(_placeholder_thumb("text/plain").getvalue()),
        # the raw string is hashed, not the blob (reference: swarm/output_processor.py:62-70)
        "sha256_hash": hashlib.sha256(string.encode()).hexdigest(),
    }


def image_grid(image_list, rows, cols) -> Image.Image:
    w, h = image_list[0].size
    grid = Image.new("RGB", size=(cols * w, rows * h))
    for i, image in enumerate(image_list[: rows * cols]):
        r, c = divmod(i, cols)
        grid.paste(image, box=(c * w, r * h))
    return grid


def post_process(image_list):
    count = len(image_list)
    if count == 1:
        return image_list[0]
    if count > MAX_GRID_IMAGES or count == 0:
        raise ValueError(
            f"Too many images ({count}) for post-processing. Maximum supported images: {MAX_GRID_IMAGES}"
        )
    rows, cols = _GRID_FOR_COUNT[count]
    return image_grid(image_list, rows, cols)


class OutputProcessor:
    """Collects a job's images; only the "primary" output name is honoured, plus any named
    side outputs (e.g. "preprocessed_input" for ControlNet)."""

    def __init__(self, output_list, main_content_type):
        self.outputs = []
        self.other_outputs = {}
        self.output_list = output_list
        self.main_content_type = main_content_type

    def add_outputs(self, images):
        self.outputs.extend(images)

    def add_other_outputs(self, name, images):
        self.other_outputs[name] = images

    def _pack(self, images):
        sheet = post_process(images)
        return make_result(image_to_buffer(sheet, self.main_content_type), sheet, self.main_content_type)

    def get_results(self):
        defer = packaging_deferred()

        def pack(images):
            images = list(images)
            return DeferredArtifact(lambda: self._pack(images)) if defer else self._pack(images)

        results = {}
        if "primary" in self.output_list:
            results["primary"] = pack(self.outputs)
        for name, images in self.other_outputs.items():
            results[name] = pack(images)
        return results
