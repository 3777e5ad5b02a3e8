"""Image sequence -> webm (reference: swarm/diffusion/video_maker.py:7-24, dead code there: nothing calls it; kept so that
`from swarm.diffusion.video_maker import make_video` keeps resolving).  Same contract -- PIL images and a duration in,
(last image, BytesIO of the video) out -- written with OpenCV (VP90, falling back to VP80) instead of moviepy, which is
not on the image."""
from __future__ import annotations

import io
import os
import tempfile

import numpy as np

from ..toolbox.video_helpers import frames_to_video


def make_video(images, duration_seconds):
    if not images:
        raise ValueError("make_video needs at least one image")
    frames = [np.asarray(img.convert("RGB")) for img in images]
    fps = max(len(frames) / float(duration_seconds), 1e-3)
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "video.webm")
        last = None
        for codec in ("VP90", "VP80"):
            try:
                frames_to_video(frames, path, codec, fps=fps)
                if os.path.getsize(path) > 0:
                    break
            except Exception as e:  # codec not built into this OpenCV
                last = e
        else:
            raise RuntimeError(f"no usable webm codec: {last}")
        with open(path, "rb") as fh:
            data = fh.read()
    # the last image (used as the thumbnail) and the video bytes
    return images[-1], io.BytesIO(data)
