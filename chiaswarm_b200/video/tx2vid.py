"""Text-to-video job callback (reference: swarm/video/tx2vid.py:17-88).  Defaults kept: 25 frames,
mp4 (h264 -> falls back to mp4v when the OpenCV build lacks an h264 encoder) or webm (VP90),
8 fps (Q12), thumbnail = frame 0.  The reference always enables model-CPU-offload; on B200 the
3-D UNet stays resident and, when the job leased a gang (parallel/gang_worker.py), the clip is split along
the frame axis across the ranks (engine/patch_engine.py::_FrameStepGraph)."""
from __future__ import annotations

import pathlib
import tempfile
from io import BytesIO

import torch

from ..output_processor import make_result
from ..parallel.context import job_gang
from ..registry import resolve
from ..toolbox.video_helpers import frames_to_video, get_frame
from ..utils.weight_cache import pipeline_cache


def txt2vid_diffusion_callback(device_identifier, model_name, **kwargs):
    scheduler_type = kwargs.pop("scheduler_type", resolve("DPMSolverMultistepScheduler"))
    kwargs.pop("pipeline_type", None)
    kwargs["num_frames"] = kwargs.pop("num_frames", 25)
    content_type = kwargs.pop("content_type", "video/mp4")
    kwargs.pop("outputs", ["primary"])
    revision = kwargs.pop("revision", "main")
    variant = kwargs.pop("variant", None)

    from .pipeline_t2v import TextToVideoPipeline

    pipeline = pipeline_cache().get(
        ("t2v", model_name, revision, variant, device_identifier),
        lambda: TextToVideoPipeline.from_pretrained(model_name, revision=revision, variant=variant,
                                                    torch_dtype=torch.float16, device=device_identifier))
    pipeline.set_scheduler(scheduler_type, use_karras_sigmas=True)
    p = pipeline(gang=job_gang(), **kwargs)  # with a gang lease: frame-split denoising + scattered VAE decode
    video_frames = p.frames

    container, codecs = ("mp4", ("h264", "avc1", "mp4v")) if content_type == "video/mp4" else ("webm", ("VP90", "VP80"))
    with tempfile.TemporaryDirectory() as tmp:
        path = str(pathlib.Path(tmp).joinpath(f"video.{container}"))
        export_to_video(video_frames, path, codecs)
        with open(path, "rb") as fh:
            video_buffer = BytesIO(fh.read())
        thumbnail = get_frame(path, 0)
    return {"primary": make_result(video_buffer, thumbnail, content_type)}, dict(pipeline.config)


def export_to_video(video_frames, output_video_path, codec) -> str:
    codecs = (codec,) if isinstance(codec, str) else tuple(codec)
    last = None
    for c in codecs:
        try:
            frames_to_video(video_frames, output_video_path, c, fps=8)
            import os

            if os.path.getsize(output_video_path) > 0:
                return output_video_path
        except Exception as e:  # codec not built into this OpenCV
            last = e
    raise RuntimeError(f"no usable video codec among {codecs}: {last}")
