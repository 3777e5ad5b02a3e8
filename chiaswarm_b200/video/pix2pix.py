"""Video-to-video: instruct-pix2pix per frame (reference: swarm/video/pix2pix.py:14-197).

Guards kept: input must be `video/*`, <= 30 MiB, <= 100 frames, resized to height 512, <= 30 fps;
Euler-ancestral with Karras sigmas; defaults 15 steps, guidance 7.5, image guidance 1.5;
`pipeline_config = {nsfw, cost = 512*512*steps*frames}`.

B200 difference: the reference runs a serial Python loop, one pipeline call per frame
(lines 53-74).  Frames are independent, so here `FRAME_BATCH` frames share one denoising run of the
resident pipeline (one batched UNet call per step instead of one per frame), and when the job leased a
gang (`parallel.context.job_gang()`) each rank takes a contiguous share of those chunks and the finished
frames are all-gathered to the gang's first rank, which encodes the video (SURVEY.md section 2.3
"frame-level parallel").  Every chunk draws its noise from a generator seeded with (job seed + index of
its first frame), so the result depends on the seed only, not on the gang size."""
from __future__ import annotations

import pathlib
import tempfile
from io import BytesIO

import torch
from PIL import Image

from .. import net
from ..output_processor import make_result
from ..parallel.context import is_leader, job_gang
from ..registry import resolve
from ..utils.weight_cache import pipeline_cache

import os

MAX_FRAMES = 100
def frame_height() -> int:
    """Working height of the frames: 512 as in the reference (swarm/video/pix2pix.py:163); SDAAS_VIDEO_HEIGHT overrides it
    (smaller for CPU tests, larger for a B200 that has the memory for it)."""
    return max(8, int(os.environ.get("SDAAS_VIDEO_HEIGHT", "512")))


FRAME_BATCH = int(os.environ.get("SDAAS_FRAME_BATCH", "8"))  # frames per denoising run (CFG x3 -> 24 UNet samples at 512 px)


def model_video_callback(device_identifier, model_name, **kwargs):
    from ..diffusion.pipelines import StableDiffusionPipeline

    pipeline = pipeline_cache().get(
        ("v2v", model_name, device_identifier),
        lambda: StableDiffusionPipeline.from_pretrained(model_name, torch_dtype=torch.float16,
                                                        device=device_identifier, family="sd_pix2pix"))
    pipeline.set_scheduler(resolve("EulerAncestralDiscreteScheduler"), use_karras_sigmas=True)

    prompt = kwargs["prompt"]
    negative_prompt = kwargs.pop("negative_prompt", "")
    guidance_scale = kwargs.pop("guidance_scale", 7.5)
    image_guidance_scale = kwargs.pop("image_guidance_scale", 1.5)
    num_inference_steps = int(kwargs.pop("num_inference_steps", 15))
    generator = kwargs.get("generator")

    with tempfile.TemporaryDirectory() as tmp:
        tmp = pathlib.Path(tmp)
        if not kwargs.get("video_uri"):
            raise ValueError("vid2vid needs a video_uri")  # (ValueError: the hive must not re-issue what cannot succeed)
        video_path = download_video(tmp, kwargs.pop("video_uri"))
        frames, fps = get_frames(tmp, video_path)
        frames = frames[:MAX_FRAMES]
        gang = job_gang()
        # fixed chunks of FRAME_BATCH frames; a gang deals out whole chunks, so a chunk's content never depends
        # on how many ranks the job leased
        chunks = [list(range(lo, min(lo + FRAME_BATCH, len(frames)))) for lo in range(0, len(frames), FRAME_BATCH)]
        if gang is not None:
            chunks = [chunks[i] for i in gang.shard(len(chunks))]
        base_seed = generator.initial_seed() if generator is not None else 0
        device = generator.device if generator is not None else torch.device(device_identifier)
        done, nsfw = [], False
        for chunk in chunks:
            print(f"frames {chunk[0]}-{chunk[-1]} of {len(frames)}")
            g = torch.Generator(device=device).manual_seed((base_seed + chunk[0]) % (2 ** 63))
            images, flagged = img2img(pipeline, prompt, guidance_scale, image_guidance_scale,
                                      [frames[i] for i in chunk], num_inference_steps, negative_prompt, frame_height(),
                                      frame_height(), g)
            nsfw = nsfw or flagged
            done.extend(im.convert("RGB") for im in images)
        if gang is not None:
            done, nsfw = _gather_frames(gang, done, nsfw)
            if not is_leader():  # one rank reports; the others are finished
                return {}, {"nsfw": nsfw, "cost": 512 * 512 * num_inference_steps * len(frames)}
        result_frames = done
        if not result_frames:
            raise ValueError("input video has no decodable frames")
        out_path = create_video(tmp, result_frames, fps)
        with open(out_path, "rb") as fh:
            video_buffer = BytesIO(fh.read())
        thumb = BytesIO()
        result_frames[0].save(thumb, format="JPEG")

    results = {"primary": make_result(video_buffer, thumb, "video/mp4")}
    return results, {"nsfw": nsfw, "cost": 512 * 512 * num_inference_steps * len(result_frames)}


def download_video(tmpdir, video_uri):
    content_type, content_length = net.probe(video_uri)
    if not content_type.startswith("video"):
        raise ValueError(f"Input does not appear to be an video.\nContent type was {content_type}.")
    if int(content_length) > net.MAX_VIDEO_BYTES:
        raise ValueError(
            f"Input video too large.\nMax size is {net.MAX_VIDEO_BYTES} bytes.\nImage was {content_length}.")
    path = str(pathlib.Path(tmpdir).joinpath("input.mp4"))
    net.fetch_to_file(video_uri, path)
    return path


def _gather_frames(gang, frames, nsfw):
    """All-gather the ranks' finished frames (rank order = frame order) as uint8 tensors."""
    import numpy as np

    dev = gang.device
    if frames:
        local = torch.from_numpy(np.stack([np.asarray(f) for f in frames])).to(dev)
        shape = torch.tensor(local.shape[1:], dtype=torch.int64, device=dev)
    else:
        local, shape = None, torch.zeros(3, dtype=torch.int64, device=dev)
    # a rank without frames (more ranks than frames) still has to bring a tensor of the right frame shape
    import torch.distributed as dist

    dist.all_reduce(shape, op=dist.ReduceOp.MAX, group=gang.group)
    if local is None:
        local = torch.zeros((0,) + tuple(int(v) for v in shape), dtype=torch.uint8, device=dev)
    every = gang.all_gather_cat(local).cpu().numpy()
    flag = gang.all_reduce_max(1.0 if nsfw else 0.0) > 0
    return [Image.fromarray(a) for a in every], flag


def img2img(pipeline, prompt, text_guidance_scale, image_guidance_scale, image, steps, neg_prompt, width, height,
            generator=None):
    """`image`: one frame or a list of equally sized frames (batched through one denoising run)."""
    batch = image if isinstance(image, (list, tuple)) else [image]

    def fit(im):
        if not isinstance(im, Image.Image):
            im = Image.open(im)
        ratio = min(height / im.height, width / im.width)
        return im.resize((max(8, int(im.width * ratio)), max(8, int(im.height * ratio))), Image.Resampling.LANCZOS)

    batch = [fit(im) for im in batch]
    many = isinstance(image, (list, tuple))
    result = pipeline([prompt] * len(batch) if many else prompt,
                      negative_prompt=[neg_prompt] * len(batch) if many else neg_prompt,
                      image=batch if many else batch[0], num_inference_steps=int(steps),
                      image_guidance_scale=image_guidance_scale, guidance_scale=text_guidance_scale,
                      generator=generator)
    return result.images, any(bool(f) for f in result.nsfw_content_detected)


def get_frames(temp_dir, video_in):
    """Decode to PIL frames at height 512 and <= 30 fps (frames are dropped to cap the rate)."""
    import cv2

    cap = cv2.VideoCapture(str(video_in))
    fps = cap.get(cv2.CAP_PROP_FPS) or 8.0
    keep_every = 1.0
    if fps > 30:
        print("video rate is over 30, resetting to 30")
        keep_every = fps / 30.0
        fps = 30.0
    frames, i, next_keep = [], 0, 0.0
    try:
        while cap.isOpened() and len(frames) < MAX_FRAMES:
            ok, frame = cap.read()
            if not ok:
                break
            if i >= next_keep:
                next_keep += keep_every
                h, w = frame.shape[:2]
                target = frame_height()
                frame = cv2.resize(frame, (max(8, int(round(w * target / h))), target), interpolation=cv2.INTER_AREA)
                frames.append(Image.fromarray(cv2.cvtColor(frame, cv2.COLOR_BGR2RGB)))
            i += 1
    finally:
        cap.release()
    return frames, fps


def create_video(temp_dir, frames, fps):
    import numpy as np

    from .tx2vid import export_to_video
    from ..toolbox.video_helpers import frames_to_video

    path = str(pathlib.Path(temp_dir).joinpath("output.mp4"))
    arrays = [np.asarray(f.convert("RGB")) for f in frames]
    last = None
    for codec in ("h264", "avc1", "mp4v"):
        try:
            frames_to_video(arrays, path, codec, fps=fps)
            import os

            if os.path.getsize(path) > 0:
                return path
        except Exception as e:
            last = e
    raise RuntimeError(f"cannot encode mp4: {last}")
