"""Workflow routing and argument rationalisation (reference: swarm/job_arguments.py:17-190).

`format_args(job) -> (callback, kwargs)` keeps the reference's routing table, defaults and
guards:

  workflow      callback                         defaults
  txt2audio     bark (model suno/bark) | audioldm steps 25, AudioLDMPipeline
  stitch        stitch                           (raw args)
  img2txt       caption                          start_image_uri -> image
  vid2vid       video pix2pix                    (raw args)
  txt2vid       txt2vid                          steps 25, num_images_per_prompt dropped
  (other)       DeepFloyd IF if model_name starts with "DeepFloyd/", else Stable Diffusion:
                steps 30, max size 1024, start image => img2img / ControlNet, mask => inpaint,
                instruct-pix2pix strength -> image_guidance_scale = strength * 5.

Pipeline / scheduler class names from the hive resolve through the closed registry
(type_helpers.get_type) instead of `getattr(diffusers, name)`.
"""
from __future__ import annotations

from PIL import Image

from . import net
from .type_helpers import get_type

max_size = 1024

DEFAULT_SCHEDULER = "DPMSolverMultistepScheduler"
DEFAULT_CONTROLNET = "lllyasviel/control_v11p_sd15_canny"


def _callbacks():
    # imported lazily: the callbacks pull in torch and the model zoo
    from .audio.audioldm import txt2audio_diffusion_callback
    from .audio.bark import bark_diffusion_callback
    from .captioning.caption_image import caption_callback
    from .diffusion.diffusion_func import diffusion_callback
    from .diffusion.diffusion_func_if import diffusion_if_callback
    from .toolbox.stitch import stitch_callback
    from .video.pix2pix import model_video_callback
    from .video.tx2vid import txt2vid_diffusion_callback

    return locals()


def format_args(job):
    args = job.copy()
    cb = _callbacks()
    workflow = args.pop("workflow", None)

    if workflow == "txt2audio":
        if args["model_name"] == "suno/bark":
            return cb["bark_diffusion_callback"], args
        return format_txt2audio_args(args)
    if workflow == "stitch":
        return cb["stitch_callback"], args
    if workflow == "img2txt":
        return format_img2txt_args(args)
    if workflow == "vid2vid":
        return cb["model_video_callback"], args
    if workflow == "txt2vid":
        return format_txt2vid_args(args)
    if args["model_name"].startswith("DeepFloyd/"):
        return cb["diffusion_if_callback"], args
    return format_stable_diffusion_args(args)


def _resolve_types(args, parameters, default_pipeline):
    args["pipeline_type"] = get_type("diffusers", parameters.pop("pipeline_type", default_pipeline))
    args["scheduler_type"] = get_type("diffusers", parameters.pop("scheduler_type", DEFAULT_SCHEDULER))


def _strip_unsupported(args, parameters):
    for name in parameters.get("unsupported_pipeline_arguments", []):
        args.pop(name, None)


def format_txt2audio_args(args):
    parameters = args.pop("parameters", {})
    args.setdefault("prompt", "")
    args.setdefault("num_inference_steps", 25)
    _resolve_types(args, parameters, "AudioLDMPipeline")
    _strip_unsupported(args, parameters)
    return _callbacks()["txt2audio_diffusion_callback"], args


def format_txt2vid_args(args):
    parameters = args.pop("parameters", {})
    args.setdefault("prompt", "")
    args.setdefault("num_inference_steps", 25)
    args.pop("num_images_per_prompt", None)
    _resolve_types(args, parameters, "DiffusionPipeline")
    return _callbacks()["txt2vid_diffusion_callback"], args


def format_img2txt_args(args):
    # `parameters` stays in args on purpose: the caption callback pops it itself
    if "start_image_uri" in args:
        args["image"] = get_image(args.pop("start_image_uri"), None)
    return _callbacks()["caption_callback"], args


def format_stable_diffusion_args(args):
    size = None
    if "height" in args and "width" in args:
        size = (args["height"], args["width"])
        if size[0] > max_size or size[1] > max_size:
            raise Exception(f"The max image size is (1024, 1024); got ({size[0]}, {size[1]}).")

    parameters = args.pop("parameters", {})
    args.setdefault("prompt", "")
    args["supports_xformers"] = parameters.get("supports_xformers", True)
    args["upscale"] = parameters.get("upscale", False)

    if "start_image_uri" in args:
        args.pop("height", None)
        args.pop("width", None)
        controlnet = parameters.get("controlnet", None)
        args["image"] = get_image(args.pop("start_image_uri"), size, controlnet)
        if controlnet is not None:
            parameters["pipeline_type"] = "StableDiffusionControlNetPipeline"
            args["controlnet_model_name"] = controlnet.get("controlnet_model_name", DEFAULT_CONTROLNET)
            args["save_preprocessed_input"] = controlnet.get("preprocess", False)
        elif "pipeline_type" not in parameters:
            parameters["pipeline_type"] = "StableDiffusionImg2ImgPipeline"
        if args["model_name"] == "timbrooks/instruct-pix2pix":
            # pix2pix takes image_guidance_scale in [1, 5] instead of strength in [0, 1]
            args["image_guidance_scale"] = args.pop("strength", 0.6) * 5

    if "mask_image_uri" in args:
        args.pop("height", None)
        args.pop("width", None)
        args["mask_image"] = get_image(args.pop("mask_image_uri"), size)

    args.setdefault("num_inference_steps", 30)
    _resolve_types(args, parameters, "DiffusionPipeline")
    _strip_unsupported(args, parameters)
    return _callbacks()["diffusion_callback"], args


def download_image(url):
    return net.open_image(url)


def get_image(uri, size, controlnet=None):
    content_type, content_length = net.probe(uri)
    if not content_type.startswith("image"):
        raise Exception(f"Input does not appear to be an image.\nContent type was {content_type}.")
    if int(content_length) > net.MAX_IMAGE_BYTES:
        raise Exception(
            f"Input image too large.\nMax size is {net.MAX_IMAGE_BYTES} bytes.\nImage was {content_length}."
        )
    image = download_image(uri)

    # `size` is (height, width) while PIL.thumbnail takes (width, height): the reference passes it
    # through unswapped (swarm/job_arguments.py:98,181-182); kept for wire compatibility (Q2).
    if size is not None and (image.height > size[0] or image.width > size[1]):
        image.thumbnail(size, Image.Resampling.LANCZOS)
    elif image.height > max_size or image.width > max_size:
        image.thumbnail((max_size, max_size), Image.Resampling.LANCZOS)

    if controlnet is not None:
        from .controlnet.input_processor import preprocess_image

        image = preprocess_image(image, controlnet)
    return image
