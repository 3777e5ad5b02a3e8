"""Job execution envelope and error taxonomy (reference: swarm/generator.py:12-95).

  * argument errors and ValueError from generation  -> error artifact + "fatal_error": True
    (the hive must not resubmit);
  * any other exception during generation           -> error artifact, no fatal flag (may be retried).
Image jobs get a 512x512 black JPEG/PNG with the message drawn on it; other content types get a
text artifact."""
from __future__ import annotations

import asyncio
import logging

from . import __version__
from .job_arguments import format_args
from .output_processor import (BatchArtifacts, DeferredArtifact, deferring_packaging, image_from_text, image_to_buffer,
                               make_result, make_text_result)


async def do_work(job, device, defer_packaging: bool = False):
    loop = asyncio.get_running_loop()
    return await loop.run_in_executor(None, synchronous_do_work_function, job, device, defer_packaging)


def _error_result(job_id, job, exc, fatal):
    content_type = job.get("content_type", "image/jpeg")
    print(exc)
    if content_type.startswith("image/"):
        artifacts, pipeline_config = exception_image(exc, content_type)
    else:
        artifacts, pipeline_config = exception_message(exc)
    result = {
        "id": job_id,
        "artifacts": artifacts,
        "nsfw": pipeline_config.get("nsfw", False),
        "worker_version": __version__,
        "pipeline_config": pipeline_config,
    }
    if fatal:
        result["fatal_error"] = True
    return result


def synchronous_do_work_function(job, device, defer_packaging: bool = False):
    """`defer_packaging`: image artifacts come back as `DeferredArtifact`s (encoding not done yet); the caller finishes
    them with `materialize_result` -- the serve loop does that on its result thread, off the GPU lane."""
    job_id = job.pop("id")
    print(f"Processing {job_id} on {device.descriptor()}")

    try:
        worker_function, kwargs = format_args(job)
    except Exception as e:  # malformed job: never resubmit
        return _error_result(job_id, job, e, fatal=True)
    return _run_formatted(job_id, job, worker_function, kwargs, device, defer_packaging)


def _envelope(job_id, job, artifacts, pipeline_config, defer_packaging):
    result = {
        "id": job_id,
        "artifacts": artifacts,
        "nsfw": pipeline_config.get("nsfw", False),
        "worker_version": __version__,
        "pipeline_config": pipeline_config,
    }
    if defer_packaging:
        result["_job"] = {"content_type": job.get("content_type", "image/jpeg")}  # for an error raised while encoding
    return result


def _run_formatted(job_id, job, worker_function, kwargs, device, defer_packaging):
    try:
        with deferring_packaging(defer_packaging):
            artifacts, pipeline_config = device(worker_function, **kwargs)
    except ValueError as e:  # generation says the request itself is unsatisfiable
        return _error_result(job_id, job, e, fatal=True)
    except Exception as e:
        logging.exception(e)
        return _error_result(job_id, job, e, fatal=False)
    return _envelope(job_id, job, artifacts, pipeline_config, defer_packaging)


# ---------------------------------------------------------------- several jobs, one denoising run
_VARYING = ("prompt", "negative_prompt", "seed")
_NOT_BATCHABLE = ("image", "mask_image", "controlnet_model_name", "upscale", "latents", "prompt_embeds")


def _batch_key(worker_function, kwargs):
    """Jobs with the same key differ only in prompt, negative prompt and seed and can share one denoising run; None for
    jobs that must run alone (inputs of their own, several images, a post stage, anything but the image callback)."""
    if getattr(worker_function, "__name__", "") != "diffusion_callback":
        return None
    if kwargs.get("num_images_per_prompt", 1) != 1 or any(kwargs.get(k) for k in _NOT_BATCHABLE):
        return None
    if not isinstance(kwargs.get("prompt", ""), str) or not isinstance(kwargs.get("negative_prompt") or "", str):
        return None
    rest = tuple(sorted((k, repr(v)) for k, v in kwargs.items() if k not in _VARYING))
    return (kwargs.get("negative_prompt") is None,) + rest


def synchronous_do_work_batch(jobs, device, defer_packaging: bool = False):
    """Several hive jobs on one lane, results in job order.  Jobs that differ only in prompt / negative prompt / seed are
    denoised as ONE batch (a B200 is launch-bound at batch 1: the step costs little more for four images than for one),
    each image from its own seed -- so each result is what the job would have produced alone; everything else runs one
    after the other as `synchronous_do_work_function` would run it.  A batch that fails is re-run job by job."""
    import torch

    results = [None] * len(jobs)
    groups, order = {}, []
    for i, job in enumerate(jobs):
        job = dict(job)
        job_id = job.pop("id")
        print(f"Processing {job_id} on {device.descriptor()}")
        try:
            worker_function, kwargs = format_args(job)
        except Exception as e:
            results[i] = _error_result(job_id, job, e, fatal=True)
            continue
        key = _batch_key(worker_function, kwargs)
        entry = (i, job_id, job, worker_function, kwargs)
        if key is None:
            order.append([entry])
        elif key in groups:
            groups[key].append(entry)
        else:
            groups[key] = group = [entry]
            order.append(group)
    for group in order:
        done = False
        if len(group) > 1:
            _, _, _, worker_function, first = group[0]
            seeds = [int(kw["seed"]) if kw.get("seed") is not None else torch.seed() for *_, kw in group]
            merged = dict(first, prompt=[kw.get("prompt", "") for *_, kw in group], seed=seeds[0], batch_seeds=seeds)
            if first.get("negative_prompt") is not None:
                merged["negative_prompt"] = [kw.get("negative_prompt") or "" for *_, kw in group]
            try:
                with deferring_packaging(defer_packaging):
                    artifacts, pipeline_config = device(worker_function, **merged)
                if not isinstance(artifacts, BatchArtifacts) or len(artifacts) != len(group):
                    raise RuntimeError("the callback did not answer per job")
                for n, (i, job_id, job, _, _) in enumerate(group):
                    cfg = dict(pipeline_config, seed=seeds[n])
                    cfg.pop("nsfw", None)
                    if artifacts.nsfw[n]:
                        cfg["nsfw"] = True
                    results[i] = _envelope(job_id, job, artifacts[n], cfg, defer_packaging)
                done = True
            except Exception as e:  # (a bad prompt, out of memory at this batch size, ...): each job gets its own verdict
                logging.warning("batch of %d jobs failed (%s): running them one by one", len(group), e)
        if not done:
            for i, job_id, job, worker_function, kwargs in group:
                results[i] = _run_formatted(job_id, job, worker_function, kwargs, device, defer_packaging)
    return results


def materialize_result(result: dict) -> dict:
    """Runs the encoding a deferred result still owes.  An encoding failure becomes the error result the synchronous path
    would have produced (ValueError: fatal, e.g. more images than the grid holds; anything else: retryable)."""
    job = result.pop("_job", None) or {}
    artifacts = result.get("artifacts") or {}
    try:
        for name, value in list(artifacts.items()):
            if isinstance(value, DeferredArtifact):
                artifacts[name] = value.resolve()
    except ValueError as e:
        return _error_result(result.get("id"), job, e, fatal=True)
    except Exception as e:
        logging.exception(e)
        return _error_result(result.get("id"), job, e, fatal=False)
    return result


def _message(e):
    return e.args[0] if len(e.args) > 0 else "error generating image"


def exception_image(e, content_type):
    message = _message(e)
    image = image_from_text(message)
    try:
        buffer = image_to_buffer(image, content_type)
    except ValueError:  # an image type the packer does not write (the very reason the job failed, perhaps): still answer
        content_type = "image/jpeg"
        buffer = image_to_buffer(image, content_type)
    return {"primary": make_result(buffer, buffer, content_type)}, {"error": message}


def exception_message(e):
    return {"primary": make_text_result(str(e))}, {"error": _message(e)}
