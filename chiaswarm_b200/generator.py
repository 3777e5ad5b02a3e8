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
from .output_processor import (DeferredArtifact, deferring_packaging, image_from_text, image_to_buffer, make_result,
                               make_text_result)


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

    try:
        with deferring_packaging(defer_packaging):
            artifacts, pipeline_config = device(worker_function, **kwargs)
    except ValueError as e:  # generation says the request itself is unsatisfiable
        return _error_result(job_id, job, e, fatal=True)
    except Exception as e:
        logging.exception(e)
        return _error_result(job_id, job, e, fatal=False)

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
    buffer = image_to_buffer(image_from_text(message), content_type)
    return {"primary": make_result(buffer, buffer, content_type)}, {"error": message}


def exception_message(e):
    return {"primary": make_text_result(str(e))}, {"error": _message(e)}
