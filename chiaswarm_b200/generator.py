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
from .output_processor import image_from_text, image_to_buffer, make_result, make_text_result


async def do_work(job, device):
    loop = asyncio.get_running_loop()
    return await loop.run_in_executor(None, synchronous_do_work_function, job, device)


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


def synchronous_do_work_function(job, device):
    job_id = job.pop("id")
    print(f"Processing {job_id} on {device.descriptor()}")

    try:
        worker_function, kwargs = format_args(job)
    except Exception as e:  # malformed job: never resubmit
        return _error_result(job_id, job, e, fatal=True)

    try:
        artifacts, pipeline_config = device(worker_function, **kwargs)
    except ValueError as e:  # generation says the request itself is unsatisfiable
        return _error_result(job_id, job, e, fatal=True)
    except Exception as e:
        logging.exception(e)
        return _error_result(job_id, job, e, fatal=False)

    return {
        "id": job_id,
        "artifacts": artifacts,
        "nsfw": pipeline_config.get("nsfw", False),
        "worker_version": __version__,
        "pipeline_config": pipeline_config,
    }


def _message(e):
    return e.args[0] if len(e.args) > 0 else "error generating image"


def exception_image(e, content_type):
    message = _message(e)
    buffer = image_to_buffer(image_from_text(message), content_type)
    return {"primary": make_result(buffer, buffer, content_type)}, {"error": message}


def exception_message(e):
    return {"primary": make_text_result(str(e))}, {"error": _message(e)}
