"""Image captioning / VQA callback (reference: swarm/captioning/caption_image.py:6-40).
transformers processor + model classes are named by the job's `parameters`; conditional
(prompted) or unconditional captioning; never raises -- failures become text artifacts."""
from __future__ import annotations

import torch

from ..output_processor import make_text_result
from ..type_helpers import get_type
from ..utils import checkpoints
from ..utils.weight_cache import pipeline_cache


def caption_callback(device_identifier, model_name, **kwargs):
    pipeline_config = {}
    results = {}
    try:
        print("Image captioning...")
        params = kwargs.pop("parameters")
        processor_type = get_type("transformers", params["processor_type"])
        model_type = get_type("transformers", params["model_type"])
        root = checkpoints.locate_checkpoint(model_name)
        source = str(root) if root is not None else model_name
        on_gpu = str(device_identifier).startswith("cuda")
        dtype = torch.float16 if on_gpu else torch.float32
        # resident between jobs (the reference reloads processor and model from disk for every caption)
        processor, model = pipeline_cache().get(
            ("caption", source, params["processor_type"], params["model_type"], str(device_identifier)),
            lambda: (processor_type.from_pretrained(source),
                     model_type.from_pretrained(source, torch_dtype=dtype).to(device_identifier).eval()))
        image = kwargs["image"]
        if kwargs.get("prompt"):
            inputs = processor(image, kwargs["prompt"], return_tensors="pt")
        else:
            inputs = processor(image, return_tensors="pt")
        inputs = inputs.to(device_identifier, dtype)
        with torch.no_grad():
            out = model.generate(**inputs)
        caption = processor.decode(out[0], skip_special_tokens=True)
        results["primary"] = make_text_result(caption)
        pipeline_config["caption"] = caption
    except Exception as e:
        print(e)
        pipeline_config["error"] = str(e)
        results["primary"] = make_text_result(str(e))
    return results, pipeline_config
