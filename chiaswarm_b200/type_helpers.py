"""Name -> class resolution (reference: swarm/type_helpers.py:1-7).

The reference does `getattr(__import__(module), name)` against diffusers/transformers. diffusers
is replaced by this framework's own pipelines and schedulers, so names from the "diffusers"
namespace resolve through a closed registry; unknown names raise ValueError, which the job
executor reports as a fatal (non-resubmittable) error instead of crashing the worker
(SURVEY.md section 7.3 item 9)."""
from __future__ import annotations

import importlib


def get_type(module_name: str, type_name: str):
    if module_name == "diffusers":
        from .registry import resolve

        return resolve(type_name)
    module = importlib.import_module(module_name)
    try:
        return getattr(module, type_name)
    except AttributeError as e:
        raise ValueError(f"{module_name} has no type named {type_name}") from e


def has_method(o, name: str) -> bool:
    return callable(getattr(o, name, None))
