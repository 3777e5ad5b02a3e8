"""First-run configuration and model warm-up CLI (reference: swarm/initialize.py:19-120).

`python -m chiaswarm_b200.initialize [--reset] [--silent]` prompts for the hive URI and token,
stores them, fetches `GET {uri}/api/models`, saves `models.json` and warms every pre-loadable
model.  The reference warms the HuggingFace cache by downloading; this worker does not assume a
network: by default it locates the diffusers-layout checkpoints locally and reports what is missing,
with `--download` it pulls the missing repositories into the HuggingFace cache first (configs, tokenizers
and safetensors weights only).  The first-run prompt really is
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
    parser.add_argument("--download", action="store_true",
                        help="fetch models that are not present locally from the HuggingFace hub (needs a network)")
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
    report = await download_diffusers(settings, fetch=args.download)
    print("To be the swarm type 'python -m chiaswarm_b200.worker'")
    return report


async def download_diffusers(settings, fetch: bool = False):
    """Warm every model the hive knows. Returns {model_name: "ok" | "downloaded" | "missing: ..."}.  `fetch`: models that
    are not on the box are pulled into the HuggingFace cache first (the reference's only mode, swarm/initialize.py:62-94:
    it downloads by instantiating every pipeline); without it nothing leaves the machine."""
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
        where = dict(revision=revision, variant=model.get("variant"), model_root=getattr(settings, "model_root", ""))
        path = locate_checkpoint(model_name, **where)
        state = "ok"
        if path is None and fetch:
            error = fetch_from_hub(model_name, revision, getattr(settings, "huggingface_token", None))
            path = locate_checkpoint(model_name, **where)
            state = "downloaded" if path is not None else f"missing: download failed ({error})"
        elif path is None:
            state = "missing: no local diffusers-layout checkpoint"
            print(f"  {model_name}: not present locally (run with --download to fetch it)")
        report[model_name] = state
    if fetch:  # the ControlNet pre-processors' own checkpoints (controlnet/{annotators,pidinet,mlsd,openpose}.py)
        for repo, files in ANNOTATOR_FILES.items():
            have = locate_checkpoint(repo, model_root=getattr(settings, "model_root", ""))
            if have is not None and all((have / f).is_file() for f in (files or ["config.json"])):
                report[repo] = "ok"
                continue
            error = fetch_from_hub(repo, "main", getattr(settings, "huggingface_token", None), patterns=files)
            report[repo] = "downloaded" if error is None else f"missing: download failed ({error})"
    print("Model download complete")
    return report


# the learned ControlNet annotators this tree implements, by repository (file names as published)
ANNOTATOR_FILES = {
    "lllyasviel/Annotators": ["ControlNetHED.pth", "sk_model.pth", "table5_pidinet.pth", "body_pose_model.pth",
                              "mlsd_large_512_fp32.pth", "scannet.pt"],
    # the two annotators that run through transformers (controlnet/input_processor.py: depth, seg): whole repositories
    "Intel/dpt-large": None,
    "openmmlab/upernet-convnext-small": None,
}

# what a worker reads from a model repository: configs, tokenizers, and the safetensors weights (fp16 variant preferred by
# the loaders); the multi-GB .ckpt / .bin duplicates most repositories also carry are left on the hub
HUB_PATTERNS = ["*.json", "*.txt", "*.model", "*.safetensors", "*.pth"]


def fetch_from_hub(model_name: str, revision: str = "main", token=None, patterns=None):
    """Pulls one repository (or the named files of it) into the HuggingFace cache.  Returns None on success, the reason
    otherwise (never raises)."""
    try:
        from huggingface_hub import snapshot_download

        snapshot_download(model_name, revision=revision, allow_patterns=list(patterns or HUB_PATTERNS),
                          token=token if isinstance(token, str) and token else None)
        return None
    except Exception as e:
        print(f"  {model_name}: download failed: {e}")
        return f"{type(e).__name__}: {e}"[:200]


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


def main() -> int:
    """Console entry point (`chiaswarm-init`, `python -m chiaswarm_b200.initialize`)."""
    asyncio.run(init())
    return 0


if __name__ == "__main__":
    main()
