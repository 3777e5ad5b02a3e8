"""Checkpoint location and loading for the diffusers / safetensors layout (SURVEY.md section
2.2e; the reference loads through `from_pretrained` per job, swarm/diffusion/diffusion_func.py:30-46,
and warms the HuggingFace cache in swarm/initialize.py:62-94).

Layout consumed: `<root>/model_index.json`, component sub-folders (`unet/`, `vae/`, `text_encoder/`,
`tokenizer/`, `scheduler/`...) each with `config.json` (`scheduler_config.json`) and weights
`diffusion_pytorch_model[.<variant>].safetensors|.bin` or `model[.<variant>].safetensors|
pytorch_model.bin`.  ControlNet repos are flat.  Search order: explicit path, `model_root`
setting, `$SDAAS_MODEL_ROOT`, HuggingFace hub cache (`models--org--name/snapshots/*`).
"""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import Dict, Optional

import torch

_WEIGHT_STEMS = ("diffusion_pytorch_model", "model", "pytorch_model")

# legacy AutoencoderKL attention parameter names -> current ones
_VAE_LEGACY = {"query": "to_q", "key": "to_k", "value": "to_v", "proj_attn": "to_out.0"}


def _hf_cache_dirs():
    dirs = []
    for env in ("HF_HUB_CACHE", "HUGGINGFACE_HUB_CACHE"):
        if os.environ.get(env):
            dirs.append(Path(os.environ[env]))
    if os.environ.get("HF_HOME"):
        dirs.append(Path(os.environ["HF_HOME"]) / "hub")
    dirs.append(Path("~/.cache/huggingface/hub").expanduser())
    return dirs


def locate_checkpoint(model_name: str, revision: str = "main", variant: Optional[str] = None,
                      model_root: str = "") -> Optional[Path]:
    """Directory holding the checkpoint, or None."""
    direct = Path(model_name).expanduser()
    if direct.is_dir() and (any(direct.glob("*.json")) or any(direct.iterdir())):
        return direct
    roots = [r for r in (model_root, os.environ.get("SDAAS_MODEL_ROOT", "")) if r]
    for root in roots:
        for cand in (Path(root) / model_name, Path(root) / model_name.replace("/", "--")):
            if cand.is_dir():
                return cand
    for cache in _hf_cache_dirs():
        repo = cache / ("models--" + model_name.replace("/", "--"))
        snaps = repo / "snapshots"
        if not snaps.is_dir():
            continue
        ref = repo / "refs" / revision
        if ref.is_file():
            cand = snaps / ref.read_text().strip()
            if cand.is_dir():
                return cand
        if (snaps / revision).is_dir():
            return snaps / revision
        found = sorted(p for p in snaps.iterdir() if p.is_dir())
        if found:
            return found[-1]
    return None


def read_json(path: Path) -> dict:
    with open(path, "r") as fh:
        return json.load(fh)


def component_config(root: Path, component: Optional[str]) -> Optional[dict]:
    folder = root / component if component else root
    for name in ("config.json", "scheduler_config.json"):
        if (folder / name).is_file():
            return read_json(folder / name)
    return None


def find_weights(folder: Path, variant: Optional[str] = None) -> Optional[Path]:
    variants = [variant] if variant else []
    variants += [None, "fp16"]
    for v in variants:
        for stem in _WEIGHT_STEMS:
            for ext in (".safetensors", ".bin"):
                p = folder / (f"{stem}.{v}{ext}" if v else f"{stem}{ext}")
                if p.is_file():
                    return p
                # sharded: <stem>.safetensors.index.json (variant: <stem>.safetensors.index.fp16.json) names the shards
                idx = folder / (f"{stem}{ext}.index.{v}.json" if v else f"{stem}{ext}.index.json")
                if idx.is_file():
                    return idx
    return None


def load_state_dict(path: Path) -> Dict[str, torch.Tensor]:
    if path.suffix == ".json":  # a shard index: {"weight_map": {tensor name: shard file}}
        shards = sorted(set(read_json(path)["weight_map"].values()))
        out: Dict[str, torch.Tensor] = {}
        for shard in shards:
            out.update(load_state_dict(path.parent / shard))
        return out
    if path.suffix == ".safetensors":
        from safetensors.torch import load_file

        return load_file(str(path))
    return torch.load(str(path), map_location="cpu", weights_only=True)


def remap_vae_keys(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    out = {}
    for k, v in sd.items():
        parts = k.split(".")
        if "attentions" in parts:
            for old, new in _VAE_LEGACY.items():
                if old in parts:
                    parts[parts.index(old)] = new
                    if v.dim() == 4:  # legacy 1x1 conv form
                        v = v[:, :, 0, 0]
                    break
        out[".".join(parts)] = v
    return out


def load_into(module: torch.nn.Module, sd: Dict[str, torch.Tensor], strict: bool = True):
    own = module.state_dict()
    fixed = {}
    for k, v in sd.items():
        if k in own and own[k].shape != v.shape and own[k].numel() == v.numel():
            v = v.reshape(own[k].shape)  # linear <-> 1x1 conv projections
        fixed[k] = v
    missing, unexpected = module.load_state_dict(fixed, strict=False)
    # position_ids buffers and the like are not parameters here
    unexpected = [k for k in unexpected if not k.endswith("position_ids")]
    if strict and (missing or unexpected):
        raise ValueError(f"checkpoint mismatch: missing={missing[:8]} unexpected={unexpected[:8]}")
    # strict=False tolerates tensors this tree does not model (logged: they may be a feature that is being ignored), never
    # parameters the checkpoint does not provide -- those would run with whatever the constructor left in them
    if missing:
        raise ValueError(f"{type(module).__name__}: the checkpoint has no weights for {len(missing)} parameter(s), e.g. "
                         f"{missing[:6]}")
    if unexpected:
        import logging

        logging.warning("%s: %d checkpoint tensor(s) are not part of this model and were ignored, e.g. %s",
                        type(module).__name__, len(unexpected), unexpected[:6])
    return missing, unexpected


def save_diffusers_layout(root: Path, components: Dict[str, torch.nn.Module], model_index: dict,
                          scheduler_config: Optional[dict] = None):
    """Writes a pipeline in the diffusers layout (used by tests to round-trip the loader)."""
    from safetensors.torch import save_file

    root.mkdir(parents=True, exist_ok=True)
    (root / "model_index.json").write_text(json.dumps(model_index, indent=2))
    for name, module in components.items():
        folder = root / name
        folder.mkdir(exist_ok=True)
        cfg = {k: (list(v) if isinstance(v, tuple) else v) for k, v in module.config.items()}
        (folder / "config.json").write_text(json.dumps(cfg, indent=2))
        stem = "model" if name.startswith("text_encoder") else "diffusion_pytorch_model"
        save_file({k: v.contiguous() for k, v in module.state_dict().items()}, str(folder / f"{stem}.safetensors"))
    if scheduler_config is not None:
        (root / "scheduler").mkdir(exist_ok=True)
        (root / "scheduler" / "scheduler_config.json").write_text(json.dumps(scheduler_config, indent=2))
