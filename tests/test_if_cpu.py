"""DeepFloyd-IF cascade through the job API with LOCAL tiny checkpoints (reference: swarm/diffusion/diffusion_func_if.py):
T5 encoder -> IF stage I UNet -> IF stage II UNet (low-resolution image noised to a level, level as class label) ->
SD x4 upscaler.  The IF UNet is also checked block by block against the layout of the diffusers checkpoint it must
load: parameter names, scale-shift modulation, added-KV attention, strict loading."""
import asyncio
import json

import numpy as np
import pytest
import torch

from tests.helpers import decode_artifact_image


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")
    monkeypatch.setenv("SDAAS_IF_SIZES", "16,32")
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    pipeline_cache().clear()


def _save_if_stage(root, seed, **overrides):
    from safetensors.torch import save_file

    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.models.unet_if import IF_TINY_CONFIG, IFUNet

    cfg = dict(IF_TINY_CONFIG, **overrides)
    unet = _init_module(IFUNet(**cfg), seed)
    (root / "unet").mkdir(parents=True)
    (root / "unet" / "config.json").write_text(json.dumps({k: (list(v) if isinstance(v, tuple) else v) for k, v in cfg.items()}))
    save_file({k: v.contiguous() for k, v in unet.state_dict().items()}, str(root / "unet" / "diffusion_pytorch_model.safetensors"))
    (root / "scheduler").mkdir()
    (root / "scheduler" / "scheduler_config.json").write_text(json.dumps({"beta_schedule": "squaredcos_cap_v2", "num_train_timesteps": 1000}))
    (root / "model_index.json").write_text(json.dumps({"_class_name": "IFPipeline"}))
    return unet


def _tiny_cascade(models):
    from tokenizers import Tokenizer, models as tmodels, pre_tokenizers, processors, trainers
    from transformers import PreTrainedTokenizerFast, T5Config, T5EncoderModel

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.utils import checkpoints

    s1 = models / "DeepFloyd" / "IF-I-tiny"
    _save_if_stage(s1, 1)
    tok = Tokenizer(tmodels.Unigram())
    tok.pre_tokenizer = pre_tokenizers.Metaspace()
    tok.train_from_iterator(["a photo of a cat", "a painting of a dog in the rain", "mountains at night"] * 5,
                            trainers.UnigramTrainer(vocab_size=60, special_tokens=["<pad>", "</s>", "<unk>"], unk_token="<unk>"))
    tok.post_processor = processors.TemplateProcessing(single="$A </s>", special_tokens=[("</s>", tok.token_to_id("</s>"))])
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, pad_token="<pad>", eos_token="</s>", unk_token="<unk>")
    fast.save_pretrained(str(s1 / "tokenizer"))
    torch.manual_seed(2)
    T5EncoderModel(T5Config(vocab_size=len(fast), d_model=24, d_kv=8, d_ff=48, num_layers=2, num_heads=3)).save_pretrained(
        str(s1 / "text_encoder"))
    _save_if_stage(models / "DeepFloyd" / "IF-II-L-v1.0", 3, in_channels=6, class_embed_type="timestep")
    x4 = StableDiffusionPipeline.from_architecture("tiny-x4")
    checkpoints.save_diffusers_layout(models / "stabilityai" / "stable-diffusion-x4-upscaler",
                                      {"unet": x4.unet, "vae": x4.vae, "text_encoder": x4.text_encoder},
                                      {"_class_name": "StableDiffusionUpscalePipeline"}, x4.scheduler_config)


def test_if_unet_matches_the_checkpoint_layout_and_loads_strictly(tmp_path):
    from chiaswarm_b200.diffusion.if_unet import IFStage
    from chiaswarm_b200.models.unet_if import IF_TINY_CONFIG, IFUNet

    net = IFUNet(**dict(IF_TINY_CONFIG, in_channels=6, class_embed_type="timestep"))
    names = set(net.state_dict())
    for expected in ("encoder_hid_proj.weight", "add_embedding.pool.positional_embedding", "add_embedding.norm2.bias",
                     "class_embedding.linear_2.weight", "time_embedding.linear_1.weight",
                     "down_blocks.0.downsamplers.0.conv2.weight", "down_blocks.0.resnets.0.time_emb_proj.weight",
                     "down_blocks.1.attentions.0.add_k_proj.weight", "down_blocks.1.attentions.0.group_norm.weight",
                     "down_blocks.1.attentions.0.norm_cross.weight", "down_blocks.1.attentions.0.to_out.0.bias",
                     "mid_block.attentions.0.to_v.weight", "up_blocks.0.upsamplers.0.conv1.weight",
                     "up_blocks.0.attentions.1.to_q.bias", "up_blocks.1.resnets.1.conv_shortcut.weight", "conv_norm_out.weight"):
        assert expected in names, expected
    # scale-shift resnets project the time embedding to 2 x channels
    assert net.down_blocks[1].resnets[0].time_emb_proj.out_features == 2 * 64
    # the stage is conditioned on the noise level: calling it without one is an error, not a silent default
    x, ctx = torch.randn(1, 6, 8, 8), torch.randn(1, 5, 24)
    with pytest.raises(ValueError):
        net(x, torch.tensor(3.0), ctx)
    assert net(x, torch.tensor(3.0), ctx, class_labels=torch.tensor([250])).shape == (1, 6, 8, 8)
    # added-KV attention: the text reaches the output
    with torch.no_grad():
        a = net(x, torch.tensor(3.0), ctx, class_labels=torch.tensor([250]))
        b = net(x, torch.tensor(3.0), ctx * 2 + 1, class_labels=torch.tensor([250]))
    assert (a - b).abs().max() > 1e-6
    # strict loading: a checkpoint with a foreign / missing key is refused as a (fatal) ValueError
    root = tmp_path / "stage"
    _save_if_stage(root, 5)
    assert IFStage.from_checkpoint(root, "cpu").unet.config["out_channels"] == 6
    from safetensors.torch import load_file, save_file

    w = load_file(str(root / "unet" / "diffusion_pytorch_model.safetensors"))
    w["down_blocks.1.attentions.0.to_k_renamed.weight"] = w.pop("down_blocks.1.attentions.0.to_k.weight")
    save_file(w, str(root / "unet" / "diffusion_pytorch_model.safetensors"))
    with pytest.raises(ValueError, match="does not match the IF UNet"):
        IFStage.from_checkpoint(root, "cpu")
    # an SD-style block list is not an IF UNet
    with pytest.raises(ValueError, match="not a DeepFloyd-IF"):
        IFUNet(**dict(IF_TINY_CONFIG, down_block_types=("DownBlock2D", "CrossAttnDownBlock2D")))


def test_if_cascade_job_end_to_end(tmp_path):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    _tiny_cascade(tmp_path / "models")
    job = {"id": "if1", "model_name": "DeepFloyd/IF-I-tiny", "prompt": "a photo of a cat", "negative_prompt": "rain",
           "num_inference_steps": 3, "guidance_scale": 5.0, "seed": 12, "content_type": "image/png"}
    r = asyncio.run(do_work(dict(job), Device("cpu")))
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    img = decode_artifact_image(r["artifacts"]["primary"])
    assert img.size == (64, 64)  # 16 -> 32 -> x2 (the tiny x4 stand-in's VAE factor)
    a = np.asarray(img, dtype=np.float32)
    assert a.std() > 0
    again = np.asarray(decode_artifact_image(asyncio.run(do_work(dict(job), Device("cpu")))["artifacts"]["primary"]), dtype=np.float32)
    assert np.array_equal(a, again)
    other = np.asarray(decode_artifact_image(asyncio.run(do_work(dict(job, prompt="mountains at night"), Device("cpu")))
                                             ["artifacts"]["primary"]), dtype=np.float32)
    assert not np.array_equal(a, other)


def test_if_without_checkpoints_is_a_fatal_job_error():
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    r = asyncio.run(do_work({"id": "if2", "model_name": "DeepFloyd/IF-I-XL-v1.0", "prompt": "x"}, Device("cpu")))
    assert r["fatal_error"] is True and "needs local checkpoints" in r["pipeline_config"]["error"]


def test_ddpm_step_is_the_learned_range_posterior():
    """`ddpm_step` against the DDPM posterior q(x_{t'} | x_t, x0) written out, with the variance interpolated in log space
    between the posterior variance (v = -1) and beta_t (v = +1), and with Imagen's dynamic thresholding."""
    import math

    from chiaswarm_b200.diffusion.if_unet import IF_SCHEDULER, ddpm_step, dynamic_threshold
    from chiaswarm_b200.ops import schedulers

    acp = np.cumprod(1.0 - schedulers._betas(IF_SCHEDULER))
    g = torch.Generator().manual_seed(0)
    x0 = torch.rand(2, 3, 8, 8, generator=g, dtype=torch.float64) * 1.6 - 0.8          # inside [-1, 1]: thresholding is a no-op
    eps = torch.randn(2, 3, 8, 8, generator=g, dtype=torch.float64)
    noise = torch.randn(2, 3, 8, 8, generator=g, dtype=torch.float64)
    t, prev = 600, 590
    a_t, a_p = acp[t], acp[prev]
    x_t = math.sqrt(a_t) * x0 + math.sqrt(1 - a_t) * eps
    beta = 1 - a_t / a_p                                                                # the step's effective beta
    mean = (math.sqrt(a_p) * beta / (1 - a_t)) * x0 + (math.sqrt(a_t / a_p) * (1 - a_p) / (1 - a_t)) * x_t
    post_var = (1 - a_p) / (1 - a_t) * beta
    for v, want_var in ((-1.0, post_var), (1.0, beta), (0.0, math.sqrt(post_var * beta))):   # log-space interpolation
        got = ddpm_step(x_t, eps, torch.full_like(x_t, v), t, prev, acp, noise)
        assert torch.allclose(got, mean + math.sqrt(want_var) * noise, atol=1e-9), v
    assert torch.allclose(ddpm_step(x_t, eps, None, t, prev, acp, noise), mean + math.sqrt(post_var) * noise, atol=1e-9)
    # the last step (prev < 0) lands on the predicted image exactly, and adds no noise
    t0 = 3
    x_last = math.sqrt(acp[t0]) * x0 + math.sqrt(1 - acp[t0]) * eps
    assert torch.allclose(ddpm_step(x_last, eps, torch.zeros_like(x0), t0, t0 - 10, acp, None), x0, atol=1e-9)
    # dynamic thresholding: the 95th percentile of |x| (clamped to [1, 1.5]) becomes the new unit
    big = torch.linspace(-3, 3, 601, dtype=torch.float64).view(1, 1, 1, -1)
    th = dynamic_threshold(big)
    assert float(th.max()) == pytest.approx(1.0) and float(th.min()) == pytest.approx(-1.0)
    assert float(th[0, 0, 0, 300 + 50]) == pytest.approx(0.5 / 1.5)                     # s hit the 1.5 cap; inner values / s
    small = big / 10
    assert torch.allclose(dynamic_threshold(small), small)                              # s floors at 1: untouched


def test_if_stage_samplers(tmp_path):
    from chiaswarm_b200.diffusion.if_unet import IFStage

    root = tmp_path / "models" / "DeepFloyd" / "IF-I-tiny"
    _save_if_stage(root, 1)
    stage = IFStage.from_checkpoint(root, "cpu")
    g = torch.Generator().manual_seed(3)
    cond, unc = torch.randn(1, 5, 24, generator=g), torch.zeros(1, 5, 24)
    kw = dict(size=16, steps=4, guidance_scale=3.0)
    a = stage.sample(cond, unc, generator=torch.Generator().manual_seed(1), **kw)
    b = stage.sample(cond, unc, generator=torch.Generator().manual_seed(1), **kw)
    c = stage.sample(cond, unc, generator=torch.Generator().manual_seed(2), **kw)
    d = stage.sample(cond, unc, generator=torch.Generator().manual_seed(1), sampler="ddim", **kw)
    assert a.shape == (1, 3, 16, 16) and torch.isfinite(a).all() and float(a.abs().max()) <= 1.0
    assert torch.equal(a, b) and not torch.allclose(a, c) and not torch.allclose(a, d)   # ancestral: the seed drives the path
    with pytest.raises(ValueError, match="unknown IF sampler"):
        stage.sample(cond, unc, sampler="euler", **kw)
