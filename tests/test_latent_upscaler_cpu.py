"""The `upscale` job flag (reference: swarm/diffusion/upscale.py, StableDiffusionLatentUpscalePipeline on
stabilityai/sd-x2-latent-upscaler): the K-diffusion UNet's checkpoint layout, its fixed resampling filters against the
dense-weight form diffusers uses, the pipeline's conditioning and sampler against a direct loop, and the whole thing
through the job API with a LOCAL tiny checkpoint (strict loading)."""
import asyncio

import numpy as np
import pytest
import torch
import torch.nn.functional as F
from PIL import Image

from tests.helpers import decode_artifact_image


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")
    monkeypatch.delenv("SDAAS_RANDOM_INIT", raising=False)
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    pipeline_cache().clear()


def _run(job):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    return asyncio.run(do_work(job, Device("cpu")))


def _tiny_pipe(seed=0):
    from chiaswarm_b200.diffusion.upscale import LatentUpscalePipeline

    return LatentUpscalePipeline.from_pretrained("tiny-x2-latent", device="cpu", random_init=True, seed=seed)


def test_k_unet_has_the_layout_of_the_diffusers_checkpoint():
    """Names and shapes of stabilityai/sd-x2-latent-upscaler/unet (a UNet2DConditionModel of K blocks)."""
    from chiaswarm_b200.models import configs
    from chiaswarm_b200.models.unet_k import KUNet

    with torch.device("meta"):
        sd = KUNet(**configs.get_architecture("sd-x2-latent-upscaler")["unet"]).state_dict()
    shape = {k: tuple(v.shape) for k, v in sd.items()}
    expect = {
        "conv_in.weight": (384, 8, 1, 1), "conv_out.weight": (5, 384, 1, 1),
        "time_proj.weight": (384,),                                    # stored Fourier frequencies
        "time_embedding.linear_1.weight": (768, 768), "time_embedding.cond_proj.weight": (768, 896),
        "down_blocks.0.resnets.0.norm1.linear.weight": (768, 768),     # AdaGroupNorm: Linear(temb, 2 * channels)
        "down_blocks.0.resnets.1.conv2.bias": (384,),
        "down_blocks.1.attentions.1.norm2.linear.bias": (768,),
        "down_blocks.1.attentions.0.attn2.norm_cross.weight": (768,),  # LayerNorm on the text, inside the attention
        "down_blocks.1.attentions.0.attn2.to_k.bias": (384,),          # attention_bias=True
        "down_blocks.2.resnets.0.conv_shortcut.weight": (768, 384, 1, 1),
        "down_blocks.2.resnets.3.conv1.weight": (768, 768, 3, 3),      # layers_per_block (2, 2, 4, 4)
        "down_blocks.3.attentions.3.attn1.to_q.weight": (768, 768),    # self-attention only at the deepest level
        "up_blocks.0.resnets.0.conv1.weight": (768, 768, 3, 3),        # deepest up block: nothing concatenated
        "up_blocks.0.attentions.0.attn1.to_out.0.weight": (768, 768),
        "up_blocks.1.resnets.0.conv1.weight": (768, 1536, 3, 3),       # skip concatenated once at the block input
        "up_blocks.1.resnets.3.conv2.weight": (384, 768, 3, 3),        # last resnet narrows to the next level's width
        "up_blocks.1.resnets.3.conv_shortcut.weight": (384, 768, 1, 1),
        "up_blocks.1.attentions.3.attn2.to_q.weight": (384, 384),
        "up_blocks.2.resnets.0.norm1.linear.weight": (1536, 768),
        "up_blocks.3.resnets.1.conv2.weight": (384, 384, 3, 3),
    }
    for k, v in expect.items():
        assert shape.get(k) == v, (k, shape.get(k), v)
    absent = ["down_blocks.0.attentions.0.attn2.to_q.weight", "down_blocks.1.attentions.0.attn1.to_q.weight",
              "down_blocks.2.resnets.0.conv_shortcut.bias", "up_blocks.1.attentions.0.attn1.to_q.weight",
              "up_blocks.3.attentions.0.attn2.to_q.weight", "mid_block.resnets.0.conv1.weight", "conv_norm_out.weight",
              "down_blocks.0.downsamplers.0.kernel", "down_blocks.0.resnets.0.norm1.weight"]
    for k in absent:
        assert k not in shape, k
    assert len([k for k in shape if ".resnets." in k and k.endswith("conv1.weight")]) == 24  # 12 down + 12 up


def test_k_unet_refuses_foreign_configurations():
    from chiaswarm_b200.models.unet_k import K_TINY_CONFIG, KUNet

    with pytest.raises(ValueError, match="not a K-diffusion"):
        KUNet(**dict(K_TINY_CONFIG, down_block_types=("DownBlock2D",) * 4))
    with pytest.raises(ValueError, match="not a K-diffusion"):
        KUNet(**dict(K_TINY_CONFIG, time_embedding_type="positional"))
    m = KUNet(**K_TINY_CONFIG)
    with pytest.raises(ValueError, match="timestep_cond"):
        m(torch.zeros(1, 8, 8, 8), 0.0, torch.zeros(1, 3, 32))


def test_fixed_resampling_filters_match_the_dense_weight_form():
    """diffusers builds KDownsample2D / KUpsample2D as dense convs with the [1 3 3 1]/8 kernel on the channel diagonal."""
    from chiaswarm_b200.models.unet_k import KDownsample2D, KUpsample2D

    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 5, 12, 10, generator=g)
    k1 = torch.tensor([[1 / 8, 3 / 8, 3 / 8, 1 / 8]])

    def dense(kernel):
        w = torch.zeros(5, 5, 4, 4)
        w[torch.arange(5), torch.arange(5)] = kernel
        return w

    down = F.conv2d(F.pad(x, (1,) * 4, "reflect"), dense(k1.T @ k1), stride=2)
    assert torch.allclose(KDownsample2D()(x), down, atol=1e-6) and down.shape[-2:] == (6, 5)
    k2 = k1 * 2
    up = F.conv_transpose2d(F.pad(x, (1,) * 4, "reflect"), dense(k2.T @ k2), stride=2, padding=3)
    assert torch.allclose(KUpsample2D()(x), up, atol=1e-6) and up.shape[-2:] == (24, 20)
    # a constant image stays constant through both (the filters sum to one per output sample)
    ones = torch.ones(1, 5, 8, 8)
    assert torch.allclose(KDownsample2D()(ones), torch.ones(1, 5, 4, 4), atol=1e-6)
    assert torch.allclose(KUpsample2D()(ones), torch.ones(1, 5, 16, 16), atol=1e-6)


def test_ada_group_norm_and_attention_block_semantics():
    from chiaswarm_b200.models.unet_k import AdaGroupNorm, KAttentionBlock

    torch.manual_seed(0)
    n = AdaGroupNorm(16, 64, 2)
    x, emb = torch.randn(2, 64, 4, 4), torch.randn(2, 16)
    scale, shift = n.linear(emb).chunk(2, dim=1)
    want = F.group_norm(x, 2, eps=1e-5) * (1 + scale[:, :, None, None]) + shift[:, :, None, None]
    assert torch.allclose(n(x, emb), want, atol=1e-6)
    blk = KAttentionBlock(64, 4, 16, 24, 16, self_attention=True)
    ctx = torch.randn(2, 5, 24)
    y = blk(x, ctx, emb)
    # by hand: x + attn1(norm1) then + attn2(norm2, LayerNorm(ctx))
    t = lambda v: v.flatten(2).transpose(1, 2)  # noqa: E731
    h = x + blk.attn1(t(blk.norm1(x, emb))).transpose(1, 2).reshape(x.shape)
    a2 = blk.attn2
    q = a2.to_q(t(blk.norm2(h, emb))).view(2, 16, 4, 16).transpose(1, 2)
    c = a2.norm_cross(ctx)
    k, v = a2.to_k(c).view(2, 5, 4, 16).transpose(1, 2), a2.to_v(c).view(2, 5, 4, 16).transpose(1, 2)
    o = torch.softmax(q @ k.transpose(-1, -2) / 4.0, dim=-1) @ v
    want = h + a2.to_out[0](o.transpose(1, 2).reshape(2, 16, 64)).transpose(1, 2).reshape(x.shape)
    assert torch.allclose(y, want, atol=1e-5)


def test_pipeline_matches_a_direct_loop():
    """Conditioning layout, Karras preconditioning and the Euler walk, restated step by step."""
    import math

    from chiaswarm_b200.ops import schedulers

    pipe = _tiny_pipe()
    low = torch.randn(1, 4, 8, 8, generator=torch.Generator().manual_seed(3))
    got = pipe(prompt="a cat", image=low, num_inference_steps=3, generator=torch.Generator().manual_seed(9),
               output_type="latent").images

    ids = pipe.tokenizer(["a cat"])
    tm = pipe.text_encoder.text_model
    h = tm.embeddings(ids)
    for layer in tm.encoder.layers:
        h = layer(h)                                                  # hidden_states[-1]: no final norm
    pooled = tm.final_layer_norm(h)[0, ids.argmax(-1)]                # pooler_output: the normed EOS state
    tc = torch.cat([torch.ones(1, 8), torch.zeros(1, 8), pooled], dim=1)
    assert tc.shape[1] == pipe.unet.config["time_cond_proj_dim"]
    cond = F.interpolate(low, scale_factor=2, mode="nearest")
    betas = np.linspace(0.0001, 0.02, 1000)
    acp = np.cumprod(1 - betas)
    table = np.sqrt((1 - acp) / acp)
    ts = np.linspace(0, 999, 3)[::-1]
    sig = np.concatenate([np.interp(ts, np.arange(1000), table), [0.0]])
    x = torch.randn(1, 4, 16, 16, generator=torch.Generator().manual_seed(9)) * float(sig[0])
    for i in range(3):
        s = float(sig[i])
        f = pipe.unet(torch.cat([x / math.sqrt(s * s + 1), cond], 1), math.log(s) / 4, h, tc)[:, :-1]
        x0 = x / (s * s + 1) + s / math.sqrt(s * s + 1) * f
        x = x + (x - x0) / s * (float(sig[i + 1]) - s)                # Euler step towards the denoised sample
    assert torch.allclose(got, x, atol=2e-4, rtol=1e-4), float((got - x).abs().max())
    plan = schedulers.make_plan("euler", 3, pipe.scheduler_config, False)
    assert np.allclose(plan.sigmas, sig) and plan.prediction_type == "sample"


def test_pipeline_batches_guides_and_checks_inputs():
    pipe = _tiny_pipe()
    im = Image.fromarray((np.random.default_rng(0).random((16, 16, 3)) * 255).astype("uint8"))
    out = pipe(prompt="a cat", image=[im, im.transpose(Image.Transpose.FLIP_LEFT_RIGHT)], num_inference_steps=2,
               generator=torch.Generator().manual_seed(1))
    assert len(out.images) == 2 and all(o.size == (32, 32) for o in out.images)
    assert np.asarray(out.images[0]).tobytes() != np.asarray(out.images[1]).tobytes()
    a = pipe(prompt="a cat", image=im, num_inference_steps=2, generator=torch.Generator().manual_seed(1),
             output_type="latent").images
    b = pipe(prompt="a cat", image=im, num_inference_steps=2, guidance_scale=4.0, negative_prompt="blurry",
             generator=torch.Generator().manual_seed(1), output_type="latent").images
    assert a.shape == b.shape == (1, 4, 16, 16) and torch.isfinite(b).all() and not torch.allclose(a, b)
    with pytest.raises(ValueError, match="batch size"):
        pipe(prompt=["a", "b", "c"], image=[im, im], num_inference_steps=1)
    with pytest.raises(ValueError, match="needs an `image`"):
        pipe(prompt="a")
    with pytest.raises(ValueError, match="cannot be upscaled"):
        pipe(prompt="a", image=torch.zeros(1, 4, 4, 6))
    # sizes the UNet cannot halve at every level are fitted down, not crashed on
    odd = pipe(prompt="a", image=im.resize((23, 19)), num_inference_steps=1).images[0]
    assert odd.size == (32, 32)


def _save_tiny_checkpoint(root, break_unet=False):
    from safetensors.torch import load_file, save_file

    from chiaswarm_b200.utils import checkpoints

    pipe = _tiny_pipe(seed=5)
    checkpoints.save_diffusers_layout(root, {"unet": pipe.unet, "vae": pipe.vae, "text_encoder": pipe.text_encoder},
                                      {"_class_name": "StableDiffusionLatentUpscalePipeline"},
                                      {"_class_name": "EulerDiscreteScheduler", "num_train_timesteps": 1000,
                                       "beta_start": 0.0001, "beta_end": 0.02, "beta_schedule": "linear",
                                       "prediction_type": "sample"})
    if break_unet:
        f = root / "unet" / "diffusion_pytorch_model.safetensors"
        w = load_file(str(f))
        w["mid_block.resnets.0.conv1.weight"] = torch.zeros(1)
        save_file(w, str(f))
    return pipe


def test_upscale_flag_through_the_job_api_with_a_local_checkpoint(tmp_path, monkeypatch):
    from chiaswarm_b200.diffusion import upscale

    src = _save_tiny_checkpoint(tmp_path / "models" / "stabilityai" / "sd-x2-latent-upscaler")
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")  # (for the tiny txt2img model only: the upscaler is on disk)
    job = {"id": "u1", "model_name": "tiny-sd", "prompt": "a castle", "seed": 4, "num_inference_steps": 2,
           "height": 32, "width": 32, "content_type": "image/png", "parameters": {"upscale": True}}
    r = _run(dict(job))
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    big = decode_artifact_image(r["artifacts"]["primary"])
    assert big.size == (64, 64) and r["pipeline_config"]["upscaler"] == "sd-x2-latent-upscaler"
    # the checkpoint on disk is the one that ran: same weights as the pipeline that wrote it
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    cached = pipeline_cache().get((upscale.UPSCALER, "cpu"), lambda: None)
    assert isinstance(cached, upscale.LatentUpscalePipeline) and not cached.config.get("_random_init")
    for k, v in src.unet.state_dict().items():
        assert torch.equal(cached.unet.state_dict()[k], v), k
    # two images per prompt: both come back upscaled (the reference returns only the first)
    r2 = _run(dict(job, id="u2", num_images_per_prompt=2))
    assert "error" not in r2["pipeline_config"], r2["pipeline_config"]
    sheet = decode_artifact_image(r2["artifacts"]["primary"])
    assert sheet.size[0] * sheet.size[1] == 2 * 64 * 64


def test_a_checkpoint_the_k_unet_does_not_describe_is_a_job_error(tmp_path, monkeypatch):
    _save_tiny_checkpoint(tmp_path / "models" / "stabilityai" / "sd-x2-latent-upscaler", break_unet=True)
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")
    r = _run({"id": "u3", "model_name": "tiny-sd", "prompt": "x", "num_inference_steps": 1, "height": 32, "width": 32,
              "parameters": {"upscale": True}})
    assert "does not match the K-diffusion UNet" in r["pipeline_config"].get("error", ""), r["pipeline_config"]


def test_without_a_checkpoint_the_stage_resizes(tmp_path, monkeypatch):
    from chiaswarm_b200.diffusion.upscale import upscale_image

    im = Image.new("RGB", (24, 16), (10, 200, 30))
    info = {}
    out = upscale_image([im, im], "cpu", "x", 2, None, info=info)
    assert [o.size for o in out] == [(48, 32), (48, 32)] and info == {"upscaler": "lanczos"}   # visible in pipeline_config
