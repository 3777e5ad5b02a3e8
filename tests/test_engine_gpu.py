"""The sm_100a engine (packed weights, recorded plan, CUDA graph) against the eager torch modules
it is built from: same seed, same scheduler, compare latents and decoded images."""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def _pipes(arch, **kw):
    os.environ["SDAAS_RANDOM_INIT"] = "1"
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    ref = StableDiffusionPipeline.from_architecture(arch, **kw)
    ref.backend = "torch"
    ref.to("cuda:0", torch.float16)
    eng = StableDiffusionPipeline(ref.unet, ref.vae, ref.text_encoder, ref.tokenizer, ref.scheduler_config, ref.config,
                                  ref.text_encoder_2, ref.tokenizer_2, ref.controlnet, ref.family, "cuda:0",
                                  torch.float16, backend="sm100")
    return ref, eng


def _compare(ref, eng, steps, tol, **kw):
    outs = []
    for p in (ref, eng):
        p.set_scheduler("dpmpp_2m", True)
        g = torch.Generator(device="cuda:0").manual_seed(11)
        outs.append(p(prompt="a photo", num_inference_steps=steps, generator=g, output_type="latent", **kw).latents)
    a, b = outs[0].float(), outs[1].float()
    assert torch.isfinite(b).all()
    err = (a - b).abs().max().item() / a.abs().max().clamp_min(1e-3).item()
    assert err < tol, f"relative error {err}"


def test_tiny_sd_engine_matches_torch():
    ref, eng = _pipes("tiny-sd")
    _compare(ref, eng, 3, 3e-2, height=64, width=64)
    _compare(ref, eng, 2, 3e-2, height=64, width=64, num_images_per_prompt=2)


def test_sd15_engine_matches_torch_and_decodes():
    ref, eng = _pipes("sd15")
    _compare(ref, eng, 2, 5e-2, height=256, width=256)
    g = torch.Generator(device="cuda:0").manual_seed(3)
    lat = torch.randn(1, 4, 32, 32, generator=g, device="cuda:0", dtype=torch.float16)
    img_ref = ref._denoiser().decode(lat)
    img_ref = ((img_ref.float() / 2 + 0.5).clamp(0, 1) * 255).round().permute(0, 2, 3, 1)
    img_eng = eng._denoiser().decode(lat).float()
    assert img_eng.shape == img_ref.shape
    assert (img_eng - img_ref).abs().mean().item() < 2.0  # uint8 levels


def test_graph_replay_is_deterministic():
    _, eng = _pipes("tiny-sd")
    eng.set_scheduler("euler_a", True)
    res = []
    for _ in range(2):
        g = torch.Generator(device="cuda:0").manual_seed(5)
        res.append(eng(prompt="x", height=64, width=64, num_inference_steps=4, generator=g, output_type="latent").latents)
    assert torch.equal(res[0], res[1])


def test_controlnet_and_img2img_engine_match_torch():
    from PIL import Image

    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.models import configs
    from chiaswarm_b200.models.controlnet import ControlNetModel

    cn = _init_module(ControlNetModel(**configs.get_controlnet("controlnet-tiny")), 9)
    ref, eng = _pipes("tiny-sd", controlnet=cn, family="sd_controlnet")
    img = Image.radial_gradient("L").resize((64, 64)).convert("RGB")
    _compare(ref, eng, 3, 3e-2, image=img)
    ref2, eng2 = _pipes("tiny-sd", family="sd_img2img")
    _compare(ref2, eng2, 4, 3e-2, image=img, strength=0.6)


def test_sdxl_shaped_engine_matches_torch():
    """text_time added embedding, linear projections, transformer depth > 1, two text encoders."""
    import chiaswarm_b200.models.configs as C

    C.ARCHITECTURES["tiny-xl"] = C._pipe(
        "StableDiffusionXLPipeline",
        dict(in_channels=4, out_channels=4, block_out_channels=(64, 128), layers_per_block=1,
             down_block_types=("DownBlock2D", "CrossAttnDownBlock2D"), up_block_types=("CrossAttnUpBlock2D", "UpBlock2D"),
             attention_head_dim=(2, 4), transformer_layers_per_block=(1, 2), cross_attention_dim=96, norm_num_groups=32,
             norm_eps=1e-5, use_linear_projection=True, addition_embed_type="text_time", addition_time_embed_dim=32,
             projection_class_embeddings_input_dim=64 + 6 * 32, sample_size=16),
        C._TINY_VAE, dict(C._TINY_CLIP, hidden_size=32),
        text_encoder_2=dict(C._TINY_CLIP, hidden_size=64, num_attention_heads=4, projection_dim=64))
    ref, eng = _pipes("tiny-xl")
    _compare(ref, eng, 3, 4e-2, height=64, width=64)


def _t2v_pipes(arch, **overrides):
    os.environ["SDAAS_RANDOM_INIT"] = "1"
    from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline

    ref = TextToVideoPipeline.from_architecture(arch, unet_overrides=overrides or None).to("cuda:0", torch.float16)
    ref.backend = "torch"
    eng = TextToVideoPipeline(ref.unet, ref.vae, ref.text_encoder, ref.tokenizer, ref.scheduler_config, ref.config)
    eng.device, eng.dtype, eng.backend = ref.device, ref.dtype, "sm100"
    return ref, eng


@pytest.mark.parametrize("overrides,frames,size", [
    ({}, 4, 64),
    (dict(block_out_channels=(64, 128), attention_head_dim=32, layers_per_block=2), 6, 96)])
def test_unet3d_engine_matches_torch(overrides, frames, size):
    """Text-to-video: the UNet3D plan (frames in the batch dim, temporal convs / frame attention as views)
    against the eager torch UNet3D it was packed from; then the frames decode to uint8 images."""
    ref, eng = _t2v_pipes("tiny-t2v", **overrides)
    outs = []
    for p, kind in ((ref, "dpmpp_2m"), (eng, "dpmpp_2m"), (ref, "euler_a"), (eng, "euler_a")):
        p.set_scheduler(kind, False)
        g = torch.Generator(device="cuda:0").manual_seed(21)
        outs.append(p(prompt="a video", num_frames=frames, height=size, width=size, num_inference_steps=3, generator=g,
                      output_type="latent").latents.float())
    for a, b in ((outs[0], outs[1]),):
        assert a.shape == b.shape and torch.isfinite(b).all()
        err = (a - b).abs().max().item() / a.abs().max().clamp_min(1e-3).item()
        assert err < 4e-2, f"relative error {err}"
    assert torch.isfinite(outs[3]).all() and outs[3].shape == outs[2].shape
    vid = eng(prompt="a video", num_frames=frames, height=size, width=size, num_inference_steps=2)
    assert len(vid.frames) == frames and vid.frames[0].shape == (size, size, 3) and vid.frames[0].dtype.name == "uint8"


def test_layernorm_fusion_matches_torch(monkeypatch):
    """SW_LN_FUSE=1: LayerNorms folded into the consuming projections, statistics from the producing GEMMs."""
    monkeypatch.setenv("SW_LN_FUSE", "1")
    ref, eng = _pipes("sd15")
    _compare(ref, eng, 2, 5e-2, height=256, width=256)
    rec = next(iter(eng._denoiser()._graphs.values())).rec
    assert not any(getattr(fn, "__name__", "") == "layernorm" for fn, _, _ in rec.ops)


@pytest.mark.parametrize("size", [(256, 256), (320, 192)])
def test_vae_encode_engine_matches_torch(size):
    """AutoencoderKL.encode on the recorded sm_100a plan (stride-2 implicit-GEMM downsamplers, chunk-free mid-block
    attention at this size) against the eager module: moments of a structured test image."""
    ref, eng = _pipes("sd15")
    h, w = size
    g = torch.Generator(device="cuda:0").manual_seed(9)
    x = (torch.rand(2, 3, h, w, generator=g, device="cuda:0") * 2 - 1).to(torch.float16)
    want = ref.vae.encode(x, sample=False).float()
    got = eng._vae_encode(x, sample=False).float()
    assert got.shape == want.shape
    assert ((got - want).abs().max() / want.abs().max().clamp_min(1e-3)).item() < 3e-2


def test_vae_attention_in_query_blocks(monkeypatch):
    """The VAE's single-head attention with the score matrix cut into blocks of query rows (what a 1024x1024 decode
    does) gives what the one-block form gives."""
    from chiaswarm_b200.engine import unet_sm100

    ref, eng = _pipes("sd15")
    g = torch.Generator(device="cuda:0").manual_seed(4)
    lat = torch.randn(1, 4, 48, 48, generator=g, device="cuda:0", dtype=torch.float16)
    whole = eng._denoiser().decode(lat).float()
    monkeypatch.setattr(unet_sm100, "SCORE_BYTES", 48 * 48 * 2 * 256)  # 256 query rows per block -> 9 blocks
    eng._engine = None
    blocks = eng._denoiser().decode(lat).float()
    assert (whole - blocks).abs().max().item() <= 1.0  # uint8 levels


@pytest.mark.parametrize("cfg", [dict(), dict(hidden_size=1280, intermediate_size=5120, num_hidden_layers=4,
                                              num_attention_heads=20, hidden_act="gelu", projection_dim=1280)])
def test_clip_text_engine_matches_module(cfg):
    """CLIP text encoder as a recorded sm_100a plan (causal flash attention, quick_gelu folded into fc1) against the
    eager module: last hidden state, penultimate hidden state (SDXL / clip-skip) and pooled output."""
    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.engine.clip_sm100 import ClipTextEngine
    from chiaswarm_b200.models.clip_text import CLIPTextModel

    te = _init_module(CLIPTextModel(**cfg), 21).to("cuda:0", torch.float16).eval()
    with torch.no_grad():
        for prm in te.parameters():  # biases / LayerNorm offsets away from their init values
            if prm.dim() == 1:
                prm.add_(torch.randn_like(prm) * 0.05)
    eng = ClipTextEngine(te, "cuda:0")
    g = torch.Generator(device="cuda:0").manual_seed(2)
    ids = torch.randint(1, 40000, (2, 77), generator=g, device="cuda:0")
    ids[0, 9:] = 49407
    ids[1, 30:] = 49407
    for which in (-1, -2):
        with torch.no_grad():
            want_h, want_p = te.float()(ids, output_hidden_state=which)
        te.half()
        got_h, got_p = eng(ids, output_hidden_state=which)
        assert ((got_h.float() - want_h).abs().max() / want_h.abs().max()).item() < 2e-2
        assert ((got_p.float() - want_p).abs().max() / want_p.abs().max()).item() < 2e-2
    assert eng.last_launches > 0


def test_controlnet_768_batch2_engine_matches_torch():
    """BASELINE config 4 at its real width (SD1.5 UNet + ControlNet, 768x768) with a batch of 2: the engine plan
    (ControlNet encoder, residual adds fused into the zero convs, UNet) against the eager modules."""
    from PIL import Image

    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.models import configs
    from chiaswarm_b200.models.controlnet import ControlNetModel

    cn = _init_module(ControlNetModel(**configs.get_controlnet("controlnet-sd15")), 9)
    ref, eng = _pipes("sd15", controlnet=cn, family="sd_controlnet")
    img = Image.radial_gradient("L").resize((768, 768)).convert("RGB")
    _compare(ref, eng, 2, 5e-2, image=img, num_images_per_prompt=2)


def test_x4_upscaler_engine_matches_torch():
    """x4 upscaler on the engine: 7 input channels (latent + noised low-resolution image as constant extra channels),
    the noise level as a class embedding in the time-embedding slot, text-only attention blocks."""
    from PIL import Image

    ref, eng = _pipes("tiny-x4")
    low = Image.radial_gradient("L").resize((32, 32)).convert("RGB")
    outs = []
    for p in (ref, eng):
        p.set_scheduler("ddim", False)
        g = torch.Generator(device="cuda:0").manual_seed(11)
        outs.append(p(prompt="sharper", image=low, num_inference_steps=3, guidance_scale=5.0, noise_level=100, generator=g,
                      output_type="latent").latents.float())
    err = ((outs[0] - outs[1]).abs().max() / outs[0].abs().max().clamp_min(1e-3)).item()
    assert torch.isfinite(outs[1]).all() and err < 4e-2, err
