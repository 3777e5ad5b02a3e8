"""Safety checker (reference: the `nsfw_content_detected` the diffusers pipeline returns at
swarm/diffusion/diffusion_func.py:97-111).  CPU tests on a tiny CLIP tower: the cosine / threshold logic against
a hand restatement, the blackout, checkpoint round trip through the diffusers folder layout, and the flag's way
to the wire through the job API."""
import json

import numpy as np
import torch
from PIL import Image

from chiaswarm_b200.diffusion.safety_checker import ClipImageProcessor, SafetyChecker, StableDiffusionSafetyChecker

TINY = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4, image_size=28,
            patch_size=14)


def _tiny_checker(seed=0):
    torch.manual_seed(seed)
    m = StableDiffusionSafetyChecker(vision_config=TINY, projection_dim=16)
    for p in m.parameters():
        if p.dim() >= 2:
            torch.nn.init.normal_(p, std=0.2)
    return m


def _images(n, seed=1, size=40):
    rng = np.random.default_rng(seed)
    return [Image.fromarray(rng.integers(0, 255, (size, size + 8, 3), dtype=np.uint8)) for _ in range(n)]


def test_processor_resizes_crops_and_normalises():
    proc = ClipImageProcessor(size=28, crop_size=28)
    x = proc(_images(2))
    assert x.shape == (2, 3, 28, 28)
    white = proc([Image.new("RGB", (64, 48), (255, 255, 255))])
    want = (1.0 - torch.tensor([0.48145466, 0.4578275, 0.40821073])) / torch.tensor([0.26862954, 0.26130258, 0.27577711])
    assert torch.allclose(white[0, :, 5, 5], want, atol=1e-5)


def test_flags_follow_the_cosine_thresholds():
    m = _tiny_checker()
    proc = ClipImageProcessor(size=28, crop_size=28)
    imgs = _images(3)
    pix = proc(imgs)
    with torch.no_grad():
        emb = torch.nn.functional.normalize(m.visual_projection(m.vision_model(pix)).float(), dim=-1)
    # concept 0 := the embedding of image 1, threshold just below a perfect match; everything else unreachable
    m.concept_embeds.data = torch.randn(17, 16)
    m.concept_embeds.data[0] = emb[1] * 3.0
    m.concept_embeds_weights.data.fill_(2.0)
    m.concept_embeds_weights.data[0] = 0.999
    m.special_care_embeds.data = torch.randn(3, 16)
    m.special_care_embeds_weights.data.fill_(2.0)
    flags = m(pix).tolist()
    cos = (emb @ torch.nn.functional.normalize(m.concept_embeds.float(), dim=-1).t())[:, 0]
    assert flags == [bool(c > 0.999) for c in cos.tolist()]
    assert flags[1] and not flags[0]
    # a special-care hit lowers the bar for the ordinary concepts by 0.01
    m.concept_embeds_weights.data[0] = 1.005
    assert not m(pix)[1]
    m.special_care_embeds.data[2] = emb[1]
    m.special_care_embeds_weights.data[2] = 0.5
    assert m(pix)[1]


def test_check_blacks_out_flagged_images():
    m = _tiny_checker()
    chk = SafetyChecker(m, ClipImageProcessor(size=28, crop_size=28))
    imgs = _images(2)
    with torch.no_grad():
        emb = m.visual_projection(m.vision_model(chk.processor(imgs)))
    m.concept_embeds.data = torch.randn(17, 16)
    m.concept_embeds.data[3] = emb[0]
    m.concept_embeds_weights.data.fill_(2.0)
    m.concept_embeds_weights.data[3] = 0.99
    m.special_care_embeds_weights.data.fill_(2.0)
    cleaned, flags = chk.check(imgs)
    assert flags == [True, False]
    assert np.asarray(cleaned[0]).max() == 0 and cleaned[0].size == imgs[0].size
    assert cleaned[1] is imgs[1]


def test_checkpoint_keys_match_the_diffusers_layout(tmp_path):
    m = _tiny_checker()
    keys = set(m.state_dict().keys())
    for k in ("vision_model.vision_model.embeddings.class_embedding",
              "vision_model.vision_model.embeddings.patch_embedding.weight",
              "vision_model.vision_model.embeddings.position_embedding.weight",
              "vision_model.vision_model.pre_layrnorm.weight",
              "vision_model.vision_model.encoder.layers.1.self_attn.q_proj.bias",
              "vision_model.vision_model.encoder.layers.0.mlp.fc1.weight",
              "vision_model.vision_model.post_layernorm.bias", "visual_projection.weight", "concept_embeds",
              "special_care_embeds", "concept_embeds_weights", "special_care_embeds_weights"):
        assert k in keys, k


def test_pipeline_loads_and_runs_the_checker_from_a_checkpoint_folder(tmp_path):
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.utils import checkpoints

    pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
    checker = _tiny_checker()
    checker.concept_embeds_weights.data.fill_(-2.0)  # every cosine beats the threshold: everything is flagged
    root = tmp_path / "tiny-sd-checked"
    index = {"_class_name": "StableDiffusionPipeline", "unet": ["diffusers", "UNet2DConditionModel"],
             "vae": ["diffusers", "AutoencoderKL"], "text_encoder": ["transformers", "CLIPTextModel"],
             "safety_checker": ["stable_diffusion", "StableDiffusionSafetyChecker"],
             "feature_extractor": ["transformers", "CLIPImageProcessor"]}
    checkpoints.save_diffusers_layout(root, {"unet": pipe.unet, "vae": pipe.vae, "text_encoder": pipe.text_encoder,
                                             "safety_checker": checker}, index, pipe.scheduler_config)
    cfg = {"vision_config": TINY, "projection_dim": 16}
    (root / "safety_checker" / "config.json").write_text(json.dumps(cfg))
    (root / "feature_extractor").mkdir()
    (root / "feature_extractor" / "preprocessor_config.json").write_text(
        json.dumps({"size": {"shortest_edge": 28}, "crop_size": {"height": 28, "width": 28}, "do_normalize": True}))
    loaded = StableDiffusionPipeline.from_pretrained(str(root), device="cpu")
    assert loaded.safety_checker is not None
    out = loaded(prompt="a", height=64, width=64, num_inference_steps=1,
                 generator=torch.Generator().manual_seed(0))
    assert out.nsfw_content_detected == [True]
    assert np.asarray(out.images[0]).max() == 0  # blacked out
    assert "nsfw_unchecked" not in loaded.config


def test_missing_checker_is_reported_not_hidden():
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
    out = pipe(prompt="a", height=64, width=64, num_inference_steps=1, generator=torch.Generator().manual_seed(0))
    assert out.nsfw_content_detected == [False]
    assert pipe.config.get("nsfw_unchecked") is True


def test_flag_reaches_the_result_envelope(monkeypatch):
    """A stub checker that flags everything: the job result carries nsfw = True and a black image."""
    import os

    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")
    from chiaswarm_b200 import generator
    from chiaswarm_b200.diffusion import diffusion_func
    from chiaswarm_b200.gpu.device import Device

    class FlagAll:
        def check(self, images):
            return [Image.new("RGB", im.size, (0, 0, 0)) for im in images], [True] * len(images)

    real = diffusion_func.pipeline_cache

    class Cache:
        def get(self, key, load):
            pipe = real().get(key, load)
            pipe.safety_checker = FlagAll()
            return pipe

    monkeypatch.setattr(diffusion_func, "pipeline_cache", lambda: Cache())
    job = {"id": "nsfw-1", "model_name": "hf-internal-testing/tiny-stable-diffusion-pipe", "prompt": "x", "height": 64,
           "width": 64, "num_inference_steps": 1, "seed": 3}
    result = generator.synchronous_do_work_function(job, Device("cpu"))
    assert result["nsfw"] is True
    assert result["pipeline_config"].get("nsfw") is True
