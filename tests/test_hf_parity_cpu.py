"""This tree's CLIP modules against the transformers implementations they replace, through real checkpoint files written
by `save_pretrained` (config.json as transformers writes it: `projection_dim` and `architectures` included)."""
import json

import pytest
import torch

from chiaswarm_b200.utils import checkpoints


def _load_ours(folder, ctor):
    cfg = checkpoints.component_config(folder.parent, folder.name)
    m = ctor(**{k: v for k, v in cfg.items() if not k.startswith("_")})
    checkpoints.load_into(m, checkpoints.load_state_dict(checkpoints.find_weights(folder)), strict=False)
    return m.eval()


@pytest.mark.parametrize("act", ["quick_gelu", "gelu"])
def test_clip_text_model_matches_transformers(tmp_path, act):
    from transformers import CLIPTextConfig, CLIPTextModel as HFText, CLIPTextModelWithProjection as HFTextProj

    from chiaswarm_b200.models.clip_text import CLIPTextModel

    cfg = CLIPTextConfig(vocab_size=120, hidden_size=32, intermediate_size=64, num_hidden_layers=3, num_attention_heads=4,
                         max_position_embeddings=16, hidden_act=act, projection_dim=24, eos_token_id=119, bos_token_id=118,
                         pad_token_id=119)
    torch.manual_seed(0)
    ids = torch.randint(0, 118, (2, 16))
    ids[:, 0] = 118
    ids[0, 9:], ids[1, 5:] = 119, 119                                    # EOS, then padding with the EOS id

    hf = HFText(cfg).eval()
    hf.save_pretrained(str(tmp_path / "plain" / "text_encoder"))
    written = json.loads((tmp_path / "plain" / "text_encoder" / "config.json").read_text())
    assert written["projection_dim"] == 24 and written["architectures"] == ["CLIPTextModel"]
    ours = _load_ours(tmp_path / "plain" / "text_encoder", CLIPTextModel)
    assert ours.text_projection is None                                  # a plain CLIPTextModel has no projection matrix
    with torch.no_grad():
        want = hf(ids, output_hidden_states=True)
        last, pooled = ours(ids, eos_token_id=119)
        penult, _ = ours(ids, output_hidden_state=-2, eos_token_id=119)
        pre_norm, _ = ours(ids, output_hidden_state=len(ours.text_model.encoder.layers), eos_token_id=119)
    assert torch.allclose(last, want.last_hidden_state, atol=1e-5)
    assert torch.allclose(pooled, want.pooler_output, atol=1e-5)
    assert torch.allclose(penult, want.hidden_states[-2], atol=1e-5)     # SDXL / clip-skip input
    assert torch.allclose(pre_norm, want.hidden_states[-1], atol=1e-5)   # the latent upscaler's text states

    hfp = HFTextProj(cfg).eval()
    hfp.save_pretrained(str(tmp_path / "proj" / "text_encoder_2"))
    ours_p = _load_ours(tmp_path / "proj" / "text_encoder_2", CLIPTextModel)
    assert ours_p.text_projection is not None
    with torch.no_grad():
        want = hfp(ids, output_hidden_states=True)
        penult, embeds = ours_p(ids, output_hidden_state=-2, eos_token_id=119)
    assert torch.allclose(embeds, want.text_embeds, atol=1e-5) and torch.allclose(penult, want.hidden_states[-2], atol=1e-5)


def test_a_checkpoint_without_weights_for_a_parameter_is_refused(tmp_path):
    from safetensors.torch import load_file, save_file
    from transformers import CLIPTextConfig, CLIPTextModel as HFText

    from chiaswarm_b200.models.clip_text import CLIPTextModel

    cfg = CLIPTextConfig(vocab_size=50, hidden_size=16, intermediate_size=32, num_hidden_layers=1, num_attention_heads=2,
                         max_position_embeddings=8)
    HFText(cfg).save_pretrained(str(tmp_path / "m" / "text_encoder"))
    f = tmp_path / "m" / "text_encoder" / "model.safetensors"
    sd = load_file(str(f))
    sd.pop("text_model.encoder.layers.0.mlp.fc2.bias")
    sd["text_model.something_new.weight"] = torch.zeros(2)              # tolerated (logged), unlike a missing parameter
    save_file(sd, str(f))
    with pytest.raises(ValueError, match="has no weights for 1 parameter"):
        _load_ours(tmp_path / "m" / "text_encoder", CLIPTextModel)


def test_clip_vision_tower_and_image_processor_match_transformers(tmp_path):
    """The safety checker's image tower (models/clip_vision.py) and its pre-processing against transformers."""
    import numpy as np
    from PIL import Image
    from transformers import CLIPImageProcessor, CLIPVisionConfig, CLIPVisionModelWithProjection as HFVision

    from chiaswarm_b200.diffusion.safety_checker import ClipImageProcessor
    from chiaswarm_b200.models.clip_vision import CLIPVisionModelWithProjection

    cfg = CLIPVisionConfig(hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4, image_size=32,
                           patch_size=8, projection_dim=20, hidden_act="quick_gelu")
    torch.manual_seed(1)
    hf = HFVision(cfg).eval()
    hf.save_pretrained(str(tmp_path / "v" / "image_encoder"))
    ours = _load_ours(tmp_path / "v" / "image_encoder", CLIPVisionModelWithProjection)
    rng = np.random.default_rng(0)
    images = [Image.fromarray((rng.random((48, 70, 3)) * 255).astype("uint8")),
              Image.fromarray((rng.random((90, 40, 3)) * 255).astype("uint8"))]
    hf_proc = CLIPImageProcessor(size={"shortest_edge": 32}, crop_size={"height": 32, "width": 32})
    want_px = hf_proc(images=images, return_tensors="pt").pixel_values
    got_px = ClipImageProcessor(size={"shortest_edge": 32}, crop_size={"height": 32, "width": 32})(images)
    assert got_px.shape == want_px.shape == (2, 3, 32, 32)
    assert float((got_px - want_px).abs().max()) < 0.05                  # same resize / crop / normalisation (8-bit rounding)
    with torch.no_grad():
        want = hf(pixel_values=want_px).image_embeds
        got = ours(want_px)
    assert torch.allclose(got, want, atol=1e-5)
