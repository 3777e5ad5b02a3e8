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
