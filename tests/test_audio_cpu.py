"""txt2audio through the job API with LOCAL tiny checkpoints (reference swarm/audio/audioldm.py,
swarm/audio/bark.py).  The AudioLDM driver is exercised end to end: CLAP text embedding (transformers, random-init tiny
config written with save_pretrained) -> this framework's UNet with the CLAP vector as a class embedding -> this framework's
VAE -> HiFi-GAN vocoder -> WAV / MP3 packaging.  Bark goes through transformers' BarkModel, which is too heavy to build
here even at toy size, so its callback is checked with a stand-in generator for the packaging path and for the error shape."""
import asyncio
import base64
import io
import json

import numpy as np
import pytest
import torch


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")


def _run(job):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    return asyncio.run(do_work(job, Device("cpu")))


def _tiny_audioldm(root):
    from tokenizers import Tokenizer, models, pre_tokenizers, trainers
    from transformers import (ClapTextConfig, ClapTextModelWithProjection, RobertaTokenizerFast, SpeechT5HifiGan,
                              SpeechT5HifiGanConfig)

    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.models.unet2d import UNet2DConditionModel
    from chiaswarm_b200.models.vae import AutoencoderKL
    from chiaswarm_b200.utils import checkpoints

    # the layout of cvssp/audioldm*/unet: the CLAP vector enters as a "simple_projection" class embedding CONCATENATED to the
    # time embedding, there is no cross-attention width (every attn2 is a second self-attention)
    unet = _init_module(UNet2DConditionModel(in_channels=4, out_channels=4, block_out_channels=(32, 64), layers_per_block=1,
                                             down_block_types=("DownBlock2D", "CrossAttnDownBlock2D"),
                                             up_block_types=("CrossAttnUpBlock2D", "UpBlock2D"), attention_head_dim=4,
                                             cross_attention_dim=None, norm_num_groups=32, sample_size=8,
                                             class_embed_type="simple_projection", class_embeddings_concat=True,
                                             projection_class_embeddings_input_dim=16), 1)
    assert unet.class_embedding.weight.shape == (128, 16)
    assert unet.down_blocks[0].resnets[0].time_emb_proj.weight.shape == (32, 256)      # twice the time-embedding width
    assert unet.down_blocks[1].attentions[0].transformer_blocks[0].attn2.to_k.weight.shape == (64, 64)
    with torch.no_grad():
        unet.class_embedding.weight.copy_(torch.randn(128, 16, generator=torch.Generator().manual_seed(5)) * 0.2)
    vae = _init_module(AutoencoderKL(in_channels=1, out_channels=1, latent_channels=4, block_out_channels=(32, 64),
                                     layers_per_block=1, norm_num_groups=32, scaling_factor=0.9), 2)
    checkpoints.save_diffusers_layout(root, {"unet": unet, "vae": vae}, {"_class_name": "AudioLDMPipeline"},
                                      {"num_train_timesteps": 1000, "beta_start": 0.0015, "beta_end": 0.0195,
                                       "beta_schedule": "scaled_linear", "prediction_type": "epsilon"})

    tok = Tokenizer(models.BPE(unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tok.train_from_iterator(["a dog barking in the rain", "soft piano music", "birds and wind"],
                            trainers.BpeTrainer(vocab_size=300, special_tokens=["<s>", "<pad>", "</s>", "<unk>", "<mask>"],
                                                initial_alphabet=pre_tokenizers.ByteLevel.alphabet()))
    (root / "tokenizer").mkdir()
    tok.save(str(root / "tokenizer" / "tokenizer.json"))
    RobertaTokenizerFast(tokenizer_file=str(root / "tokenizer" / "tokenizer.json"), model_max_length=32).save_pretrained(
        str(root / "tokenizer"))
    torch.manual_seed(3)
    ClapTextModelWithProjection(ClapTextConfig(vocab_size=300, hidden_size=32, num_hidden_layers=2, num_attention_heads=2,
                                               intermediate_size=64, max_position_embeddings=40, projection_dim=16,
                                               pad_token_id=1)).save_pretrained(str(root / "text_encoder"))
    SpeechT5HifiGan(SpeechT5HifiGanConfig(model_in_dim=16, sampling_rate=16000, upsample_initial_channel=32,
                                          upsample_rates=[4, 4], upsample_kernel_sizes=[8, 8], resblock_kernel_sizes=[3],
                                          resblock_dilation_sizes=[[1, 3]])).save_pretrained(str(root / "vocoder"))


def _wav(artifact):
    import scipy.io.wavfile

    assert artifact["content_type"] in ("audio/wav", "audio/mpeg")
    raw = base64.b64decode(artifact["blob"])
    if artifact["content_type"] == "audio/mpeg":
        return None, raw
    rate, data = scipy.io.wavfile.read(io.BytesIO(raw))
    return rate, data


def test_audioldm_job_end_to_end(tmp_path):
    _tiny_audioldm(tmp_path / "models" / "cvssp" / "audioldm-tiny")
    job = {"id": "a1", "workflow": "txt2audio", "model_name": "cvssp/audioldm-tiny", "prompt": "a dog barking",
           "num_inference_steps": 2, "audio_length_in_s": 0.064, "seed": 7, "content_type": "audio/wav"}
    r = _run(dict(job))
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert r["pipeline_config"]["_class_name"] == "AudioLDMPipeline" and r["pipeline_config"]["seed"] == 7
    rate, data = _wav(r["artifacts"]["primary"])
    assert rate == 16000 and data.dtype == np.float32
    assert data.shape[0] == int(0.064 * 16000) and np.isfinite(data).all() and float(np.abs(data).max()) > 0
    # same seed, same waveform; another prompt, another waveform (the CLAP vector conditions the UNet)
    _, again = _wav(_run(dict(job))["artifacts"]["primary"])
    assert np.array_equal(data, again)
    _, other = _wav(_run(dict(job, prompt="soft piano music"))["artifacts"]["primary"])
    assert not np.array_equal(data, other)


def test_audioldm_without_checkpoint_is_a_fatal_job_error():
    r = _run({"id": "a2", "workflow": "txt2audio", "model_name": "cvssp/audioldm", "prompt": "x"})
    assert r["fatal_error"] is True and "not available locally" in r["pipeline_config"]["error"]


def test_bark_callback_packages_audio(monkeypatch):
    from chiaswarm_b200.audio import bark

    seen = {}

    def fake_generate(prompt, device, model_name="suno/bark"):
        seen.update(prompt=prompt, device=device, model_name=model_name)
        t = np.arange(2400, dtype=np.float32) / bark.SAMPLE_RATE
        return np.sin(2 * np.pi * 440 * t).astype(np.float32)

    monkeypatch.setattr(bark, "generate_audio", fake_generate)
    r = _run({"id": "b1", "workflow": "txt2audio", "model_name": "suno/bark", "prompt": "hello there",
              "content_type": "audio/wav"})
    assert "error" not in r["pipeline_config"] and seen["prompt"] == "hello there" and seen["model_name"] == "suno/bark"
    rate, data = _wav(r["artifacts"]["primary"])
    assert rate == bark.SAMPLE_RATE and data.shape[0] == 2400


def test_bark_without_checkpoint_is_a_fatal_job_error():
    r = _run({"id": "b2", "workflow": "txt2audio", "model_name": "suno/bark", "prompt": "hi"})
    assert r["fatal_error"] is True and "not available locally" in r["pipeline_config"]["error"]
