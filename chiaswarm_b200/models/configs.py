"""Built-in architecture configs.

There is no network on the worker, so the component `config.json` contents of the architectures
named by the benchmark configs (BASELINE.json) ship here; a locally present diffusers-layout
checkpoint always wins over these.  They are recalled from the upstream repos (SURVEY.md section
2.5 -- [UPSTREAM], re-verify against real config.json files when available).  `random_init=True`
pipelines built from them are what the timing runs use (random weights of the named architecture).
"""
from __future__ import annotations

import copy

_SD_SCHED = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                 prediction_type="epsilon", steps_offset=1)

_SD15_UNET = dict(
    in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
    down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
    up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"),
    attention_head_dim=8, cross_attention_dim=768, norm_num_groups=32, norm_eps=1e-5, use_linear_projection=False,
    sample_size=64)
_SD_VAE = dict(in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(128, 256, 512, 512),
               layers_per_block=2, norm_num_groups=32, scaling_factor=0.18215)
_CLIP_L = dict(vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12, num_attention_heads=12,
               max_position_embeddings=77, hidden_act="quick_gelu", layer_norm_eps=1e-5)
_CLIP_H = dict(vocab_size=49408, hidden_size=1024, intermediate_size=4096, num_hidden_layers=23, num_attention_heads=16,
               max_position_embeddings=77, hidden_act="gelu", layer_norm_eps=1e-5)
_CLIP_BIGG = dict(vocab_size=49408, hidden_size=1280, intermediate_size=5120, num_hidden_layers=32,
                  num_attention_heads=20, max_position_embeddings=77, hidden_act="gelu", layer_norm_eps=1e-5,
                  projection_dim=1280)

_SD21_UNET = dict(_SD15_UNET, attention_head_dim=(5, 10, 20, 20), cross_attention_dim=1024, use_linear_projection=True,
                  sample_size=96)

_SDXL_UNET = dict(
    in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280), layers_per_block=2,
    down_block_types=("DownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D"),
    up_block_types=("CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "UpBlock2D"),
    attention_head_dim=(5, 10, 20), transformer_layers_per_block=(1, 2, 10), cross_attention_dim=2048,
    norm_num_groups=32, norm_eps=1e-5, use_linear_projection=True, addition_embed_type="text_time",
    addition_time_embed_dim=256, projection_class_embeddings_input_dim=2816, sample_size=128)

_TINY_UNET = dict(
    in_channels=4, out_channels=4, block_out_channels=(32, 64), layers_per_block=2,
    down_block_types=("DownBlock2D", "CrossAttnDownBlock2D"), up_block_types=("CrossAttnUpBlock2D", "UpBlock2D"),
    attention_head_dim=8, cross_attention_dim=32, norm_num_groups=32, norm_eps=1e-5, use_linear_projection=False,
    sample_size=8)
_TINY_VAE = dict(in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(32, 64), layers_per_block=1,
                 norm_num_groups=32, scaling_factor=0.18215)
_TINY_CLIP = dict(vocab_size=1000, hidden_size=32, intermediate_size=37, num_hidden_layers=5, num_attention_heads=4,
                  max_position_embeddings=77, hidden_act="quick_gelu", layer_norm_eps=1e-5)

_ZEROSCOPE_UNET3D = dict(
    in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
    down_block_types=("CrossAttnDownBlock3D", "CrossAttnDownBlock3D", "CrossAttnDownBlock3D", "DownBlock3D"),
    up_block_types=("UpBlock3D", "CrossAttnUpBlock3D", "CrossAttnUpBlock3D", "CrossAttnUpBlock3D"),
    attention_head_dim=64, cross_attention_dim=1024, norm_num_groups=32, norm_eps=1e-5, sample_size=32)

_TINY_UNET3D = dict(
    in_channels=4, out_channels=4, block_out_channels=(32, 64), layers_per_block=1,
    down_block_types=("CrossAttnDownBlock3D", "DownBlock3D"), up_block_types=("UpBlock3D", "CrossAttnUpBlock3D"),
    attention_head_dim=8, cross_attention_dim=32, norm_num_groups=32, norm_eps=1e-5, sample_size=8)

_CONTROLNET_SD15 = dict(_SD15_UNET, conditioning_channels=3, conditioning_embedding_out_channels=(16, 32, 96, 256))

# stabilityai/sd-x2-latent-upscaler: the K-diffusion UNet of models/unet_k.py.  8 input channels (4 noisy latents + the
# nearest-2x low-resolution latents), 5 outputs (4 + a variance channel the sampler drops), time embedding conditioned on
# [128 noise-level features ; 768 CLIP pooled]; Euler sampler over a linear-beta schedule that predicts the sample
_K_UNET = dict(
    in_channels=8, out_channels=5, block_out_channels=(384, 384, 768, 768), layers_per_block=(2, 2, 4, 4),
    down_block_types=("KDownBlock2D", "KCrossAttnDownBlock2D", "KCrossAttnDownBlock2D", "KCrossAttnDownBlock2D"),
    up_block_types=("KCrossAttnUpBlock2D", "KCrossAttnUpBlock2D", "KCrossAttnUpBlock2D", "KUpBlock2D"),
    mid_block_type=None, attention_head_dim=64, cross_attention_dim=768, only_cross_attention=(True, True, True, False),
    act_fn="gelu", norm_num_groups=None, norm_eps=1e-5, resnet_time_scale_shift="scale_shift",
    time_embedding_type="fourier", timestep_post_act="gelu", time_cond_proj_dim=896, conv_in_kernel=1, conv_out_kernel=1)
_K_SCHED = dict(num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                prediction_type="sample", steps_offset=0)


def _pipe(cls, unet, vae, text, sched=None, **extra):
    d = dict(_class_name=cls, unet=unet, vae=vae, text_encoder=text, scheduler=dict(sched or _SD_SCHED))
    d.update(extra)
    return d


ARCHITECTURES = {
    "tiny-sd": _pipe("StableDiffusionPipeline", _TINY_UNET, _TINY_VAE, _TINY_CLIP),
    "tiny-pix2pix": _pipe("StableDiffusionInstructPix2PixPipeline", dict(_TINY_UNET, in_channels=8), _TINY_VAE, _TINY_CLIP),
    "sd15": _pipe("StableDiffusionPipeline", _SD15_UNET, _SD_VAE, _CLIP_L),
    "sd15-inpaint": _pipe("StableDiffusionInpaintPipeline", dict(_SD15_UNET, in_channels=9), _SD_VAE, _CLIP_L),
    "instruct-pix2pix": _pipe("StableDiffusionInstructPix2PixPipeline", dict(_SD15_UNET, in_channels=8), _SD_VAE, _CLIP_L),
    "sd21": _pipe("StableDiffusionPipeline", _SD21_UNET, _SD_VAE, _CLIP_H,
                  dict(_SD_SCHED, prediction_type="v_prediction")),
    "sd21-base": _pipe("StableDiffusionPipeline", dict(_SD21_UNET, sample_size=64), _SD_VAE, _CLIP_H),
    "sdxl": _pipe("StableDiffusionXLPipeline", _SDXL_UNET, dict(_SD_VAE, scaling_factor=0.13025), _CLIP_L,
                  text_encoder_2=_CLIP_BIGG),
    # the SDXL wiring at toy size: two text encoders (the second with a projection), hidden states concatenated
    # (32 + 48 = cross-attention width 80), pooled vector + six size numbers as the added time embedding (48 + 6 * 8 = 96)
    "tiny-sdxl": _pipe("StableDiffusionXLPipeline",
                       dict(_TINY_UNET, cross_attention_dim=80, use_linear_projection=True, transformer_layers_per_block=(1, 2),
                            addition_embed_type="text_time", addition_time_embed_dim=8, projection_class_embeddings_input_dim=96),
                       dict(_TINY_VAE, scaling_factor=0.13025), _TINY_CLIP,
                       text_encoder_2=dict(_TINY_CLIP, hidden_size=48, num_attention_heads=4, intermediate_size=64,
                                           projection_dim=48, hidden_act="gelu")),
    "zeroscope": _pipe("TextToVideoSDPipeline", _ZEROSCOPE_UNET3D, _SD_VAE, _CLIP_H),
    "tiny-t2v": _pipe("TextToVideoSDPipeline", _TINY_UNET3D, _TINY_VAE, _TINY_CLIP),
    # stabilityai/stable-diffusion-x4-upscaler: 7 input channels (4 latent + the noised low-resolution RGB image), the
    # noise level enters as a class label, first three levels attend to the text only, VAE with a 4x factor
    "sd-x4-upscaler": _pipe("StableDiffusionUpscalePipeline",
                            dict(in_channels=7, out_channels=4, block_out_channels=(256, 512, 512, 1024), layers_per_block=2,
                                 down_block_types=("DownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D"),
                                 up_block_types=("CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "UpBlock2D"),
                                 attention_head_dim=8, cross_attention_dim=1024, norm_num_groups=32, norm_eps=1e-5,
                                 use_linear_projection=True, only_cross_attention=(True, True, True, False),
                                 num_class_embeds=1000, sample_size=128),
                            dict(_SD_VAE, block_out_channels=(128, 256, 512), scaling_factor=0.08333), _CLIP_H,
                            dict(_SD_SCHED, beta_start=0.0001, beta_end=0.02, prediction_type="v_prediction")),
    "tiny-x4": _pipe("StableDiffusionUpscalePipeline",
                     dict(_TINY_UNET, in_channels=7, use_linear_projection=True, only_cross_attention=(False, True),
                          num_class_embeds=1000),
                     dict(_TINY_VAE, scaling_factor=0.08333), _TINY_CLIP,
                     dict(_SD_SCHED, beta_start=0.0001, beta_end=0.02, prediction_type="v_prediction")),
    "sd-x2-latent-upscaler": _pipe("StableDiffusionLatentUpscalePipeline", _K_UNET, _SD_VAE, _CLIP_L, _K_SCHED),
    "tiny-x2-latent": _pipe("StableDiffusionLatentUpscalePipeline",
                            dict(_K_UNET, block_out_channels=(32, 32, 64, 64), layers_per_block=(1, 1, 2, 2),
                                 attention_head_dim=16, cross_attention_dim=32, time_cond_proj_dim=48),
                            _TINY_VAE, _TINY_CLIP, _K_SCHED),
}

_CONTROLNET_TINY = dict(_TINY_UNET, conditioning_channels=3, conditioning_embedding_out_channels=(16, 32))

CONTROLNETS = {"controlnet-sd15": _CONTROLNET_SD15, "controlnet-tiny": _CONTROLNET_TINY}

# model_name (as the hive sends it) -> architecture key
_NAME_RULES = [
    ("tiny-t2v", "tiny-t2v"), ("tiny-text-to-video", "tiny-t2v"),
    ("tiny-stable-diffusion", "tiny-sd"), ("tiny-sd", "tiny-sd"),
    ("controlnet-tiny", "tiny-sd"),
    ("tiny-pix2pix", "tiny-pix2pix"),
    ("instruct-pix2pix", "instruct-pix2pix"),
    ("inpainting", "sd15-inpaint"), ("-inpaint", "sd15-inpaint"),
    ("tiny-sdxl", "tiny-sdxl"), ("stable-diffusion-xl", "sdxl"), ("sdxl", "sdxl"),
    ("tiny-x2-latent", "tiny-x2-latent"), ("sd-x2-latent-upscaler", "sd-x2-latent-upscaler"),
    ("tiny-x4", "tiny-x4"), ("x4-upscaler", "sd-x4-upscaler"),
    ("stable-diffusion-2-1-base", "sd21-base"), ("stable-diffusion-2-base", "sd21-base"),
    ("stable-diffusion-2", "sd21"),
    ("zeroscope", "zeroscope"), ("text-to-video", "zeroscope"), ("modelscope", "zeroscope"),
    ("stable-diffusion-v1", "sd15"), ("sd15", "sd15"), ("sd-v1", "sd15"),
]


def architecture_for(model_name: str):
    """Best-effort architecture key for a hive model name; None when unknown."""
    low = model_name.lower()
    if low in ARCHITECTURES:
        return low
    for needle, key in _NAME_RULES:
        if needle in low:
            return key
    return None


def get_architecture(key: str) -> dict:
    return copy.deepcopy(ARCHITECTURES[key])


def get_controlnet(model_name: str) -> dict:
    key = "controlnet-tiny" if "tiny" in model_name.lower() else "controlnet-sd15"
    return copy.deepcopy(CONTROLNETS[key])
