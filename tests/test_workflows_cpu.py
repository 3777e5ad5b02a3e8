"""CPU tests of the non-txt2img workflows through the job API: ControlNet (canny), inpaint,
instruct-pix2pix routing, text-to-video, stitch, and the long-tail callbacks' error shapes."""
import asyncio
import base64
import io

import numpy as np
import pytest
from PIL import Image

from tests.helpers import decode_artifact_image, png_data_uri


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    pipeline_cache().clear()


def _run(job):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    return asyncio.run(do_work(job, Device("cpu")))


def _checker(size=(64, 64)):
    arr = np.zeros((size[1], size[0], 3), np.uint8)
    arr[: size[1] // 2, : size[0] // 2] = 255
    arr[size[1] // 2:, size[0] // 2:] = 200
    buf = io.BytesIO()
    Image.fromarray(arr).save(buf, format="PNG")
    return "data:image/png;base64," + base64.b64encode(buf.getvalue()).decode()


def test_controlnet_canny_job_emits_preprocessed_input():
    r = _run({"id": "c1", "model_name": "tiny-sd", "prompt": "edges", "start_image_uri": _checker(), "seed": 2,
              "num_inference_steps": 2,
              "parameters": {"controlnet": {"type": "canny", "preprocess": True,
                                            "controlnet_model_name": "lllyasviel/controlnet-tiny"}}})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert set(r["artifacts"]) == {"primary", "preprocessed_input"}
    edges = np.asarray(decode_artifact_image(r["artifacts"]["preprocessed_input"]).convert("L"))
    assert edges.max() > 200 and edges.mean() < 60  # thin white edges on black
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (64, 64)
    assert "controlnet_annotator" not in r["pipeline_config"]          # canny is exact: nothing stood in
    # a learned annotator that is not on this worker: the stand-in is named in the result, not only in the log
    r = _run({"id": "c1b", "model_name": "tiny-sd", "prompt": "pose", "start_image_uri": _checker(), "seed": 2,
              "num_inference_steps": 1,
              "parameters": {"controlnet": {"type": "openpose", "preprocess": True,
                                            "controlnet_model_name": "lllyasviel/controlnet-tiny"}}})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert r["pipeline_config"]["controlnet_annotator"].startswith("stand-in for openpose: an empty skeleton")


def test_controlnet_changes_the_image():
    base = {"id": "c2", "model_name": "tiny-sd", "prompt": "x", "seed": 3, "num_inference_steps": 2}
    a = _run(dict(base, start_image_uri=_checker(), parameters={"controlnet": {"type": "canny", "preprocess": True,
                                                                               "controlnet_model_name": "controlnet-tiny"}}))
    b = _run(dict(base, height=64, width=64))
    assert a["artifacts"]["primary"]["sha256_hash"] != b["artifacts"]["primary"]["sha256_hash"]


def test_inpaint_legacy_keeps_unmasked_region_close():
    mask = np.zeros((64, 64), np.uint8)
    mask[:, 32:] = 255
    buf = io.BytesIO()
    Image.fromarray(mask).save(buf, format="PNG")
    mask_uri = "data:image/png;base64," + base64.b64encode(buf.getvalue()).decode()
    r = _run({"id": "i1", "model_name": "tiny-sd", "prompt": "fill", "start_image_uri": png_data_uri((64, 64)),
              "mask_image_uri": mask_uri, "seed": 4, "num_inference_steps": 3,
              "parameters": {"pipeline_type": "StableDiffusionInpaintPipelineLegacy"}})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (64, 64)


def test_txt2vid_job_returns_a_video():
    r = _run({"id": "v1", "workflow": "txt2vid", "model_name": "tiny-t2v", "prompt": "waves", "num_frames": 3,
              "num_inference_steps": 2, "height": 32, "width": 32, "seed": 5})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    art = r["artifacts"]["primary"]
    assert art["content_type"] == "video/mp4" and len(base64.b64decode(art["blob"])) > 100
    thumb = Image.open(io.BytesIO(base64.b64decode(art["thumbnail"])))
    assert max(thumb.size) <= 100


def test_stitch_geometry_and_image_map():
    from chiaswarm_b200.toolbox import stitch

    uris = [png_data_uri((200, 100), (i * 40, 0, 0)) for i in range(5)]
    r = _run({"id": "s1", "workflow": "stitch", "model_name": "stitcher",
              "jobs": [{"resultUri": u, "fileName": f"f{i}.jpg", "model_name": f"m{i}"} for i, u in enumerate(uris)]})
    sheet = decode_artifact_image(r["artifacts"]["primary"])
    assert sheet.size == (3 * stitch.thumb_size, 3 * stitch.thumb_size)  # ceil(sqrt(5)) = 3
    amap = r["pipeline_config"]["image_map"]
    assert [a["coords"] for a in amap][:4] == ["0,0,144,144", "144,0,288,144", "288,0,432,144", "0,144,144,288"]
    assert amap[2]["alt"] == "m2" and amap[2]["filename"] == "f2.jpg" and amap[0]["shape"] == "rect"
    thumbs = stitch.resize_images([Image.new("RGB", (200, 100)), Image.new("RGB", (50, 300))])
    assert thumbs[0].size == (144, 72) and thumbs[1].size == (24, 144)


def test_long_tail_callbacks_fail_in_the_reference_error_shape():
    r = _run({"id": "a1", "workflow": "txt2audio", "model_name": "cvssp/audioldm", "prompt": "rain",
              "content_type": "audio/mpeg"})
    assert r["fatal_error"] is True and r["artifacts"]["primary"]["content_type"] == "application/json"
    r = _run({"id": "a2", "workflow": "txt2audio", "model_name": "suno/bark", "prompt": "hi", "content_type": "audio/mpeg"})
    assert r["fatal_error"] is True and "not available" in r["pipeline_config"]["error"]
    r = _run({"id": "if1", "model_name": "DeepFloyd/IF-I-XL-v1.0", "prompt": "frog"})
    assert r["fatal_error"] is True and "DeepFloyd" in r["pipeline_config"]["error"]
    # captioning never raises: the error becomes the caption artifact
    r = _run({"id": "cap", "workflow": "img2txt", "model_name": "Salesforce/blip-image-captioning-base",
              "start_image_uri": png_data_uri((32, 32)),
              "parameters": {"processor_type": "BlipProcessor", "model_type": "BlipForConditionalGeneration"}})
    assert "fatal_error" not in r and "error" in r["pipeline_config"]


def test_controlnet_preprocessors_all_types_produce_rgb():
    from chiaswarm_b200.controlnet.input_processor import preprocess_image

    img = decode_artifact_image({"blob": _checker((96, 64)).split(",")[1]})
    for kind in ("canny", "mlsd", "depth", "normalbae", "seg", "lineart", "openpose", "pix2pix", "scribble", "softedge",
                 "shuffle", "tile"):
        out = preprocess_image(img, {"type": kind, "preprocess": True})
        assert out.mode == "RGB", kind
    assert preprocess_image(img, {"type": "tile", "preprocess": True}).size[1] == 1024
    assert preprocess_image(img, {"type": "canny", "preprocess": False}) is img
    with pytest.raises(Exception, match="Unknown controlnet type"):
        preprocess_image(img, {"type": "nope", "preprocess": True})


def _tiny_video(path, frames=11, size=(64, 48), fps=8.0):
    import cv2

    w = cv2.VideoWriter(str(path), cv2.VideoWriter_fourcc(*"mp4v"), fps, size)
    assert w.isOpened()
    for i in range(frames):
        img = np.zeros((size[1], size[0], 3), np.uint8)
        img[:, : 4 + 5 * i] = (40 + 15 * i, 200 - 10 * i, 90)
        w.write(img)
    w.release()
    return "file://" + str(path)


def _video_frames(artifact, tmp_path):
    import cv2

    out = tmp_path / "result.mp4"
    out.write_bytes(base64.b64decode(artifact["blob"]))
    cap = cv2.VideoCapture(str(out))
    frames = []
    while True:
        ok, f = cap.read()
        if not ok:
            break
        frames.append(f)
    cap.release()
    return frames


def test_vid2vid_job_batches_frames_and_returns_a_video(tmp_path, monkeypatch):
    """Video-to-video (instruct-pix2pix per frame, swarm/video/pix2pix.py): 5 frames go through the pipeline in
    chunks of FRAME_BATCH instead of one call per frame, and every frame comes back at the working height (512 in
    production; 96 here to keep the CPU run short)."""
    from chiaswarm_b200.video import pix2pix

    monkeypatch.setenv("SDAAS_VIDEO_HEIGHT", "96")

    calls = []
    real = pix2pix.img2img

    def spy(pipeline, prompt, tg, ig, image, *a, **k):
        calls.append(len(image) if isinstance(image, (list, tuple)) else 1)
        return real(pipeline, prompt, tg, ig, image, *a, **k)

    monkeypatch.setattr(pix2pix, "img2img", spy)
    monkeypatch.setattr(pix2pix, "FRAME_BATCH", 3)
    uri = _tiny_video(tmp_path / "in.mp4", frames=5)
    r = _run({"id": "vv", "workflow": "vid2vid", "model_name": "tiny-pix2pix", "prompt": "make it night", "video_uri": uri,
              "num_inference_steps": 2, "seed": 4})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert calls == [3, 2]  # five frames in two denoising runs, not five
    art = r["artifacts"]["primary"]
    assert art["content_type"] == "video/mp4"
    frames = _video_frames(art, tmp_path)
    assert len(frames) == 5 and frames[0].shape[:2] == (72, 96)  # 4:3 frames fitted into the height x height box
    assert r["pipeline_config"]["cost"] == 512 * 512 * 2 * 5


def test_vid2vid_guards(tmp_path):
    r = _run({"id": "vg", "workflow": "vid2vid", "model_name": "tiny-pix2pix", "prompt": "x",
              "video_uri": png_data_uri((32, 32))})
    assert "does not appear to be an video" in r["pipeline_config"]["error"]


def test_x4_upscaler_pipeline_matches_a_direct_unet_loop():
    """`StableDiffusionUpscalePipeline` semantics (stage III of the IF cascade): the low-resolution image is noised
    to `noise_level`, rides as three extra UNet input channels, the level enters as a class label, blocks flagged
    only_cross_attention attend to the text; output is 4x (the VAE of this tiny arch: 2x) the input size."""
    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline, pil_to_tensor

    pipe = StableDiffusionPipeline.from_architecture("tiny-x4").to("cpu")
    assert pipe.family == "sd_upscale" and pipe.unet.class_embedding is not None
    blocks = pipe.unet.down_blocks[1].attentions[0].transformer_blocks
    assert blocks[0].only_cross_attention and blocks[0].attn1.to_k.in_features == 32
    pipe.set_scheduler("ddim", False)
    low = Image.radial_gradient("L").resize((16, 16)).convert("RGB")
    g = torch.Generator().manual_seed(3)
    out = pipe(prompt="sharper", image=low, num_inference_steps=3, guidance_scale=4.0, noise_level=100, generator=g)
    assert out.images[0].size == (32, 32)  # tiny VAE factor 2; the real x4 VAE gives 4x
    a = np.asarray(out.images[0], dtype=np.float32)
    # another noise level is another class embedding and another amount of noise on the conditioning image
    g = torch.Generator().manual_seed(3)
    b = np.asarray(pipe(prompt="sharper", image=low, num_inference_steps=3, guidance_scale=4.0, noise_level=900,
                        generator=g).images[0], dtype=np.float32)
    assert np.abs(a - b).mean() > 0.5
    # a tensor input (what the IF cascade hands over) gives the same image as the PIL input
    g = torch.Generator().manual_seed(3)
    c = np.asarray(pipe(prompt="sharper", image=pil_to_tensor(low, "cpu", torch.float32), num_inference_steps=3,
                        guidance_scale=4.0, noise_level=100, generator=g).images[0], dtype=np.float32)
    assert np.abs(a - c).max() <= 1.0


def test_hed_annotator_loads_the_reference_checkpoint_layout(tmp_path, monkeypatch):
    """The in-tree HED network has the parameter names of `lllyasviel/Annotators/ControlNetHED.pth`, is picked up from
    the local model root by the soft-edge / scribble pre-processors, and returns an RGB edge map of the input size."""
    import torch

    from chiaswarm_b200.controlnet import annotators, input_processor

    net = annotators.HEDNet()
    names = set(net.state_dict())
    assert {"norm", "block1.convs.0.weight", "block1.convs.1.bias", "block3.convs.2.weight", "block5.projection.weight"} <= names
    assert len(names) == 1 + 2 * (2 + 2 + 3 + 3 + 3) + 2 * 5
    torch.manual_seed(0)
    root = tmp_path / "models" / "lllyasviel" / "Annotators"
    root.mkdir(parents=True)
    torch.save(net.state_dict(), root / "ControlNetHED.pth")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    annotators._cache.clear()
    img = Image.radial_gradient("L").resize((96, 64)).convert("RGB")
    soft = input_processor.preprocess_image(img, {"preprocess": True, "type": "softedge"})
    scr = input_processor.preprocess_image(img, {"preprocess": True, "type": "scribble"})
    assert annotators._cache["hed"] is not None  # the learned detector ran, not the Sobel stand-in
    assert soft.size == img.size and soft.mode == "RGB" and scr.size == img.size
    assert set(np.unique(np.asarray(scr))) <= {0, 255}
    annotators._cache.clear()


def test_lineart_annotator_loads_the_reference_checkpoint_layout(tmp_path, monkeypatch):
    """The in-tree lineart generator has the parameter names of `lllyasviel/Annotators/sk_model.pth` (informative-drawings
    Generator(3, 1, 3)), is picked up from the local model root, and returns white-on-black lines of the input size."""
    import torch

    from chiaswarm_b200.controlnet import annotators, input_processor

    net = annotators.LineartNet()
    sd = net.state_dict()
    expect = {"model0.1.weight": (64, 3, 7, 7), "model1.0.weight": (128, 64, 3, 3), "model1.3.weight": (256, 128, 3, 3),
              "model2.0.conv_block.1.weight": (256, 256, 3, 3), "model2.2.conv_block.5.bias": (256,),
              "model3.0.weight": (256, 128, 3, 3), "model3.3.weight": (128, 64, 3, 3), "model4.1.weight": (1, 64, 7, 7)}
    for k, shape in expect.items():
        assert tuple(sd[k].shape) == shape, k
    assert len(sd) == 2 * (1 + 2 + 2 * 3 + 2 + 1)  # instance norms carry no parameters
    torch.manual_seed(0)
    root = tmp_path / "models" / "lllyasviel" / "Annotators"
    root.mkdir(parents=True)
    torch.save(sd, root / "sk_model.pth")
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    annotators._cache.clear()
    img = Image.radial_gradient("L").resize((98, 63)).convert("RGB")  # not a multiple of 4: fitted, then restored
    out = input_processor.preprocess_image(img, {"preprocess": True, "type": "lineart"})
    assert annotators._cache["lineart"] is not None  # the learned generator ran, not the Sobel stand-in
    assert out.size == img.size and out.mode == "RGB"
    # a file that is not this network is refused, not half-loaded
    annotators._cache.clear()
    torch.save({"model0.1.weight": torch.zeros(64, 3, 7, 7)}, root / "sk_model.pth")
    with pytest.raises(RuntimeError):
        annotators.load_lineart()
    annotators._cache.clear()


def test_eta_and_prompt_embeds_are_honoured():
    """Two pipeline arguments that used to be accepted and dropped: DDIM `eta` and precomputed `prompt_embeds`."""
    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.registry import resolve

    pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
    kw = dict(prompt="a red cube", height=32, width=32, num_inference_steps=3, output_type="latent")
    gen = lambda: torch.Generator().manual_seed(11)  # noqa: E731
    pipe.set_scheduler(resolve("DDIMScheduler"), use_karras_sigmas=False)
    det = pipe(generator=gen(), **kw).latents
    sto = pipe(generator=gen(), eta=1.0, **kw).latents
    assert torch.equal(det, pipe(generator=gen(), **kw).latents)
    assert torch.equal(sto, pipe(generator=gen(), eta=1.0, **kw).latents) and not torch.allclose(det, sto)
    pipe.set_scheduler(resolve("DPMSolverMultistepScheduler"))
    base = pipe(generator=gen(), **kw).latents
    assert torch.equal(base, pipe(generator=gen(), eta=1.0, **kw).latents)  # eta is DDIM's knob only
    cond, _ = pipe.encode_prompt(["a red cube"], [""], 1)
    via_embeds = pipe(generator=gen(), prompt_embeds=cond, **dict(kw, prompt="")).latents
    assert torch.allclose(via_embeds, base, atol=1e-6)
    other, _ = pipe.encode_prompt(["a blue sphere"], [""], 1)
    assert not torch.allclose(pipe(generator=gen(), prompt_embeds=other, **dict(kw, prompt="")).latents, base)
    import pytest as _pytest

    with _pytest.raises(ValueError, match="do not fit"):
        pipe(generator=gen(), prompt_embeds=cond[:, :5], **kw)


def test_force_zeros_for_empty_prompt_is_honoured():
    """SDXL's model_index flag: no negative prompt -> zero unconditional text states (not the encoding of "")."""
    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    pipe = StableDiffusionPipeline.from_architecture("tiny-sd")
    kw = dict(prompt="a red cube", height=32, width=32, num_inference_steps=2, output_type="latent")
    gen = lambda: torch.Generator().manual_seed(5)  # noqa: E731
    plain = pipe(generator=gen(), **kw).latents
    seen = {}
    real_run = None

    pipe.config["force_zeros_for_empty_prompt"] = True
    den = pipe._denoiser()
    real_run = den.run

    def spy(**k):
        seen["embeds"] = k["embeds"].clone()
        return real_run(**k)

    den.run = spy
    zeroed = pipe(generator=gen(), **kw).latents
    assert not seen["embeds"][:1].any() and seen["embeds"][1:].any()      # [uncond = 0, cond]
    assert not torch.allclose(zeroed, plain)
    with_negative = pipe(generator=gen(), negative_prompt="", **kw).latents  # an explicit negative prompt is encoded
    assert seen["embeds"][:1].any() and torch.allclose(with_negative, plain, atol=1e-6)


def test_vae_force_upcast_runs_the_vae_in_fp32():
    """`force_upcast` (vae/config.json of the original SDXL VAE): an fp16 pipeline decodes / encodes through an fp32 copy."""
    import torch

    from chiaswarm_b200.diffusion import pipelines as P

    pipe = P.StableDiffusionPipeline.from_architecture("tiny-sd")
    assert not P.vae_needs_fp32(pipe)
    pipe.vae.config["force_upcast"] = True
    assert not P.vae_needs_fp32(pipe)          # an fp32 pipeline (CPU) needs no second copy
    pipe.dtype = torch.float16                 # what a CUDA pipeline runs at
    assert P.vae_needs_fp32(pipe)
    z = torch.randn(1, 4, 8, 8, generator=torch.Generator().manual_seed(0))
    want = pipe.vae.decode(z)
    got = P.TorchDenoiser(pipe).decode(z.half())
    assert got.dtype == torch.float32 and torch.allclose(got, pipe.vae.decode(z.half().float()), atol=1e-6)
    assert torch.allclose(got, want, atol=5e-2)
    copy = pipe.__dict__["_vae_fp32"]
    assert copy is not pipe.vae and next(copy.parameters()).dtype == torch.float32
    img = torch.rand(1, 3, 16, 16) * 2 - 1
    enc = pipe._vae_encode(img.half(), sample=False)
    assert enc.dtype == torch.float16 and torch.allclose(enc.float(), pipe.vae.encode(img.half().float(), None, False), atol=2e-3)
    pipe.to("cpu", torch.float32)
    assert "_vae_fp32" not in pipe.__dict__


def test_sdxl_wiring_at_toy_size():
    """`tiny-sdxl`: two text encoders (penultimate states concatenated, pooled vector from the projecting second one) and
    the size micro-conditioning reach the UNet; SDXL's force_zeros_for_empty_prompt zeroes states AND pooled vector."""
    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    r = _run({"id": "xl", "model_name": "tiny-sdxl", "prompt": "a lighthouse", "height": 64, "width": 32, "seed": 5,
              "num_inference_steps": 2})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (32, 64)
    assert r["pipeline_config"]["_class_name"] == "StableDiffusionXLPipeline"

    pipe = StableDiffusionPipeline.from_architecture("tiny-sdxl")
    seen = {}
    den = pipe._denoiser()
    real = den.run

    def spy(**k):
        seen.update(embeds=k["embeds"].clone(), added={n: v.clone() for n, v in k["added"].items()})
        return real(**k)

    den.run = spy
    kw = dict(prompt="a lighthouse", num_inference_steps=1, output_type="latent")
    gen = lambda: torch.Generator().manual_seed(2)  # noqa: E731
    a = pipe(height=64, width=32, generator=gen(), **kw).latents
    assert seen["embeds"].shape == (2, 77, 80)                                     # [uncond, cond] x (32 + 48)
    assert seen["added"]["text_embeds"].shape == (2, 48)
    assert seen["added"]["time_ids"].tolist() == [[64.0, 32.0, 0.0, 0.0, 64.0, 32.0]] * 2
    h1, _ = pipe.text_encoder(pipe.tokenizer(["a lighthouse"]), output_hidden_state=-2)
    h2, pooled = pipe.text_encoder_2(pipe.tokenizer_2(["a lighthouse"]), output_hidden_state=-2)
    assert torch.allclose(seen["embeds"][1], torch.cat([h1, h2], -1)[0], atol=1e-6)
    assert torch.allclose(seen["added"]["text_embeds"][1], pooled[0], atol=1e-6)
    # the size numbers condition the UNet: same latent shape, other micro-conditioning, other result
    pipe.unet.add_embedding.linear_2.weight.data.normal_(0, 0.5)                   # (random init leaves this path faint)
    a = pipe(height=64, width=32, generator=gen(), **kw).latents
    ids = seen["added"]["time_ids"].clone()
    real_embed = pipe.unet.time_embed

    def other_size(t, batch, dtype, added_cond=None, class_labels=None):
        return real_embed(t, batch, dtype, dict(added_cond, time_ids=ids * 2), class_labels)

    pipe.unet.time_embed = other_size
    b = pipe(height=64, width=32, generator=gen(), **kw).latents
    pipe.unet.time_embed = real_embed
    assert a.shape == b.shape and not torch.allclose(a, b)
    pipe.config["force_zeros_for_empty_prompt"] = True
    pipe(height=64, width=32, generator=gen(), **kw)
    assert not seen["embeds"][0].any() and not seen["added"]["text_embeds"][0].any()
    assert seen["embeds"][1].any() and seen["added"]["text_embeds"][1].any()
