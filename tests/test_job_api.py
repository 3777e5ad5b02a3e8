"""CPU tests of the job contract: routing table and defaults of
format_args, output packaging, settings precedence, the error taxonomy of the result envelope, and
the tiny random-init SD txt2img 64x64 / 2-step plumbing config from BASELINE.json."""
import asyncio
import base64
import hashlib
import io
import json

import pytest
from PIL import Image

from tests.helpers import decode_artifact_image, png_data_uri


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")


# ------------------------------------------------------------------ routing
def test_routing_table():
    from chiaswarm_b200 import job_arguments as ja

    cb = ja._callbacks()
    assert ja.format_args({"workflow": "txt2audio", "model_name": "suno/bark"})[0] is cb["bark_diffusion_callback"]
    f, a = ja.format_args({"workflow": "txt2audio", "model_name": "cvssp/audioldm"})
    assert f is cb["txt2audio_diffusion_callback"] and a["num_inference_steps"] == 25 and a["prompt"] == ""
    assert a["pipeline_type"].name == "AudioLDMPipeline"
    assert ja.format_args({"workflow": "stitch", "model_name": "x", "jobs": []})[0] is cb["stitch_callback"]
    assert ja.format_args({"workflow": "vid2vid", "model_name": "x"})[0] is cb["model_video_callback"]
    f, a = ja.format_args({"workflow": "txt2vid", "model_name": "zeroscope", "num_images_per_prompt": 4})
    assert f is cb["txt2vid_diffusion_callback"] and "num_images_per_prompt" not in a and a["num_inference_steps"] == 25
    assert ja.format_args({"model_name": "DeepFloyd/IF-I-XL-v1.0"})[0] is cb["diffusion_if_callback"]
    f, a = ja.format_args({"model_name": "runwayml/stable-diffusion-v1-5", "workflow": "txt2img"})
    assert f is cb["diffusion_callback"]
    assert a["num_inference_steps"] == 30 and a["supports_xformers"] is True and a["upscale"] is False
    assert a["scheduler_type"].name == "DPMSolverMultistepScheduler" and a["pipeline_type"].name == "DiffusionPipeline"


def test_sd_size_guard_and_unsupported_args():
    from chiaswarm_b200 import job_arguments as ja

    with pytest.raises(Exception, match="max image size"):
        ja.format_args({"model_name": "m", "height": 2048, "width": 512})
    _, a = ja.format_args({"model_name": "m", "height": 2048})  # only checked when both are present
    assert a["height"] == 2048
    _, a = ja.format_args({"model_name": "m", "eta": 0.3, "parameters": {"unsupported_pipeline_arguments": ["eta"],
                                                                            "scheduler_type": "EulerDiscreteScheduler"}})
    assert "eta" not in a and a["scheduler_type"].kind == "euler"
    with pytest.raises(ValueError):
        ja.format_args({"model_name": "m", "parameters": {"scheduler_type": "NoSuchScheduler"}})


def test_start_image_routes_to_img2img_controlnet_and_pix2pix():
    from chiaswarm_b200 import job_arguments as ja

    uri = png_data_uri((96, 64))
    _, a = ja.format_args({"model_name": "m", "start_image_uri": uri, "height": 64, "width": 64})
    assert a["pipeline_type"].name == "StableDiffusionImg2ImgPipeline"
    assert "height" not in a and "width" not in a and a["image"].size[0] <= 64
    _, a = ja.format_args({"model_name": "m", "start_image_uri": uri,
                           "parameters": {"controlnet": {"type": "canny", "preprocess": True}}})
    assert a["pipeline_type"].name == "StableDiffusionControlNetPipeline"
    assert a["controlnet_model_name"] == "lllyasviel/control_v11p_sd15_canny" and a["save_preprocessed_input"] is True
    assert a["image"].mode == "RGB"
    _, a = ja.format_args({"model_name": "timbrooks/instruct-pix2pix", "start_image_uri": uri, "strength": 0.4})
    assert a["image_guidance_scale"] == pytest.approx(2.0) and "strength" not in a
    _, a = ja.format_args({"model_name": "m", "start_image_uri": uri, "mask_image_uri": uri})
    assert "mask_image" in a


def test_input_guards():
    from chiaswarm_b200 import job_arguments as ja

    with pytest.raises(Exception, match="does not appear to be an image"):
        ja.get_image("data:text/plain;base64," + base64.b64encode(b"hi").decode(), None)
    big = "data:image/png;base64," + base64.b64encode(b"\0" * (3 * 1048576 + 1)).decode()
    with pytest.raises(Exception, match="too large"):
        ja.get_image(big, None)
    im = ja.get_image(png_data_uri((2000, 1500)), None)
    assert max(im.size) <= 1024


# ------------------------------------------------------------------ output packaging
def test_make_result_and_text_result():
    from chiaswarm_b200 import output_processor as op

    img = Image.new("RGB", (300, 200), (1, 2, 3))
    buf = op.image_to_buffer(img, "image/png")
    r = op.make_result(buf, buf, "image/png")
    assert r["content_type"] == "image/png"
    assert r["sha256_hash"] == hashlib.sha256(base64.b64decode(r["blob"])).hexdigest()
    thumb = Image.open(io.BytesIO(base64.b64decode(r["thumbnail"])))
    assert thumb.format == "JPEG" and max(thumb.size) <= 100
    t = op.make_text_result("a cat")
    assert json.loads(base64.b64decode(t["blob"])) == {"caption": "a cat"}
    assert t["sha256_hash"] == hashlib.sha256(b"a cat").hexdigest() and t["content_type"] == "application/json"
    with pytest.raises(ValueError):
        op.image_to_buffer(img, "image/gif")
    with pytest.raises(ValueError):
        op.image_to_buffer(img, "video/mp4")


@pytest.mark.parametrize("count,grid", [(1, (1, 1)), (2, (1, 2)), (3, (2, 2)), (4, (2, 2)), (5, (2, 3)), (6, (2, 3)),
                                        (7, (3, 3)), (9, (3, 3))])
def test_grid_geometry(count, grid):
    from chiaswarm_b200 import output_processor as op

    out = op.post_process([Image.new("RGB", (16, 8), (i * 20, 0, 0)) for i in range(count)])
    assert out.size == (grid[1] * 16, grid[0] * 8)
    if count > 1:
        assert out.getpixel((16 + 1, 1))[0] == 20  # second image sits to the right of the first


def test_grid_limit_and_processor_outputs():
    from chiaswarm_b200 import output_processor as op

    with pytest.raises(ValueError, match="Too many images"):
        op.post_process([Image.new("RGB", (4, 4))] * 10)
    proc = op.OutputProcessor(["primary"], "image/jpeg")
    proc.add_outputs([Image.new("RGB", (32, 32))] * 2)
    proc.add_other_outputs("preprocessed_input", [Image.new("RGB", (8, 8))])
    res = proc.get_results()
    assert set(res) == {"primary", "preprocessed_input"}
    assert decode_artifact_image(res["primary"]).size == (64, 32)
    assert op.OutputProcessor(["other"], "image/jpeg").get_results() == {}


# ------------------------------------------------------------------ settings / logging
def test_settings_precedence(tmp_path, monkeypatch):
    from chiaswarm_b200 import settings as st

    s = st.load_settings()
    assert s.sdaas_uri == "http://localhost:9511" and s.worker_name == "worker" and s.log_level == "WARN"
    assert not st.settings_exist()
    s.sdaas_token, s.sdaas_uri = "tok", "https://hive.example"
    st.save_settings(s)
    assert st.settings_exist()
    assert st.load_settings().sdaas_token == "tok"
    monkeypatch.setenv("SDAAS_TOKEN", "envtok")
    monkeypatch.setenv("SDAAS_WORKERNAME", "b200-0")
    s2 = st.load_settings()
    assert (s2.sdaas_token, s2.worker_name, s2.sdaas_uri) == ("envtok", "b200-0", "https://hive.example")
    st.get_settings_full_path().write_text("{not json")
    assert st.load_settings().sdaas_uri == "http://localhost:9511"
    assert st.resolve_path("log/x.log").parent.is_dir()


def test_logging_levels(tmp_path):
    import logging

    from chiaswarm_b200.log_setup import setup_logging

    h = setup_logging(tmp_path / "a.log", "WARN")
    assert logging.getLogger().level == logging.WARNING
    assert h.maxBytes == 50 * 1024 * 1024 and h.backupCount == 7
    setup_logging(tmp_path / "a.log", "nonsense")
    assert logging.getLogger().level == logging.INFO
    logging.getLogger().removeHandler(logging.getLogger().handlers[-1])


# ------------------------------------------------------------------ envelope / taxonomy / e2e
def _run(job, device=None):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    return asyncio.run(do_work(job, device or Device("cpu")))


def test_tiny_sd_txt2img_end_to_end():
    """BASELINE.json config 1: tiny random-init SD, txt2img 64x64, 2 steps, CPU, through the job API."""
    from chiaswarm_b200 import __version__

    job = {"id": "j1", "model_name": "hf-internal-testing/tiny-stable-diffusion-pipe", "prompt": "a red cube",
           "height": 64, "width": 64, "num_inference_steps": 2, "seed": 1234, "outputs": ["primary"]}
    r = _run(dict(job))
    assert r["id"] == "j1" and r["worker_version"] == __version__ and r["nsfw"] is False
    assert "fatal_error" not in r and "error" not in r["pipeline_config"]
    assert r["pipeline_config"]["seed"] == 1234
    img = decode_artifact_image(r["artifacts"]["primary"])
    assert img.size == (64, 64) and img.format == "JPEG"
    r2 = _run(dict(job))
    assert r2["artifacts"]["primary"]["sha256_hash"] == r["artifacts"]["primary"]["sha256_hash"]  # seeded
    r3 = _run(dict(job, seed=99))
    assert r3["artifacts"]["primary"]["sha256_hash"] != r["artifacts"]["primary"]["sha256_hash"]


def test_tiny_sd_batch_grid_png_img2img_and_schedulers():
    job = {"id": "j2", "model_name": "tiny-sd", "prompt": "x", "height": 64, "width": 64, "num_inference_steps": 3,
           "num_images_per_prompt": 2, "content_type": "image/png", "seed": 5}
    r = _run(job)
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (128, 64)
    for sched in ("EulerAncestralDiscreteScheduler", "EulerDiscreteScheduler", "DDIMScheduler"):
        r = _run({"id": "j3", "model_name": "tiny-sd", "prompt": "x", "start_image_uri": png_data_uri((64, 64)),
                  "strength": 0.7, "num_inference_steps": 4, "seed": 1, "parameters": {"scheduler_type": sched}})
        assert "error" not in r["pipeline_config"], r["pipeline_config"]
        assert decode_artifact_image(r["artifacts"]["primary"]).size == (64, 64)


def test_error_taxonomy():
    # malformed arguments -> fatal, image artifact carries the message
    r = _run({"id": "e1", "model_name": "m", "height": 4096, "width": 64})
    assert r["fatal_error"] is True and "max image size" in r["pipeline_config"]["error"]
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (512, 512)
    # ValueError during generation -> fatal (unknown model cannot be satisfied by resubmitting)
    r = _run({"id": "e2", "model_name": "nobody/unknown-arch"})
    assert r["fatal_error"] is True and "not available" in r["pipeline_config"]["error"]
    # non-image content type -> text artifact
    r = _run({"id": "e3", "model_name": "nobody/unknown-arch", "content_type": "audio/mpeg", "workflow": "txt2audio"})
    assert r["artifacts"]["primary"]["content_type"] == "application/json"

    # any other exception -> error artifact without the fatal flag
    from chiaswarm_b200 import generator

    def boom(device_identifier, model_name, **kw):
        raise RuntimeError("transient")

    orig = generator.format_args
    generator.format_args = lambda job: (boom, {"model_name": "m"})
    try:
        r = _run({"id": "e4"})
    finally:
        generator.format_args = orig
    assert "fatal_error" not in r and r["pipeline_config"]["error"] == "transient"


def test_device_is_exclusive_and_seeds():
    import threading

    from chiaswarm_b200.gpu.device import Device

    d = Device("cpu")
    started, release = threading.Event(), threading.Event()

    def slow(ident, model_name, **kw):
        started.set()
        release.wait(5)
        return {}, {"generator_seed": kw["generator"].initial_seed()}

    out = {}
    t = threading.Thread(target=lambda: out.update(r=d(slow, model_name="m", seed=7)))
    t.start()
    started.wait(5)
    with pytest.raises(Exception, match="busy"):
        d(slow, model_name="m")
    release.set()
    t.join()
    assert out["r"][1] == {"generator_seed": 7, "seed": 7}
    _, cfg = d(lambda i, m, **kw: ({}, {}), model_name="m")
    assert isinstance(cfg["seed"], int)


def test_deferred_packaging_gives_the_same_result_off_the_lane():
    """The serve loop takes results with their image encoding still owed and finishes them on the result thread."""
    from chiaswarm_b200.generator import materialize_result, synchronous_do_work_function
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.output_processor import DeferredArtifact, packaging_deferred

    job = {"id": "d1", "model_name": "tiny-sd", "prompt": "p", "height": 64, "width": 64, "num_inference_steps": 1, "seed": 9,
           "content_type": "image/png"}
    dev = Device("cpu")
    plain = synchronous_do_work_function(dict(job), dev)
    lazy = synchronous_do_work_function(dict(job), dev, defer_packaging=True)
    assert isinstance(lazy["artifacts"]["primary"], DeferredArtifact) and not packaging_deferred()
    assert lazy["pipeline_config"]["seed"] == 9 and lazy["id"] == "d1"
    done = materialize_result(lazy)
    assert "_job" not in done and done == plain                       # same PNG, thumbnail and hash
    assert materialize_result(dict(plain)) == plain                   # results that owe nothing pass through
    # an error that only shows while encoding (more images than the contact grid holds) becomes the same fatal result
    many = dict(job, id="d2", num_images_per_prompt=10)
    want = synchronous_do_work_function(dict(many), dev)
    got = materialize_result(synchronous_do_work_function(dict(many), dev, defer_packaging=True))
    assert want["fatal_error"] is True and "Too many images" in want["pipeline_config"]["error"]
    assert got["fatal_error"] is True and got["pipeline_config"] == want["pipeline_config"] and got["id"] == "d2"
    assert got["artifacts"]["primary"]["content_type"] == "image/png"


def test_device_makes_its_gpu_current_for_the_callback(monkeypatch):
    """One process serving several GPUs (`python -m swarm.worker`): the callback of lane k runs with cuda:k current, so
    the native kernels launch on that GPU's stream, not on device 0's."""
    import contextlib

    import torch

    from chiaswarm_b200.gpu import device as dev_mod

    entered = []

    @contextlib.contextmanager
    def fake_device(idx):
        entered.append(("enter", idx))
        yield
        entered.append(("exit", idx))

    class FakeGen:
        def __init__(self, device=None):
            self.device = device

        def manual_seed(self, s):
            return self

    monkeypatch.setattr(torch.cuda, "mem_get_info", lambda i=None: (1 << 34, 180 << 30))
    monkeypatch.setattr(torch.cuda, "device", fake_device)
    monkeypatch.setattr(torch.cuda, "get_device_name", lambda i=None: "B200")
    monkeypatch.setattr(torch, "Generator", FakeGen)
    d = dev_mod.Device(3)

    def cb(identifier, model_name, **kw):
        assert entered == [("enter", 3)] and identifier == "cuda:3" and kw["generator"].device == "cuda:3"
        return {}, {}

    d(cb, model_name="m", seed=1)
    assert entered == [("enter", 3), ("exit", 3)]
    entered.clear()
    dev_mod.Device("cpu")(lambda identifier, model_name, **kw: ({}, {}), model_name="m", seed=1)
    assert entered == []


def test_jobs_that_differ_only_in_prompt_and_seed_share_one_denoising_run(monkeypatch):
    """generator.synchronous_do_work_batch: compatible jobs run as one batch, each image from its own seed (so each result
    is what the job gives alone); incompatible and malformed jobs in the same call keep their own path and verdict."""
    import numpy as np

    from chiaswarm_b200.diffusion import pipelines
    from chiaswarm_b200.generator import synchronous_do_work_batch, synchronous_do_work_function
    from chiaswarm_b200.gpu.device import Device

    base = {"model_name": "tiny-sd", "height": 64, "width": 64, "num_inference_steps": 2, "content_type": "image/png"}
    jobs = [dict(base, id="a", prompt="a cat", seed=3),
            dict(base, id="wide", prompt="a fox", seed=4, width=32),                      # other shape: runs alone
            dict(base, id="b", prompt="a dog", negative_prompt="", seed=5),               # negative "" vs None: own group
            dict(base, id="bad", prompt="x", height=5000),                                 # malformed: fatal, others unaffected
            dict(base, id="c", prompt="a bird", seed=6),
            dict(base, id="d", prompt="a fish", seed=7, num_images_per_prompt=2)]          # a grid of its own: runs alone
    calls = []
    real = pipelines.StableDiffusionPipeline.__call__

    def spy(self, *a, **k):
        calls.append(k.get("prompt"))
        return real(self, *a, **k)

    monkeypatch.setattr(pipelines.StableDiffusionPipeline, "__call__", spy)
    dev = Device("cpu")
    got = synchronous_do_work_batch([dict(j) for j in jobs], dev)
    assert [r["id"] for r in got] == ["a", "wide", "b", "bad", "c", "d"]
    assert calls.count(["a cat", "a bird"]) == 1 and len(calls) == 4                       # a + c shared one run
    assert got[3]["fatal_error"] is True and "max image size" in got[3]["pipeline_config"]["error"]
    assert [got[i]["pipeline_config"]["seed"] for i in (0, 1, 2, 4, 5)] == [3, 4, 5, 6, 7]
    calls.clear()
    for j, r in zip(jobs, got):
        if j["id"] == "bad":
            continue
        alone = synchronous_do_work_function(dict(j), dev)
        a = np.asarray(decode_artifact_image(alone["artifacts"]["primary"]), dtype=np.int16)
        b = np.asarray(decode_artifact_image(r["artifacts"]["primary"]), dtype=np.int16)
        assert a.shape == b.shape and int(np.abs(a - b).max()) <= 1, j["id"]
        assert set(r["artifacts"]["primary"]) == {"blob", "content_type", "thumbnail", "sha256_hash"}

    # a batch that fails as a whole is re-run job by job: each job still gets its own verdict
    def broken(self, *a, **k):
        if isinstance(k.get("prompt"), list):
            raise RuntimeError("out of memory at this batch size")
        return real(self, *a, **k)

    monkeypatch.setattr(pipelines.StableDiffusionPipeline, "__call__", broken)
    again = synchronous_do_work_batch([dict(jobs[0]), dict(jobs[4])], dev)
    assert [r["id"] for r in again] == ["a", "c"] and all("error" not in r["pipeline_config"] for r in again)


def test_awkward_but_legal_requests_are_answered():
    """Sizes the UNet cannot halve at every level come back at the requested geometry; a missing prompt is the empty
    prompt; an image type the packer cannot write and non-positive sizes are fatal errors WITH an artifact (the error image
    falls back to JPEG) instead of an exception that loses the job."""
    r = _run({"id": "s", "model_name": "tiny-sd", "prompt": "p", "height": 70, "width": 42, "num_inference_steps": 1, "seed": 1})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    assert decode_artifact_image(r["artifacts"]["primary"]).size == (42, 70)
    r = _run({"id": "n", "model_name": "tiny-sd", "prompt": None, "height": 32, "width": 32, "num_inference_steps": 1})
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    r = _run({"id": "w", "model_name": "tiny-sd", "prompt": "p", "height": 32, "width": 32, "num_inference_steps": 1,
              "content_type": "image/webp"})
    assert r["fatal_error"] is True and "Invalid image format" in r["pipeline_config"]["error"]
    assert r["artifacts"]["primary"]["content_type"] == "image/jpeg"
    for bad in ({"height": -32, "width": 32}, {"height": 32, "width": 32, "num_inference_steps": 0}):
        r = _run({"id": "b", "model_name": "tiny-sd", "prompt": "p", "num_inference_steps": 1, **bad})
        assert r["fatal_error"] is True and "must be positive" in r["pipeline_config"]["error"]
