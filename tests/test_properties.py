"""Property tests (hypothesis) of the host-side algebra the GPU paths rest on: the gang
allocator's invariants under random lease / release / failure sequences, the gang-size policy, the weight rewrites that
delete kernels from the plan (nearest-2x folded into four 2x2 convolutions; k-block-major packing), the samplers on a
problem with a known answer, the job router on arbitrary job dictionaries, and the result packaging geometry."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F
from hypothesis import HealthCheck, assume, given, settings
from hypothesis import strategies as st
from PIL import Image

FAST = settings(max_examples=40, deadline=None, derandomize=True, database=None, suppress_health_check=[HealthCheck.too_slow])  # same examples every run


# ---------------------------------------------------------------- gang allocator
@FAST
@given(world=st.sampled_from([1, 2, 4, 8]),
       ops=st.lists(st.tuples(st.sampled_from(["lease", "release", "dead"]), st.integers(0, 7), st.sampled_from([1, 2, 4, 8])),
                    max_size=40))
def test_rank_group_invariants(world, ops):
    """Whatever the order of leases, releases and rank deaths: gangs are aligned power-of-two blocks, never overlap, never
    contain a dead rank, and the free count is what is neither leased nor dead."""
    from chiaswarm_b200.gpu.device_pool import RankGroup

    pool = RankGroup(world)
    held, dead = [], set()
    for op, k, size in ops:
        if op == "lease":
            if size > world:
                with pytest.raises(ValueError):
                    pool.lease(size, timeout=0)
                continue
            try:
                gang = pool.lease(size, timeout=0)
            except ValueError:
                raise
            except Exception as e:  # the reference's pool raises a bare Exception("busy")
                assert str(e) == "busy"
                # refusal is only legitimate when no aligned block of that size is wholly free and alive
                busy = {r for g in held for r in g.ranks} | dead
                assert all(any(r in busy for r in block) for block in pool.aligned_subsets(size))
                continue
            assert gang.size == size and gang.ranks[0] % size == 0
            assert list(gang.ranks) == list(range(gang.ranks[0], gang.ranks[0] + size))
            assert not (set(gang.ranks) & dead)
            for g in held:
                assert not (set(g.ranks) & set(gang.ranks))
            held.append(gang)
        elif op == "release" and held:
            pool.release(held.pop(k % len(held)))
        elif op == "dead":
            r = k % world
            lost = pool.mark_dead(r)
            dead.add(r)
            for g in lost:  # the leases that contain the rank are reported; they stay leased until released
                assert r in g.ranks and any(h.lease_id == g.lease_id for h in held)
        leased = {r for g in held for r in g.ranks}
        assert pool.free_count() == world - len(leased | dead)
    for g in held:
        pool.release(g)
    assert pool.free_count() == world - len(dead) and pool.dead_ranks() == sorted(dead)


@FAST
@given(world=st.sampled_from([1, 2, 4, 8]), mode=st.sampled_from(["dp", "patch", "tp", "auto", "bogus"]),
       gang_size=st.integers(0, 16), w=st.integers(1, 64), h=st.integers(1, 64), n_img=st.integers(1, 9),
       workflow=st.sampled_from([None, "txt2vid", "vid2vid", "txt2audio", "img2txt", "stitch"]))
def test_gang_size_policy_is_always_leasable(world, mode, gang_size, w, h, n_img, workflow):
    from chiaswarm_b200.parallel.gang_worker import plan_gang_size
    from chiaswarm_b200.settings import Settings

    s = Settings()
    s.parallel_mode, s.gang_size = mode, gang_size
    job = {"id": "p", "model_name": "tiny-sd", "width": w * 32, "height": h * 32, "num_images_per_prompt": n_img}
    if workflow:
        job["workflow"] = workflow
    n = plan_gang_size(job, s, world)
    assert n in (1, 2, 4, 8) and n <= world
    if mode == "dp" or world == 1 or workflow in ("txt2audio", "img2txt", "stitch"):
        assert n == 1


# ---------------------------------------------------------------- weight rewrites
@FAST
@given(cin=st.integers(1, 5), cout=st.integers(1, 5), h=st.integers(1, 6), w=st.integers(1, 6), seed=st.integers(0, 2 ** 16))
def test_folded_upsample_equals_upsample_then_conv(cin, cout, h, w, seed):
    """conv3x3(nearest2x(x)) == four 2x2 phase convolutions on x with pre-summed taps, interleaved (W6)."""
    from chiaswarm_b200.ops.packing import fold_upsample2x

    g = torch.Generator().manual_seed(seed)
    x = torch.randn(1, cin, h, w, generator=g, dtype=torch.float64)
    wgt = torch.randn(cout, cin, 3, 3, generator=g, dtype=torch.float64)
    want = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), wgt, padding=1)
    got = torch.empty_like(want)
    xp = F.pad(x, (1, 1, 1, 1))
    for (a, b), packed in fold_upsample2x(wgt).items():
        k = packed.reshape(cout, 2, 2, cin).permute(0, 3, 1, 2)      # [cout, (tap_i, tap_j, cin)] -> torch conv layout
        # phase (a, b) reads x[y - 1 + a + i, x - 1 + b + j]: a 2x2 window whose top-left starts at offset (a, b) of xp
        got[:, :, a::2, b::2] = F.conv2d(xp[:, :, a:a + h + 1, b:b + w + 1], k)
    assert torch.allclose(got, want, atol=1e-5)  # (taps are summed in fp32 whatever the weight dtype)


@FAST
@given(cout=st.integers(1, 9), taps=st.sampled_from([1, 4, 9]), c1=st.integers(1, 150), c2=st.integers(0, 70),
       seed=st.integers(0, 2 ** 16))
def test_kmajor_packing_round_trips(cout, taps, c1, c2, seed):
    from chiaswarm_b200.ops.packing import pack_kmajor, unpack_kmajor

    wgt = torch.randn(cout, taps * (c1 + c2), generator=torch.Generator().manual_seed(seed))
    packed = pack_kmajor(wgt, taps, c1, c2)
    assert packed.shape == (taps * ((c1 + 63) // 64 + (c2 + 63) // 64), cout, 64)
    assert torch.equal(unpack_kmajor(packed, taps, c1, c2), wgt)
    assert float(packed.abs().sum()) == pytest.approx(float(wgt.abs().sum()), rel=1e-5)  # padding is zeros


# ---------------------------------------------------------------- samplers
def _toy(plan, x0, noise):
    """A 'model' that knows the answer: with x_t = x0 + sigma * noise (sigma space) or its variance-preserving twin, the
    exact epsilon / v / sample prediction makes every sampler land on x0."""
    def predict(x, i):
        s = float(plan.sigmas[i])
        if plan.kind in ("euler", "euler_a"):
            eps = (x - x0) / s
        else:
            a = 1.0 / math.sqrt(s * s + 1.0)
            eps = (x - a * x0) / (s * a)
        if plan.prediction_type == "epsilon":
            return eps
        if plan.prediction_type == "sample":
            return x0.clone()
        a = 1.0 / math.sqrt(s * s + 1.0)
        return a * eps - (s * a) * x0        # v = alpha * eps - sigma_t * x0
    return predict


@FAST
@given(kind=st.sampled_from(["dpmpp_2m", "euler", "euler_a", "ddim"]), steps=st.integers(1, 40),
       pred=st.sampled_from(["epsilon", "v_prediction", "sample"]), karras=st.booleans(), seed=st.integers(0, 999))
def test_every_sampler_recovers_a_known_sample(kind, steps, pred, karras, seed):
    from chiaswarm_b200.ops import schedulers

    plan = schedulers.make_plan(kind, steps, {"prediction_type": pred}, karras)
    assert plan.num_steps == steps and len(plan.sigmas) == steps + 1 and plan.sigmas[-1] == 0.0
    assert np.all(np.diff(plan.sigmas) < 0) and np.all(np.diff(plan.timesteps) <= 0)
    assert np.isfinite(plan.coef).all()
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(1, 4, 4, 4, generator=g, dtype=torch.float64)
    noise = torch.randn(1, 4, 4, 4, generator=g, dtype=torch.float64)
    x = plan.add_noise(x0, noise, 0)
    predict = _toy(plan, x0, noise)
    x0_prev = None
    for i in range(steps):
        fresh = torch.randn(x.shape, generator=g, dtype=torch.float64) if plan.needs_noise else None
        x, x0_prev, _ = schedulers.apply_step_torch(plan, i, x, predict(x, i), x0_prev, fresh)
        x, x0_prev = x.double(), x0_prev.double()
        assert torch.allclose(x0_prev, x0, atol=2e-3), (i, float((x0_prev - x0).abs().max()))
    if kind != "ddim":  # (DDIM ends at alphas_cumprod[0], not at zero noise: set_alpha_to_one=False in the SD configs)
        assert torch.allclose(x, x0, atol=5e-3), float((x - x0).abs().max())


@FAST
@given(steps=st.integers(2, 40), start=st.integers(0, 39), kind=st.sampled_from(["dpmpp_2m", "euler", "euler_a", "ddim"]))
def test_img2img_slices_keep_the_tail_of_the_plan(steps, start, kind):
    from chiaswarm_b200.ops import schedulers

    assume(start < steps)
    plan = schedulers.make_plan(kind, steps, None, kind != "ddim")
    if kind == "ddim":
        return  # (DDIM img2img starts from add_noise at a table timestep; it has no slice)
    tail = plan.slice_from(start)
    assert tail.num_steps == steps - start
    assert np.array_equal(tail.sigmas, plan.sigmas[start:]) and np.array_equal(tail.timesteps, plan.timesteps[start:])
    assert np.array_equal(tail.coef[1:], plan.coef[start + 1:])   # only the first row may change (multistep restart)
    if kind in ("euler", "euler_a"):
        assert tail.first_scale == pytest.approx(1 / math.sqrt(plan.sigmas[start] ** 2 + 1))


# ---------------------------------------------------------------- job router and packaging
_JSON = st.recursive(st.none() | st.booleans() | st.integers(-5, 2000) | st.floats(-2, 2, allow_nan=False) | st.text(max_size=8),
                     lambda c: st.lists(c, max_size=3) | st.dictionaries(st.text(max_size=6), c, max_size=3), max_leaves=6)


_TYPE_NAMES = ["EulerDiscreteScheduler", "DDIMScheduler", "StableDiffusionPipeline", "StableDiffusionImg2ImgPipeline",
               "os.system", "__import__", "NoSuchScheduler"]


@FAST
@given(workflow=st.sampled_from([None, "txt2img", "img2img", "txt2audio", "txt2vid", "vid2vid", "stitch", "img2txt", "nonsense"]),
       model=st.sampled_from(["tiny-sd", "suno/bark", "cvssp/audioldm", "DeepFloyd/IF-I-XL-v1.0", "timbrooks/instruct-pix2pix"]),
       height=st.one_of(st.none(), st.integers(8, 2048)), width=st.one_of(st.none(), st.integers(8, 2048)),
       extra=st.dictionaries(st.sampled_from(["prompt", "negative_prompt", "guidance_scale", "num_inference_steps", "strength",
                                              "num_images_per_prompt", "seed", "content_type", "outputs"]), _JSON, max_size=4),
       params=st.fixed_dictionaries({}, optional={
           "upscale": st.booleans(), "supports_xformers": st.booleans(),
           "unsupported_pipeline_arguments": st.lists(st.sampled_from(["strength", "negative_prompt", "nothing"]), max_size=2),
           "scheduler_type": st.sampled_from(_TYPE_NAMES), "pipeline_type": st.sampled_from(_TYPE_NAMES)}))
def test_router_never_resolves_outside_its_tables(workflow, model, height, width, extra, params):
    """Arbitrary job dictionaries either route to one of the known callbacks with types from the closed registry, or are
    refused for a reason the contract names (size cap, unknown type name); the caller's job dict is never mutated."""
    import copy

    from chiaswarm_b200 import job_arguments
    from chiaswarm_b200.registry import pipeline_names, scheduler_names

    job = {"id": "r", "model_name": model, **extra}
    if workflow:
        job["workflow"] = workflow
    if height is not None:
        job["height"] = height
    if width is not None:
        job["width"] = width
    if params:
        job["parameters"] = dict(params)
    before = copy.deepcopy(job)
    known_types = set(pipeline_names()) | set(scheduler_names())
    resolves_types = workflow in ("txt2vid",) or (workflow == "txt2audio" and model != "suno/bark") or (
        workflow not in ("txt2audio", "stitch", "img2txt", "vid2vid", "txt2vid") and not model.startswith("DeepFloyd/"))
    sd_path = workflow not in ("txt2audio", "stitch", "img2txt", "vid2vid", "txt2vid") and not model.startswith("DeepFloyd/")
    too_big = sd_path and height is not None and width is not None and max(height, width) > 1024
    bad_type = resolves_types and any(params.get(k) is not None and params[k] not in known_types
                                      for k in ("scheduler_type", "pipeline_type"))
    try:
        cb, args = job_arguments.format_args(job)
    except Exception as e:
        assert too_big or bad_type, (type(e), e)
        assert "max image size" in str(e) or "Unknown pipeline or scheduler type" in str(e)
    else:
        assert not too_big and not bad_type
        assert cb in set(job_arguments._callbacks().values()) and "workflow" not in args
        for key in ("pipeline_type", "scheduler_type"):
            if key in args:
                assert type(args[key]).__module__ == "chiaswarm_b200.registry", args[key]
    job_now = {k: v for k, v in job.items() if k != "parameters"}
    assert job_now == {k: v for k, v in before.items() if k != "parameters"}  # (the top level is copied, never edited)


@FAST
@given(count=st.integers(1, 9), w=st.integers(8, 40), h=st.integers(8, 40))
def test_grid_geometry(count, w, h):
    from chiaswarm_b200 import output_processor as op

    images = [Image.new("RGB", (w, h), (i * 20, 0, 0)) for i in range(count)]
    sheet = op.post_process(images)
    if count == 1:
        assert sheet is images[0]
        return
    rows, cols = op._GRID_FOR_COUNT[count]
    assert rows * cols >= count and sheet.size == (cols * w, rows * h)
    for i in range(count):  # image i sits at (row, col) = divmod(i, cols)
        r, c = divmod(i, cols)
        assert sheet.getpixel((c * w + w // 2, r * h + h // 2))[0] == i * 20


def test_grid_refuses_more_than_nine():
    from chiaswarm_b200 import output_processor as op

    with pytest.raises(ValueError, match="Too many images"):
        op.post_process([Image.new("RGB", (8, 8))] * 10)
    with pytest.raises(ValueError):
        op.post_process([])
