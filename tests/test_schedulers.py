"""Scheduler coefficient tables against independent fp64 restatements of the formulas
(SURVEY.md appendix A.1-A.4)."""
import math

import numpy as np
import pytest
import torch

from chiaswarm_b200.ops import schedulers as S


def _run_table(plan, x, eps_fn, noise=None):
    x = torch.tensor(x, dtype=torch.float64)
    x0_prev = None
    for i in range(plan.num_steps):
        out = torch.tensor(eps_fn(x.numpy(), i), dtype=torch.float64)
        e0, e1, c0, c1, c2, c3, s_in, _ = (float(v) for v in plan.coef[i].astype(np.float64))
        x0 = e0 * x + e1 * out
        xn = c0 * x + c1 * x0
        if c2:
            xn = xn + c2 * x0_prev
        if c3:
            xn = xn + c3 * torch.tensor(noise[i])
        x, x0_prev = xn, x0
    return x.numpy()


@pytest.mark.parametrize("steps", [2, 10, 30, 50])
@pytest.mark.parametrize("pred", ["epsilon", "v_prediction"])
def test_dpmpp_2m_table_matches_formulas(steps, pred):
    plan = S.make_plan("dpmpp_2m", steps, {"prediction_type": pred}, use_karras_sigmas=True)
    assert plan.coef.shape == (steps, 8) and np.isfinite(plan.coef).all()
    assert plan.sigmas[-1] == 0 and np.all(np.diff(plan.sigmas) < 0)
    rng = np.random.default_rng(0)
    x = rng.standard_normal(16)
    w = rng.standard_normal(16)
    eps_fn = lambda xx, i: 0.3 * xx + 0.1 * w * math.cos(i)
    got = _run_table(plan, x, eps_fn)
    ref = S.reference_dpmpp_2m(plan, x, eps_fn)
    assert np.allclose(got, ref, rtol=2e-4, atol=2e-4)


def test_karras_sigmas_and_timesteps():
    plan = S.make_plan("dpmpp_2m", 50, None, True)
    acp = plan.alphas_cumprod
    smax, smin = math.sqrt((1 - acp[-1]) / acp[-1]), math.sqrt((1 - acp[0]) / acp[0])
    assert plan.sigmas[0] == pytest.approx(smax) and plan.sigmas[-2] == pytest.approx(smin)
    rho = 7.0
    i = 17
    expect = (smax ** (1 / rho) + i / 49 * (smin ** (1 / rho) - smax ** (1 / rho))) ** rho
    assert plan.sigmas[i] == pytest.approx(expect)
    assert plan.timesteps[0] == 999 and plan.timesteps[-1] <= 1 and np.all(np.diff(plan.timesteps) <= 0)


def test_euler_and_ancestral_rows():
    plan = S.make_plan("euler", 8, None, True)
    s, sn = plan.sigmas[3], plan.sigmas[4]
    x, eps = 0.7, -0.2
    x0 = x - s * eps
    d = (x - x0) / s
    e0, e1, c0, c1, *_ = plan.coef[3]
    got = c0 * x + c1 * (e0 * x + e1 * eps)
    assert got == pytest.approx(x + d * (sn - s), rel=1e-5)
    assert plan.init_noise_sigma == pytest.approx(plan.sigmas[0])
    assert plan.first_scale == pytest.approx(1 / math.sqrt(plan.sigmas[0] ** 2 + 1))
    assert plan.coef[3][6] == pytest.approx(1 / math.sqrt(sn * sn + 1), rel=1e-6)

    pa = S.make_plan("euler_a", 8, None, True)
    assert pa.needs_noise
    up = math.sqrt(sn ** 2 * (s ** 2 - sn ** 2) / s ** 2)
    down = math.sqrt(sn ** 2 - up ** 2)
    e0, e1, c0, c1, c2, c3, *_ = pa.coef[3]
    nz = 0.4
    got = c0 * x + c1 * (e0 * x + e1 * eps) + c3 * nz
    assert got == pytest.approx(x + d * (down - s) + up * nz, rel=1e-5)
    assert pa.coef[-1][5] == 0  # last step lands on sigma 0: no noise


def test_ddim_rows_and_guidance():
    plan = S.make_plan("ddim", 10, None, False)
    acp = plan.alphas_cumprod
    t = int(plan.timesteps[2])
    prev = t - 100
    x, eps = 0.3, 0.9
    x0 = (x - math.sqrt(1 - acp[t]) * eps) / math.sqrt(acp[t])
    expect = math.sqrt(acp[prev]) * x0 + math.sqrt(1 - acp[prev]) * eps
    e0, e1, c0, c1, *_ = plan.coef[2].astype(np.float64)
    assert c0 * x + c1 * (e0 * x + e1 * eps) == pytest.approx(expect, rel=1e-5)

    out = torch.tensor([[1.0], [3.0]])
    assert S.guided(out, 2, 7.5).item() == pytest.approx(1 + 7.5 * 2)
    out3 = torch.tensor([[5.0], [2.0], [1.0]])  # [text+img, img, uncond]
    assert S.guided(out3, 3, 7.5, 1.5).item() == pytest.approx(1 + 7.5 * (5 - 2) + 1.5 * (2 - 1))


def test_img2img_slice_restarts_first_order():
    plan = S.make_plan("dpmpp_2m", 20, None, True)
    sl = plan.slice_from(8)
    assert sl.num_steps == 12 and sl.coef[0][4] == 0 and plan.coef[8][4] != 0
    x0 = torch.ones(4)
    nz = torch.full((4,), 2.0)
    s = plan.sigmas[8]
    a = 1 / math.sqrt(s * s + 1)
    assert torch.allclose(sl.add_noise(x0, nz, 0), a * x0 + s * a * nz)


def test_ddim_eta_adds_the_ddpm_variance():
    """eta > 0: x_prev = sqrt(a_p) x0 + sqrt(1 - a_p - s^2) eps + s z with s = eta sqrt((1-a_p)/(1-a_t)) sqrt(1 - a_t/a_p)."""
    import math

    det = S.make_plan("ddim", 10, None, False)
    sto = S.make_plan("ddim", 10, {"eta": 1.0}, False)
    assert not det.needs_noise and sto.needs_noise and np.all(det.coef[:, 5] == 0)
    acp = det.alphas_cumprod
    ratio = 1000 // 10
    for i, t in enumerate(det.timesteps.astype(int)):
        a_t, a_p = acp[t], (acp[t - ratio] if t - ratio >= 0 else acp[0])
        s = math.sqrt((1 - a_p) / (1 - a_t) * (1 - a_t / a_p))
        assert sto.coef[i, 5] == pytest.approx(s, rel=1e-5, abs=1e-7)
        # the deterministic and the noise parts together keep the marginal variance 1 - a_p
        k = sto.coef[i, 2] * math.sqrt(1 - a_t)
        assert k * k + s * s == pytest.approx(1 - a_p, rel=1e-4)
    # with a perfect epsilon model the x0 estimate stays exact whatever the noise
    g = torch.Generator().manual_seed(0)
    x0 = torch.randn(1, 4, 4, 4, generator=g, dtype=torch.float64)
    eps = torch.randn(1, 4, 4, 4, generator=g, dtype=torch.float64)
    x = sto.add_noise(x0, eps, 0)
    for i in range(sto.num_steps):
        s_i = float(sto.sigmas[i])
        a = 1 / math.sqrt(s_i * s_i + 1)
        pred = (x - a * x0) / (s_i * a)
        x, x0_hat, _ = S.apply_step_torch(sto, i, x, pred, None, torch.randn(x.shape, generator=g, dtype=torch.float64))
        x = x.double()
        assert torch.allclose(x0_hat.double(), x0, atol=1e-3)


def test_euler_initial_noise_scale_follows_the_timestep_spacing():
    """diffusers: init_noise_sigma = sigma_max for linspace / trailing spacing, sqrt(sigma_max^2 + 1) for leading (SDXL)."""
    plain = S.make_plan("euler", 20, None, True)
    lead = S.make_plan("euler", 20, {"timestep_spacing": "leading"}, True)
    assert plain.init_noise_sigma == pytest.approx(plain.sigmas[0])
    assert lead.init_noise_sigma == pytest.approx(math.sqrt(lead.sigmas[0] ** 2 + 1))
    assert np.array_equal(plain.sigmas, lead.sigmas) and np.array_equal(plain.coef, lead.coef)
    assert S.make_plan("dpmpp_2m", 20, {"timestep_spacing": "leading"}, True).init_noise_sigma == 1.0


def _ode_error(kind, steps, karras=True, s_data=0.6):
    """Final error of a sampler on a problem whose probability-flow ODE has a closed form: data ~ N(0, s^2), so the exact
    noise prediction is eps(x, sigma) = sigma x / (s^2 + sigma^2) (x in sigma space) and the exact solution is
    x(sigma) = x(sigma_max) sqrt((s^2 + sigma^2) / (s^2 + sigma_max^2))."""
    plan = S.make_plan(kind, steps, None, karras)
    g = torch.Generator().manual_seed(0)
    n = torch.randn(4096, generator=g, dtype=torch.float64)
    s_max = float(plan.sigmas[0])
    sigma_space = kind in ("euler", "euler_a")
    x = n * s_max if sigma_space else n * s_max / math.sqrt(s_max ** 2 + 1)       # the same start, in the plan's own space
    x0_prev = None
    for i in range(steps):
        s = float(plan.sigmas[i])
        xs = x if sigma_space else x * math.sqrt(s * s + 1)
        eps = s * xs / (s_data ** 2 + s * s)
        x, x0_prev, _ = S.apply_step_torch(plan, i, x, eps, x0_prev, None)
        x, x0_prev = x.double(), x0_prev.double()
    exact = n * s_max * s_data / math.sqrt(s_data ** 2 + s_max ** 2)
    return float((x - exact).abs().max() / exact.abs().max())


def test_samplers_converge_to_the_exact_ode_solution_at_their_order():
    """Not a restatement of the update formulas: the tables are run against the closed-form solution of the ODE they
    discretise.  DPM-Solver++(2M) converges at second order, Euler at first; all errors are small at 40 steps."""
    e2m = [_ode_error("dpmpp_2m", n) for n in (10, 20, 40, 80)]
    eul = [_ode_error("euler", n) for n in (10, 20, 40, 80)]
    assert e2m[2] < 3e-3 and eul[2] < 5e-2 and e2m[2] < eul[2] / 10
    for coarse, fine in zip(e2m, e2m[1:]):
        assert coarse / fine > 3.0, e2m                     # ~4x per halving of the step: second order
    for coarse, fine in zip(eul, eul[1:]):
        assert 1.6 < coarse / fine < 2.6, eul               # ~2x: first order
    # DDIM (uniform timesteps, ends at alphas_cumprod[0] rather than zero noise): first-order decrease as well
    dd = [_ode_error("ddim", n, karras=False) for n in (25, 50, 100)]
    assert dd[0] > dd[1] > dd[2] and dd[2] < 5e-2, dd
