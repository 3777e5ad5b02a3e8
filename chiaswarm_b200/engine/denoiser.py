"""SM100Denoiser: the production denoising loop on B200.

Per (batch, latent size, n_cond, mode) it records ONE static plan of sm_100a kernel launches --
[ControlNet encoder] -> UNet -> [ControlNet residual adds] -> fused CFG + scheduler update ->
step-counter increment -- captures it into a CUDA graph, and replays that graph once per
denoising step.  All per-step variation (timestep, scheduler coefficients, ancestral noise) lives
in device tables indexed by a device-side step counter, so the host does nothing between steps.
The VAE decode is a second captured plan ending in a uint8 HWC image.

Replaces diffusers' Python denoising loop and its ~1-2 k eager launches per step
(SURVEY.md section 3.2 "HOT LOOP"; reference call site swarm/diffusion/diffusion_func.py:96).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import torch

from ..ops import native as nat
from .layers import FM, Recorder, _r8
from .unet_sm100 import ControlNetPlan, UNetPlan, VaeDecoderPlan

MAX_STEPS = 1024
NOISE_STEPS_MIN = 64  # rows of a captured per-step noise table (ancestral samplers, DDIM eta): a graph built for a 15-step job
#                       also serves a 50-step job of the same shape; longer jobs re-record (single GPU) or are refused (gangs)


def _use_graphs() -> bool:
    return os.environ.get("SDAAS_CUDA_GRAPH", "1") not in ("0", "false", "")


def capture_mode() -> str:
    """cudaStreamCaptureMode of the plan captures.  "global" (the default, one lane per process: launcher, gang ranks,
    bench) makes an allocation on ANY thread during a capture an error; a process that serves several GPUs from lane
    threads (`python -m swarm.worker`, worker.py sets SW_GRAPH_CAPTURE_MODE) needs "thread_local", or lane A's capture
    fails whenever lane B allocates."""
    mode = os.environ.get("SW_GRAPH_CAPTURE_MODE", "global")
    return mode if mode in ("global", "thread_local", "relaxed") else "global"


class _StepGraph:
    """Static buffers + recorded plan (+ CUDA graph) for one problem shape."""

    def __init__(self, eng: "SM100Denoiser", batch, h, w, n_cond, extra_ch, has_control, has_noise, aug: bool):
        dev = eng.device
        up = eng.unet_plan
        self.batch, self.h, self.w, self.n_cond = batch, h, w, n_cond
        n = batch * n_cond
        rows = batch * h * w
        self.rows = rows
        self.lat_c = eng.lat_c
        self.x = torch.zeros(rows, self.lat_c, dtype=torch.float32, device=dev)
        self.x0_prev = torch.zeros(rows, self.lat_c, dtype=torch.float32, device=dev)
        self.model_in = torch.zeros(n, h, w, up.in_pad, dtype=torch.float16, device=dev)
        self.t_table = torch.zeros(MAX_STEPS, dtype=torch.float32, device=dev)
        self.coef = torch.zeros(MAX_STEPS, 8, dtype=torch.float32, device=dev)
        self.step = torch.zeros(1, dtype=torch.int32, device=dev)
        self.noise = torch.zeros(0, device=dev)
        self.has_noise = has_noise
        self.guidance = torch.zeros(2, dtype=torch.float32, device=dev)  # (g_text, g_image), filled per job
        self.aug_emb = torch.zeros(n, up.boc[0] * 4, dtype=torch.float16, device=dev) if aug else None
        self.control_feat: Optional[FM] = None
        self.rec: Optional[Recorder] = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.has_control = has_control
        self.eng = eng

    def build(self, max_noise_steps: int):
        eng, up = self.eng, self.eng.unet_plan
        n = self.batch * self.n_cond
        if self.has_noise:
            self.noise = torch.zeros(max_noise_steps, self.rows, self.lat_c, dtype=torch.float32, device=eng.device)
        rec = Recorder(eng.device)
        x_in = FM(self.model_in, n, self.h, self.w, up.in_pad)
        temb = up.record_time_embedding(rec, n, self.t_table, self.step, self.aug_emb)
        cn_feats = cn_mid = None
        if self.has_control:
            cp = eng.control_plan
            self.control_feat = rec.fm(n, self.h, self.w, cp.boc[0])  # filled per job by the cond embedder
            ctemb = cp.record_time_embedding(rec, n, self.t_table, self.step, None)
            cn_mid, cn_feats = cp.record_encoder(rec, x_in, ctemb, cond_feat=self.control_feat)
        mid, skips = up.record_encoder(rec, x_in, temb)
        if self.has_control:
            eng.control_plan.record_residual_adds(rec, cn_feats, cn_mid, skips, mid)
        eps = up.record_decoder(rec, mid, skips, temb)
        rec.add(self._sched, eps)
        rec.add(nat.advance_counter, self.step)
        self.rec = rec
        return self

    def _sched(self, eps: FM):
        nat.sched_step(eps=eps.t, n_cond=self.n_cond, rows=self.rows, c=self.lat_c, eps_ld=eps.c,
                       g_text=0.0, g_image=0.0, guidance=self.guidance, x=self.x, x0_prev=self.x0_prev,
                       noise=self.noise if self.has_noise else None, x_model_in=self.model_in,
                       xin_ld=self.model_in.shape[-1], coef=self.coef, step_idx=self.step)

    def capture(self):
        # one eager pass first: sets kernel attributes, loads modules, validates every launch
        self.rec.run()
        torch.cuda.synchronize()
        if _use_graphs():
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode=capture_mode()):
                self.rec.run()
            self.graph = g
        self.launches_per_step = len(self.rec)

    def replay(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.rec.run()


class SM100Denoiser:
    def __init__(self, pipe):
        self.pipe = pipe
        self.device = pipe.device
        nat.lib()  # a missing native library on a GPU box is an error, never a silent fallback
        self.lat_c = pipe.vae.config["latent_channels"]
        if hasattr(pipe.unet, "transformer_in"):  # UNet3DConditionModel: frames ride in the batch dimension
            from .unet3d_sm100 import UNet3DPlan

            self.unet_plan = UNet3DPlan(pipe.unet, self.device)
        else:
            self.unet_plan = UNetPlan(pipe.unet, self.device)
        self.control_plan = ControlNetPlan(pipe.controlnet, self.device) if pipe.controlnet is not None else None
        self.vae_plan = None
        self._graphs: Dict[tuple, _StepGraph] = {}  # insertion order = LRU order
        self.max_graphs = 8
        self._vae_graphs: Dict[tuple, tuple] = {}
        self.last_launches = 0

    # ------------------------------------------------------------------ denoise
    @torch.no_grad()
    def run(self, x, embeds, plan, n_cond, g_text, g_image, extra, control, control_scale, added, noise, inpaint,
            num_frames: int = 1):
        """`num_frames` > 1 (UNet3D): `x` is [clips*frames, c, h, w] and `embeds` holds one row per frame."""
        if inpaint is not None:
            # legacy 4-channel inpainting re-imposes the known region between steps (host-side blend)
            from ..diffusion.pipelines import TorchDenoiser

            return TorchDenoiser(self.pipe).run(x, embeds, plan, n_cond, g_text, g_image, extra, control,
                                                control_scale, added, noise, inpaint)
        batch, c, h, w = x.shape
        steps = plan.num_steps
        if steps > MAX_STEPS:
            raise ValueError(f"at most {MAX_STEPS} denoising steps are supported")
        extra_ch = 0 if extra is None else extra.shape[1]
        # shapes only: guidance lives in a device slot and the ControlNet scale in the zero-conv weights, both
        # refreshed per job, so a new guidance value never re-records or re-captures a plan
        key = (batch, h, w, n_cond, extra_ch, control is not None, noise is not None, added is not None, num_frames)
        sg = self._graphs.pop(key, None)
        if sg is not None and noise is not None and sg.noise.shape[0] < steps:
            sg.graph = sg.rec = None  # the captured noise table is too short for this job: re-record with a longer one
            sg = None
        if sg is not None:
            self._graphs[key] = sg  # most recently used goes last
        up = self.unet_plan
        if num_frames > 1 or hasattr(up, "num_frames"):
            if not hasattr(up, "num_frames") or batch % num_frames:
                raise ValueError("num_frames needs a UNet3D plan and a batch that is a multiple of it")
            up.num_frames = num_frames
        up.set_context(embeds.to(torch.float16))
        if self.control_plan is not None and control is not None:
            self.control_plan.set_context(embeds.to(torch.float16))
        if sg is None:
            sg = _StepGraph(self, batch, h, w, n_cond, extra_ch, control is not None, noise is not None,
                            added is not None)
            sg.build(max(steps, NOISE_STEPS_MIN) if noise is not None else 0)
            self._load_job(sg, x, plan, extra, control, added, noise, g_text, g_image, control_scale)
            sg.capture()
            self._graphs[key] = sg
            if len(self._graphs) > self.max_graphs:
                self._evict(next(iter(self._graphs)))  # least recently used
        self._load_job(sg, x, plan, extra, control, added, noise, g_text, g_image, control_scale)
        for _ in range(steps):
            sg.replay()
        self.last_launches = sg.launches_per_step * steps
        return sg.x.view(batch, h, w, c).permute(0, 3, 1, 2).contiguous()

    def _evict(self, key):
        sg = self._graphs.pop(key, None)
        if sg is not None:
            sg.graph = None
            sg.rec = None

    def _load_job(self, sg: _StepGraph, x, plan, extra, control, added, noise, g_text=7.5, g_image=1.5,
                  control_scale=1.0):
        batch, c, h, w = x.shape
        steps = plan.num_steps
        sg.guidance.copy_(torch.tensor([float(g_text), float(g_image)], dtype=torch.float32))
        if self.control_plan is not None and control is not None:
            self.control_plan.set_scale(float(control_scale))
        x_nhwc = x.float().permute(0, 2, 3, 1).reshape(sg.rows, c)
        sg.x.copy_(x_nhwc)
        sg.x0_prev.zero_()
        sg.step.zero_()
        sg.t_table[:steps].copy_(torch.from_numpy(plan.timesteps.astype(np.float32)))
        sg.coef[:steps].copy_(torch.from_numpy(plan.coef))
        if noise is not None:
            if sg.noise.shape[0] < steps:
                raise ValueError(f"this plan was captured with a {sg.noise.shape[0]}-step noise table; {steps} steps of a "
                                 "noise-driven sampler do not fit it")
            sg.noise[:steps].copy_(noise.float().permute(0, 1, 3, 4, 2).reshape(steps, sg.rows, c))
        first = (x_nhwc * plan.first_scale).to(torch.float16).view(batch, h, w, c)
        mi = sg.model_in.view(sg.n_cond, batch, h, w, -1)
        mi[..., :c] = first
        if extra is not None:
            e = extra.to(torch.float16).permute(0, 2, 3, 1).reshape(sg.n_cond, batch, h, w, -1)
            mi[..., c:c + e.shape[-1]] = e
        if added is not None:
            aug = self._aug_embedding(added)
            sg.aug_emb.copy_(aug)
        if control is not None:
            self._embed_control(sg, control)

    def _aug_embedding(self, added):
        """SDXL text_time embedding: add_embedding(concat(pooled text, sinusoidal time ids)); step
        invariant, so evaluated once per job with the module itself."""
        from ..models.unet2d import timestep_embedding

        unet = self.pipe.unet
        cfg = unet.config
        if "class_labels" in added:  # x4 upscaler: the noise level of the low-resolution input, as a class embedding
            return unet.class_embedding(added["class_labels"]).to(torch.float16)
        tids = added["time_ids"]
        tid_emb = timestep_embedding(tids.flatten(), cfg["addition_time_embed_dim"]).reshape(tids.shape[0], -1)
        vec = torch.cat([added["text_embeds"].float(), tid_emb], dim=-1).to(self.pipe.dtype)
        return unet.add_embedding(vec).to(torch.float16)

    def _embed_control(self, sg: _StepGraph, control):
        """ControlNet conditioning embedder (step invariant): one recorded plan per input shape, kept with the engine
        and re-run per job -- not re-recorded, and no scratch of its own."""
        cp = self.control_plan
        n, _, H, W = control.shape
        plans = self.__dict__.setdefault("_cond_plans", {})
        entry = plans.get((n, H, W))
        if entry is None:
            cond = torch.zeros(n, H, W, cp.cond_in_pad, dtype=torch.float16, device=self.device)
            rec = Recorder(self.device)
            feat = cp.record_cond_embedding(rec, FM(cond, n, H, W, cp.cond_in_pad))
            entry = plans[(n, H, W)] = (cond, rec, feat)
            if len(plans) > 4:
                plans.pop(next(iter(plans)))
        cond, rec, feat = entry
        nat.nchw_to_nhwc(control.contiguous(), cond, cp.cond_in_pad)
        rec.run()
        sg.control_feat.t.copy_(feat.t)

    # ------------------------------------------------------------------ VAE encode
    @torch.no_grad()
    def encode_moments(self, x):
        """x: images in [-1, 1], NCHW. Returns (mean, logvar) NCHW like AutoencoderKL.encode_moments, computed by the
        recorded sm_100a plan (one plan per input shape, kept with the engine)."""
        from .unet_sm100 import VaeEncoderPlan

        if getattr(self, "vae_enc_plan", None) is None:
            self.vae_enc_plan = VaeEncoderPlan(self.pipe.vae, self.device)
            self._vae_enc = {}
        vp = self.vae_enc_plan
        n, c, h, w = x.shape
        entry = self._vae_enc.get((n, h, w))
        if entry is None:
            xin = torch.zeros(n, h, w, vp.in_pad, dtype=torch.float16, device=self.device)
            rec = Recorder(self.device)
            out = vp.record(rec, FM(xin, n, h, w, vp.in_pad))
            entry = self._vae_enc[(n, h, w)] = (xin, out, rec)
            if len(self._vae_enc) > 4:
                self._vae_enc.pop(next(iter(self._vae_enc)))
        xin, out, rec = entry
        nat.nchw_to_nhwc(x.to(torch.float16).contiguous(), xin, vp.in_pad)
        rec.run()
        m = out.t[..., :vp.moments].permute(0, 3, 1, 2).to(x.dtype)
        return m.chunk(2, dim=1)

    # ------------------------------------------------------------------ VAE decode
    @torch.no_grad()
    def decode(self, z):
        """z: latents already divided by the scaling factor, NCHW. Returns uint8 [n, H, W, 3] -- or, for a VAE whose config
        asks for fp32 (`force_upcast`: the original SDXL VAE overflows fp16), float NCHW from the eager fp32 module."""
        from ..diffusion.pipelines import vae_fp32, vae_needs_fp32

        if vae_needs_fp32(self.pipe):
            return vae_fp32(self.pipe).decode(z.float())
        if self.vae_plan is None:
            self.vae_plan = VaeDecoderPlan(self.pipe.vae, self.device)
        n, c, h, w = z.shape
        key = (n, h, w)
        entry = self._vae_graphs.get(key)
        if entry is None:
            vp = self.vae_plan
            zin = torch.zeros(n, h, w, vp.lat_pad, dtype=torch.float16, device=self.device)
            rec = Recorder(self.device)
            img = vp.record(rec, FM(zin, n, h, w, vp.lat_pad))
            out = torch.empty(n, img.h, img.w, 3, dtype=torch.uint8, device=self.device)
            rec.add(nat.image_to_u8, img.t, out, n * img.h * img.w, img.c)
            nat.nchw_to_nhwc(z.contiguous(), zin, vp.lat_pad)
            rec.run()
            torch.cuda.synchronize()
            graph = None
            if _use_graphs():
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, capture_error_mode=capture_mode()):
                    rec.run()
            entry = (zin, out, rec, graph)
            self._vae_graphs[key] = entry
            if len(self._vae_graphs) > 4:
                self._vae_graphs.pop(next(iter(self._vae_graphs)))
        zin, out, rec, graph = entry
        nat.nchw_to_nhwc(z.contiguous(), zin, self.vae_plan.lat_pad)
        if graph is not None:
            graph.replay()
        else:
            rec.run()
        self.last_vae_launches = len(rec)
        return out
