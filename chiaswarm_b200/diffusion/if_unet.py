"""DeepFloyd-IF pixel-space stages (reference use: swarm/diffusion/diffusion_func_if.py:16-27).

IF's UNets are diffusers `UNet2DConditionModel`s built from `ResnetDownsampleBlock2D` /
`SimpleCrossAttnDownBlock2D` blocks with `resnet_time_scale_shift="scale_shift"`, added-KV
attention and a learned-variance head (6 output channels).  This loader maps those checkpoints
onto this framework's blocks where the block types coincide and refuses -- with a ValueError that
the job envelope reports as fatal -- configurations it does not implement, instead of producing
wrong pixels."""
from __future__ import annotations

import torch

from ..models.unet2d import UNet2DConditionModel
from ..ops import schedulers
from ..utils import checkpoints

_SUPPORTED_DOWN = {"DownBlock2D", "CrossAttnDownBlock2D"}
_SUPPORTED_UP = {"UpBlock2D", "CrossAttnUpBlock2D"}


class IFStage:
    def __init__(self, unet, sched_cfg, device):
        self.unet, self.sched_cfg, self.device = unet, sched_cfg, torch.device(device)

    @classmethod
    def from_checkpoint(cls, root, device):
        cfg = checkpoints.component_config(root, "unet")
        bad = [t for t in cfg.get("down_block_types", []) if t not in _SUPPORTED_DOWN]
        bad += [t for t in cfg.get("up_block_types", []) if t not in _SUPPORTED_UP]
        if bad or cfg.get("resnet_time_scale_shift", "default") != "default":
            raise ValueError(
                "DeepFloyd IF checkpoint uses UNet blocks this worker does not implement yet "
                f"({sorted(set(bad))}, resnet_time_scale_shift={cfg.get('resnet_time_scale_shift')})")
        unet = UNet2DConditionModel(**{k: v for k, v in cfg.items() if not k.startswith("_")})
        checkpoints.load_into(unet, checkpoints.load_state_dict(checkpoints.find_weights(root / "unet")), strict=False)
        sched = checkpoints.component_config(root, "scheduler") or {"beta_schedule": "squaredcos_cap_v2"}
        return cls(unet.to(device, torch.float16).eval(), sched, device)

    @torch.no_grad()
    def sample(self, cond, unc, size, steps, guidance_scale, generator=None, low_res=None):
        plan = schedulers.make_plan("ddim", int(steps), {**schedulers.SD_SCHEDULER_CONFIG, **self.sched_cfg}, False)
        c_in = self.unet.config["in_channels"]
        x = torch.randn((1, 3, size, size), generator=generator, device=self.device, dtype=torch.float16).float()
        embeds = torch.cat([unc, cond]).to(torch.float16)
        x0_prev = None
        for i in range(plan.num_steps):
            inp = x
            if low_res is not None and c_in == 6:
                up = torch.nn.functional.interpolate(low_res.float(), size=(size, size), mode="bilinear")
                inp = torch.cat([x, up], dim=1)
            t = torch.tensor(float(plan.timesteps[i]), device=self.device)
            out = self.unet(torch.cat([inp] * 2).to(torch.float16), t, embeds)[:, :3]
            out = schedulers.guided(out.float(), 2, float(guidance_scale))
            x, x0_prev, _ = schedulers.apply_step_torch(plan, i, x, out, x0_prev, None)
        return x.clamp(-1, 1)
