"""Single-GPU engine-vs-eager check at the LOCAL shapes a 4-way split produces (SDXL strip = 256x1024 image,
zeroscope share = 6 frames): separates kernel bugs at those shapes from bugs in the cross-GPU exchanges."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
import torch
from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

dev = "cuda:0"
what = sys.argv[1] if len(sys.argv) > 1 else "both"
if what in ("sdxl", "both"):
    pipe = StableDiffusionPipeline.from_architecture("sdxl").to(dev, torch.float16)
    pipe.set_scheduler("dpmpp_2m", True)
    outs = []
    for backend in ("torch", "sm100"):
        pipe.backend = backend
        pipe._engine = None
        g = torch.Generator(device=dev).manual_seed(3)
        outs.append(pipe(prompt="strip", height=256, width=1024, num_inference_steps=2, guidance_scale=1.0, generator=g,
                         output_type="latent").latents.float())
    err = ((outs[0] - outs[1]).abs().max() / outs[0].abs().max()).item()
    print(f"sdxl 256x1024 engine vs eager rel_err {err:.3e}", flush=True)
    del pipe
    torch.cuda.empty_cache()
if what in ("t2v", "both"):
    from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline
    pipe = TextToVideoPipeline.from_architecture("zeroscope").to(dev, torch.float16)
    pipe.set_scheduler("dpmpp_2m", True)
    outs = []
    for backend in ("torch", "sm100"):
        pipe.backend = backend
        g = torch.Generator(device=dev).manual_seed(3)
        outs.append(pipe(prompt="six frames", height=320, width=576, num_frames=6, num_inference_steps=2, guidance_scale=1.0,
                         generator=g, output_type="latent").latents.float())
    err = ((outs[0] - outs[1]).abs().max() / outs[0].abs().max()).item()
    print(f"zeroscope 6 frames engine vs eager rel_err {err:.3e} finite={bool(torch.isfinite(outs[1]).all())}", flush=True)
