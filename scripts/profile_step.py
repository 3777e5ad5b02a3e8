"""Runs a few denoising steps + one VAE decode without CUDA graphs so that ncu sees every kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ["SDAAS_CUDA_GRAPH"] = "0"
import torch

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

arch = sys.argv[1] if len(sys.argv) > 1 else "sd15"
size = int(sys.argv[2]) if len(sys.argv) > 2 else 512
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
pipe = StableDiffusionPipeline.from_architecture(arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
g = torch.Generator(device="cuda:0").manual_seed(0)
out = pipe(prompt="profile", height=size, width=size, num_inference_steps=steps, generator=g, output_type="device")
torch.cuda.synchronize()
print("done", out.device_images.shape)
