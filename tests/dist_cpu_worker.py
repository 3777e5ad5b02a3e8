"""Worker for tests/test_patch_parallel_cpu.py: runs under torch.multiprocessing with gloo."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main(rank, world, port, out_path):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from chiaswarm_b200.diffusion.pipelines import _init_module
    from chiaswarm_b200.models import configs
    from chiaswarm_b200.models.unet2d import UNet2DConditionModel
    from chiaswarm_b200.parallel import runtime
    from chiaswarm_b200.models.vae import AutoencoderKL
    from chiaswarm_b200.parallel.patch import PatchOps, PatchParallelUNet, PatchParallelVAEDecoder

    gang = runtime.init("gloo")
    torch.manual_seed(0)
    unet = _init_module(UNet2DConditionModel(**configs.get_architecture("tiny-sd")["unet"]), 3).eval()
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 4, 16, 8, generator=g)
    ctx = torch.randn(2, 7, 32, generator=g)
    with torch.no_grad():
        ref = unet(x, torch.tensor(500.0), ctx)
    pp = PatchParallelUNet(unet, gang)
    got = pp(x, torch.tensor(500.0), ctx)
    err = (got - ref).abs().max().item() / ref.abs().max().item()
    # primitive checks
    ops = PatchOps(gang)
    conv = torch.nn.Conv2d(4, 6, 3, stride=2, padding=1)
    with torch.no_grad():
        full = conv(x)
        mine = ops.gather_rows(ops.conv(conv, pp.split(x)))
    err2 = (full - mine).abs().max().item()
    # the image decoder, row-split: halos through the up-sampling convs, distributed GroupNorm, mid-block attention
    vae = _init_module(AutoencoderKL(**configs.get_architecture("tiny-sd")["vae"]), 5).eval()
    z = torch.randn(2, 4, 8, 6, generator=g)
    with torch.no_grad():
        img = vae.decode(z)
        img_pp = PatchParallelVAEDecoder(vae, gang)(z)
    err3 = (img - img_pp).abs().max().item() / img.abs().max().item()
    # same seed -> same initial latents on every rank
    lat = torch.randn(1, 4, 8, 8, generator=torch.Generator().manual_seed(77))
    parts = [torch.empty_like(lat) for _ in range(world)]
    dist.all_gather(parts, lat)
    same = all(torch.equal(p, lat) for p in parts)
    if rank == 0:
        with open(out_path, "w") as fh:
            fh.write(f"{err} {err2} {int(same)} {err3}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
