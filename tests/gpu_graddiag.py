"""Per-parameter gradient agreement (cosine, norm ratio) of the CUDA path vs the fp32 oracle (diagnostic).
    python tests/gpu_graddiag.py H ch B"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.diff3d import diffusion_loss  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.xunet import XUNet  # noqa: E402
from oracle import diffusion_ref as dr, xunet_ref as xr  # noqa: E402


def main():
    H, ch, B = (int(a) for a in sys.argv[1:4])
    kw = dict(H=H, W=H, ch=ch)
    dev = "cuda:0"
    ocfg = xr.Cfg(**kw)
    sd = xr.init_state_dict(ocfg, seed=11)
    img, R, T, K = synthetic.make_views(B, H, seed=12)
    g = torch.Generator().manual_seed(13)
    logsnr = dr.logsnr_schedule_cosine(torch.rand(B, generator=g))
    noise = torch.randn(img[:, 0].shape, generator=g)
    cm = torch.tensor([i % 2 == 0 for i in range(B)])
    batch = dr.make_batch(img[:, 0], logsnr, dr.q_sample(img[:, 1], logsnr, noise), R, T, K)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    sdo = {k: v.to(dev).requires_grad_(True) for k, v in sd.items()}
    eo = xr.xunet_forward(sdo, ocfg, {k: v.to(dev) for k, v in batch.items()}, cm.to(dev))
    dr.loss_fn(noise.to(dev), eo).backward()
    m = XUNet(**kw).to(dev)
    m.load_state_dict(sd)
    m.eval()
    e = m({k: v.to(dev) for k, v in batch.items()}, cond_mask=cm.to(dev))
    diffusion_loss(noise.to(dev), e).backward()
    tot = sum(sdo[n].grad.norm().item() ** 2 for n in sdo) ** 0.5
    for n, p in m.named_parameters():
        a, b = p.grad.flatten().float(), sdo[n].grad.flatten().float()
        cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
        print(f"{n:70s} cos {cos:8.5f}  |g| {a.norm().item():.3e} ref {b.norm().item():.3e}  share {b.norm().item() / tot:.2e}")


if __name__ == "__main__":
    main()
