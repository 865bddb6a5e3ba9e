"""Per-block relative error of the CUDA path against the fp32 oracle (diagnostic, not a test).
    python tests/gpu_precision.py [H ch B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.xunet import XUNet  # noqa: E402
from oracle import diffusion_ref as dr, xunet_ref as xr  # noqa: E402


def main():
    H, ch, B = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (16, 32, 2)))
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
    sdo = {k: v.to(dev) for k, v in sd.items()}
    taps = {}
    with torch.no_grad():
        eo = xr.xunet_forward(sdo, ocfg, {k: v.to(dev) for k, v in batch.items()}, cm.to(dev), taps=taps)
    m = XUNet(**kw).to(dev)
    m.load_state_dict(sd)
    m.eval()
    with torch.no_grad():
        e = m({k: v.to(dev) for k, v in batch.items()}, cond_mask=cm.to(dev))
    eng = m._engines[B]
    rel = lambda a, b: ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()
    pe = eng._bufs["pose_emb"].view(B, 2, H, H, 144).permute(0, 1, 4, 2, 3)
    print(f"pose_emb           {rel(pe, taps['pose_emb']):.3e}")
    print(f"logsnr_emb         {rel(eng._bufs['lemb'].view(B, 2, -1), taps['logsnr_emb']):.3e}")
    for name, d in eng.saved["acts"].items():
        if name in taps:
            t = d["final"].t if "final" in d else d["out"].t
            N, hh, ww, C = t.shape
            ours = t.view(B, 2, hh, ww, C).permute(0, 1, 4, 2, 3)
            print(f"{name:45s} {rel(ours, taps[name]):.3e}")
    print(f"eps                {rel(e, eo):.3e}")


if __name__ == "__main__":
    main()
