"""One launch of each non-GEMM kernel at its benchmark shape (batch 128 of CFG-64), as the target of
    ncu --set full --import-source on -k regex:'attn_|gn_|pose_embedding|loss|adam|q_sample' -o gpurun_out/r02_kernels python tests/ncu_targets.py
(the summaries under profiles/ come from that report).  Prints nothing of interest itself."""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402


def main():
    ops, dev = CudaOps("cuda:0"), "cuda:0"
    bf = torch.bfloat16
    # attention: 16x16 level (L=256, d=64) and 8x8 level (L=64, d=128), 256 images
    for N, L, heads, d in ((256, 256, 4, 64), (256, 64, 4, 128)):
        C = heads * d
        qkv = torch.randn(N, L, 3 * C, device=dev).to(bf)
        out = torch.empty(N, L, C, device=dev, dtype=bf)
        probs = torch.empty(N, heads, L, L, device=dev, dtype=bf)
        dout = torch.randn(N, L, C, device=dev).to(bf)
        dqkv = torch.empty(N, L, 3 * C, device=dev, dtype=bf)
        ops.attn_fwd(qkv, heads, True, 1.0 / math.sqrt(d), out, probs=probs)
        ops.attn_bwd(qkv, probs, out, dout, dqkv, heads, True, 1.0 / math.sqrt(d))
        lse = torch.empty(N, heads, L, device=dev)                     # flash-style: row statistics out, P recomputed in the backward
        ops.attn_fwd(qkv, heads, True, 1.0 / math.sqrt(d), out, lse=lse)
        ops.attn_bwd(qkv, None, out, dout, dqkv, heads, True, 1.0 / math.sqrt(d), lse=lse)
    # GroupNorm family at the 64x64 level, 128 channels
    N, HW, C = 256, 4096, 128
    x = torch.randn(N, HW, C, device=dev).to(bf)
    dy = torch.randn(N, HW, C, device=dev).to(bf)
    ss = torch.randn(N * HW, 2 * C, device=dev).to(bf)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    sums = torch.zeros(N, 32, 2, device=dev)
    out = torch.empty(N, HW, C, device=dev, dtype=bf)
    ops.gn_stats(x, None, sums)
    ops.gn_apply(x, None, sums, gamma, beta, 1, out)
    ops.gn_apply(x, None, sums, gamma, beta, 2, out, ss=ss, F=C, drop_p=0.1, seed=1)
    for mode in (1, 2):
        dg, db, ws = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(N, 32, 2, device=dev)
        dx = torch.empty_like(out)
        ops.gn_bwd(x, None, sums, gamma, beta, mode, dy, dg, db, ws, dx, ss=ss if mode == 2 else None, F=C, dss=torch.empty_like(ss) if mode == 2 else None,
                   drop_p=0.1 if mode == 2 else 0.0, seed=1, add0=dy)
    # pose embedding (K1), batch 128
    B = 128
    _, R, T, K = synthetic.make_views(B, 64, seed=0, device=dev)
    pe = torch.empty(2 * B, 64, 64, 144, device=dev, dtype=bf)
    pos = torch.randn(144, 64, 64, device=dev)
    ops.pose_embedding(K, R, T, torch.ones(B, dtype=torch.uint8, device=dev), pos, torch.randn(144, device=dev), torch.randn(144, device=dev), pe)
    # diffusion glue, batch 128
    z, noise, xx, xr = (torch.randn(B, 3, 64, 64, device=dev) for _ in range(4))
    zo, xo, dp = torch.empty_like(z), torch.empty_like(z), torch.empty_like(z)
    ops.q_sample(z, noise, torch.randn(B, device=dev), zo, xx, xr, torch.ones(B, dtype=torch.uint8, device=dev), xo)
    loss = torch.zeros(1, device=dev)
    ops.loss(noise, z, dp, 0, ops.loss_scratch(), loss)
    # Adam over the CFG-64 arena
    n = 136_700_000
    p, g, m, v = (torch.zeros(n, device=dev) for _ in range(4))
    ops.adam(p, g, m, v, 1e-4, 0.9, 0.99, 1e-8, 1)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
