"""Sweep of the image-chunk budget of the two-pass GroupNorm launches (b3d_set_gn_chunk_mb): does the second pass hit L2?
    python tests/bench_gn_chunk.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_gn import timeit  # noqa: E402
from distributed_3d_diffusion_pytorch_b200 import _lib  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402


def main():
    ops, dev = CudaOps("cuda:0"), "cuda:0"
    shapes = [(256, 4096, 128, 0), (256, 4096, 256, 128), (256, 1024, 256, 0), (256, 1024, 512, 256), (256, 256, 256, 0)]
    budgets = [0, 4, 8, 16, 24, 32, 48, 64]
    for N, HW, C0, C1 in shapes:
        C = C0 + C1
        x0 = torch.randn(N, HW, C0, device=dev).bfloat16()
        x1 = torch.randn(N, HW, C1, device=dev).bfloat16() if C1 else None
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        sums = torch.zeros(N, 32, 2, device=dev)
        ss = torch.randn(N * HW, 2 * C, device=dev).bfloat16()
        out = torch.empty(N, HW, C, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(N, HW, C, device=dev).bfloat16()
        dg, db, ws = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(N, 32, 2, device=dev)
        dx0 = torch.empty(N, HW, C0, device=dev, dtype=torch.bfloat16)
        dx1 = torch.empty(N, HW, C1, device=dev, dtype=torch.bfloat16) if C1 else None
        dss = torch.empty(N * HW, 2 * C, device=dev, dtype=torch.bfloat16)
        add0 = torch.randn(N, HW, C, device=dev).bfloat16()
        mb = N * HW * C * 2 / 2 ** 20
        print(f"N={N} HW={HW} C={C0}+{C1} ({mb:.0f} MB per tensor); ms per call by chunk budget {budgets} MB")
        rows = {"fwd silu": [], "fwd film": [], "bwd silu": [], "bwd film": []}
        for b in budgets:
            _lib.set_gn_chunk_mb(b)
            for mode, key in ((1, "fwd silu"), (2, "fwd film")):
                def f():
                    sums.zero_()
                    ops.gn_stats_apply(x0, x1, sums, gamma, beta, mode, out, ss=ss if mode == 2 else None, F=C, drop_p=0.1 if mode == 2 else 0.0, seed=1)
                rows[key].append(timeit(f, iters=3, reps=4))
            for mode, key in ((1, "bwd silu"), (2, "bwd film")):
                def f():
                    ws.zero_()
                    ops.gn_bwd(x0, x1, sums, gamma, beta, mode, dy, dg, db, ws, dx0, dx1, ss=ss if mode == 2 else None, F=C,
                               dss=dss if mode == 2 else None, drop_p=0.1 if mode == 2 else 0.0, seed=1, add0=add0)
                rows[key].append(timeit(f, iters=3, reps=4))
        _lib.set_gn_chunk_mb(0)
        for k, v in rows.items():
            print(f"   {k}: " + "  ".join(f"{t:7.3f}" for t in v))


if __name__ == "__main__":
    main()
