"""Microbenchmark of the pose-embedding kernel K1 (HBM-bound: writes 2*144*H*W bf16 per pair).  python tests/bench_pose.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_gn import timeit  # noqa: E402
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402

if __name__ == "__main__":
    ops, dev = CudaOps("cuda:0"), "cuda:0"
    for B, S in ((128, 64), (16, 64), (32, 128)):
        _, R, T, K = synthetic.make_views(B, S, seed=0, device=dev)
        pe = torch.empty(2 * B, S, S, 144, device=dev, dtype=torch.bfloat16)
        pos, fe, oe = torch.randn(144, S, S, device=dev), torch.randn(144, device=dev), torch.randn(144, device=dev)
        mask = torch.ones(B, dtype=torch.uint8, device=dev)
        t = timeit(lambda: ops.pose_embedding(K, R, T, mask, pos, fe, oe, pe))
        gb = pe.numel() * 2 / 1e9
        print(f"pose_embedding B={B} {S}x{S}: {t * 1e3:7.1f} us  {gb / t * 1e3:7.0f} GB/s of {gb * 1e3:.0f} MB written")
