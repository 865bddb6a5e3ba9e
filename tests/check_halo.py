"""GPU check + timing of the halo-patch convolution form against the tile-load form (same descriptors, b3d_set_conv_halo).
    python tests/check_halo.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gemm_cases as gc  # noqa: E402
from distributed_3d_diffusion_pytorch_b200 import _lib  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402

CASES = [
    ("single CTA 128->128 @64x64 N=2", gc.conv_case("c", 2, 64, 64, [128], [9], 128, resid=True, out_scale=0.7071)),
    ("single CTA 64->256 @32x32 N=2", gc.conv_case("c", 2, 32, 32, [64], [9], 256)),
    ("single CTA 8->128 stem @32x32", gc.conv_case("c", 2, 32, 32, [8], [9], 128)),
    ("single CTA 3 sources [9,1,1] @16x16", gc.conv_case("c", 2, 16, 16, [128, 64, 64], [9, 1, 1], 128, rowvec=True)),
    ("two accumulators 128->128 @64x64 N=32", gc.conv_case("c", 32, 64, 64, [128], [9], 128)),
    ("pair x2 128->128 @64x64 N=256 (+resid)", gc.conv_case("c", 256, 64, 64, [128], [9], 128, resid=True, out_scale=0.7071)),
    ("pair x2 256+128 -> 128 @64x64 N=256", gc.conv_case("c", 256, 64, 64, [256, 128], [9, 9], 128)),
    ("pair wide 256->256 @32x32 N=256", gc.conv_case("c", 256, 32, 32, [256], [9], 256)),
    ("pair wide 128->256 @64x64 N=256", gc.conv_case("c", 256, 64, 64, [128], [9], 256)),
    ("pair wide over 2 images 256->256 @16x16 N=256", gc.conv_case("c", 256, 16, 16, [256], [9], 256)),
    ("pair wide over 2 images 768->256 @16x16 N=256", gc.conv_case("c", 256, 16, 16, [512, 256], [9, 9], 256)),
    ("pair x2 conv2+skip [9,1,1] 128 @64x64 N=256", gc.conv_case("c", 256, 64, 64, [128, 256, 128], [9, 1, 1], 128, rowvec=True, out_scale=0.7071)),
]


def timeit(ops, d, n=5):
    for _ in range(2):
        ops.gemm(d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        ops.gemm(d)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    ops = CudaOps("cuda:0")
    sel = [int(a) for a in sys.argv[1:]] or range(len(CASES))
    for i in sel:
        name, case = CASES[i]
        d, outs = case.build(ops)
        res = {}
        for halo in (0, 1, 2):
            _lib.set_conv_halo(halo)
            for o in outs:
                o.zero_()
            ops.gemm(d)
            torch.cuda.synchronize()
            res[halo] = ([o.clone() for o in outs], timeit(ops, d))
        _lib.set_conv_halo(0)
        err = [max(((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item() for a, b in zip(res[h][0], res[0][0])) for h in (1, 2)]
        flop = ops.gemm_flop(d)
        print(f"{name:48s} rel diff {err[0]:.1e} / {err[1]:.1e}   tile-load {res[0][1]:.3f} ms ({flop / res[0][1] / 1e9:5.0f} TF/s)   halo pitch 2048 "
              f"{res[1][1]:.3f} ms ({flop / res[1][1] / 1e9:5.0f})   packed {res[2][1]:.3f} ms ({flop / res[2][1] / 1e9:5.0f})", flush=True)
        del d, outs, res
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
