"""Microbenchmark of the HBM-bound GroupNorm kernels at the shapes of the model (achieved GB/s vs algorithmic bytes).
    python tests/bench_gn.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402


def timeit(fn, iters=5, reps=10):
    """ms per call, timed over CUDA-graph replays of `reps` calls (the Python launch cost would hide the small kernels)"""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(reps):
                fn()
        g.replay()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            g.replay()
        e1.record(st)
        st.synchronize()
    return e0.elapsed_time(e1) / (iters * reps)


def main():
    ops = CudaOps("cuda:0")
    dev = "cuda:0"
    shapes = [(256, 4096, 128, 0), (256, 4096, 256, 128), (256, 1024, 256, 0), (256, 256, 256, 0), (256, 64, 512, 0), (256, 64, 512, 512),
              (32, 4096, 128, 0), (32, 1024, 256, 0), (32, 256, 256, 0), (32, 64, 512, 0), (32, 64, 512, 512)]   # N = 32: the 8-GPU shard
    if len(sys.argv) > 1:
        shapes = shapes[:int(sys.argv[1])]
    for N, HW, C0, C1 in shapes:
        C = C0 + C1
        x0 = torch.randn(N, HW, C0, device=dev).bfloat16()
        x1 = torch.randn(N, HW, C1, device=dev).bfloat16() if C1 else None
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        sums = torch.zeros(N, 32, 2, device=dev)
        ops.gn_stats(x0, x1, sums)
        ss = torch.randn(N * HW, 2 * C, device=dev).bfloat16()
        out = torch.empty(N, HW, C, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(N, HW, C, device=dev).bfloat16()
        dg, db, ws = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(N, 32, 2, device=dev)
        dx0 = torch.empty(N, HW, C0, device=dev, dtype=torch.bfloat16)
        dx1 = torch.empty(N, HW, C1, device=dev, dtype=torch.bfloat16) if C1 else None
        dss = torch.empty(N * HW, 2 * C, device=dev, dtype=torch.bfloat16)
        add0 = torch.randn(N, HW, C, device=dev).bfloat16()
        el = N * HW * C * 2 / 1e9           # GB per full read or write of the bf16 tensor
        t = timeit(lambda: ops.gn_stats(x0, x1, sums))
        print(f"N={N} HW={HW} C={C0}+{C1}: stats      {t:7.3f} ms  {el / t * 1e3:7.0f} GB/s (1 read)")
        for mode, name, rw in ((1, "apply silu", 2), (2, "apply film", 4)):
            t = timeit(lambda: ops.gn_apply(x0, x1, sums, gamma, beta, mode, out, ss=ss if mode == 2 else None, F=C, drop_p=0.1 if mode == 2 else 0.0, seed=1))
            print(f"                              {name:10s} {t:7.3f} ms  {rw * el / t * 1e3:7.0f} GB/s ({rw} tensor passes)")
        for mode, name, rw in ((1, "bwd silu", 6), (2, "bwd film", 12)):
            def f():
                ops.gn_bwd(x0, x1, sums, gamma, beta, mode, dy, dg, db, ws, dx0, dx1, ss=ss if mode == 2 else None, F=C,
                           dss=dss if mode == 2 else None, drop_p=0.1 if mode == 2 else 0.0, seed=1, add0=add0)
            t = timeit(f)
            print(f"                              {name:10s} {t:7.3f} ms  {rw * el / t * 1e3:7.0f} GB/s ({rw} tensor passes as implemented)")


if __name__ == "__main__":
    main()
