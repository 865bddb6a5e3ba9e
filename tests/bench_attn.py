"""Microbenchmark of the fused attention kernels at the CFG-64 shapes (graph-timed).
    python tests/bench_attn.py"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402


def timeit(fn, reps=10, iters=5):
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
    return e0.elapsed_time(e1) / (reps * iters) * 1e3


def main():
    ops = CudaOps("cuda:0")
    dev = "cuda:0"
    for N, L, heads, d in ((256, 256, 4, 64), (256, 64, 4, 128), (32, 256, 4, 64), (32, 64, 4, 128)):
        C = heads * d
        qkv = torch.randn(N, L, 3 * C, device=dev).bfloat16()
        out = torch.empty(N, L, C, device=dev, dtype=torch.bfloat16)
        Lp = (L + 7) // 8 * 8
        probs = torch.empty(N, heads, L, Lp, device=dev, dtype=torch.bfloat16)
        dout = torch.randn(N, L, C, device=dev).bfloat16()
        dqkv = torch.empty(N, L, 3 * C, device=dev, dtype=torch.bfloat16)
        sc = 1.0 / math.sqrt(d)
        flop_f = 4.0 * N * heads * L * L * d
        t0 = timeit(lambda: ops.attn_fwd(qkv, heads, True, sc, out, probs=None))
        t1 = timeit(lambda: ops.attn_fwd(qkv, heads, True, sc, out, probs=probs[..., :L]))
        t2 = timeit(lambda: ops.attn_bwd(qkv, probs[..., :L], out, dout, dqkv, heads, True, sc))
        lse = torch.empty(N, heads, L, device=dev)
        t3 = timeit(lambda: ops.attn_fwd(qkv, heads, True, sc, out, lse=lse))
        t4 = timeit(lambda: ops.attn_bwd(qkv, None, out, dout, dqkv, heads, True, sc, lse=lse))
        print(f"N={N} L={L} heads={heads} d={d}: fwd {t0:7.1f} us ({flop_f / t0 / 1e6:6.1f} TF/s)  fwd+P {t1:7.1f} us  bwd(P) {t2:7.1f} us "
              f"({2.5 * flop_f / t2 / 1e6:6.1f} TF/s)  fwd+lse {t3:7.1f} us  bwd(lse, P recomputed) {t4:7.1f} us", flush=True)


if __name__ == "__main__":
    main()
