"""BASELINE configs[4]: microbenchmark sweep of the building blocks -- conv3x3 (forward = data-gradient shape, weight gradient),
the FiLM GEMM, GroupNorm (+SiLU) forward / backward and the cross-frame attention forward / backward -- at the four
resolution levels of CFG-64 (64^2 x 128 ch, 32^2 x 256, 16^2 x 256, 8^2 x 512) for 1 / 16 / 128 pairs (2 / 32 / 256 images).
Graph-timed (10 launches per replay, 5 replays), inputs created on the device.  One line per (op, level, batch):
time per launch and achieved TFLOP/s (tensor-bound ops) or GB/s of algorithmic bytes (HBM-bound ops).

    python tests/bench_sweep.py [> profiles/rNN_microbench_sweep.txt]
"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import gemm as G  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402

DEV = "cuda:0"
BF = torch.bfloat16


def timeit(fn, reps=10, iters=5):
    """us per call"""
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


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(BF)


def line(op, res, N, us, flop=None, nbytes=None):
    rate = f"{flop / us / 1e6:8.1f} TFLOP/s" if flop else f"{nbytes / us / 1e3:8.0f} GB/s   "
    print(f"{op:28s} {res:2d}x{res:<2d} images={N:3d}  {us:9.1f} us  {rate}", flush=True)


def attempt(fn):
    try:
        fn()
    except Exception as e:      # keep the sweep going: one unsupported shape must not cost the rest
        print(f"  (skipped: {type(e).__name__}: {str(e)[:120]})", flush=True)
    torch.cuda.empty_cache()


def main():
    ops = CudaOps(DEV)
    levels = [(64, 128), (32, 256), (16, 256), (8, 512)]
    batches = [2, 32, 256]
    for res, C in levels:
        for N in batches:
            M = N * res * res

            def conv():
                x, w, out = rnd(N, res, res, C), rnd(C, 9 * G.ceil64(C), scale=0.02), torch.empty(N, res, res, C, device=DEV, dtype=BF)
                bias = torch.zeros(C, device=DEV)
                d = G.conv_desc([x], [9], w, C, out, bias=bias)
                line(f"conv3x3 {C}->{C} fwd/dgrad", res, N, timeit(lambda: ops.gemm(d)), flop=2.0 * M * C * C * 9)

            def wgrad():
                dy, x = rnd(N, res, res, C, scale=0.25), rnd(N, res, res, C, scale=0.25)
                out = torch.zeros(C, 9 * G.ceil64(C), device=DEV)
                d = G.wgrad_desc(dy, x, 9, out, 0, G.ceil64(C))
                line(f"conv3x3 {C}->{C} wgrad", res, N, timeit(lambda: ops.gemm(d)), flop=2.0 * M * C * C * 9)

            def film():
                x, w, out = rnd(M, 1024), rnd(2 * C, 1024, scale=0.03), torch.empty(M, 2 * C, device=DEV, dtype=BF)
                d = G.dense_desc(x, w, out, 2 * C, bias=torch.zeros(2 * C, device=DEV))
                line(f"FiLM dense 1024->{2 * C}", res, N, timeit(lambda: ops.gemm(d)), flop=2.0 * M * 1024 * 2 * C)

            def gn():
                HW = res * res
                x, dy = rnd(N, HW, C), rnd(N, HW, C)
                gamma, beta = torch.ones(C, device=DEV), torch.zeros(C, device=DEV)
                sums = torch.zeros(N, 32, 2, device=DEV)
                out = torch.empty(N, HW, C, device=DEV, dtype=BF)
                dg, db, ws = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV), torch.zeros(N, 32, 2, device=DEV)
                dx = torch.empty(N, HW, C, device=DEV, dtype=BF)
                el = N * HW * C * 2.0

                def fwd():
                    sums.zero_()
                    ops.gn_stats_apply(x, None, sums, gamma, beta, 1, out)
                line(f"GroupNorm+SiLU fwd C={C}", res, N, timeit(fwd), nbytes=2 * el)          # algorithmic: one read, one write

                def bwd():
                    ws.zero_()
                    ops.gn_bwd(x, None, sums, gamma, beta, 1, dy, dg, db, ws, dx, None)
                line(f"GroupNorm+SiLU bwd C={C}", res, N, timeit(bwd), nbytes=3 * el)          # algorithmic: x, dy read, dx written

            def attn():
                L, heads = res * res, 4
                d_ = C // heads
                qkv, out = rnd(N, L, 3 * C), torch.empty(N, L, C, device=DEV, dtype=BF)
                dout, dqkv = rnd(N, L, C), torch.empty(N, L, 3 * C, device=DEV, dtype=BF)
                lse = torch.empty(N, heads, L, device=DEV)
                sc = 1.0 / math.sqrt(d_)
                fl = 4.0 * N * heads * L * L * d_
                line(f"cross-attn fwd L={L} d={d_}", res, N, timeit(lambda: ops.attn_fwd(qkv, heads, True, sc, out, lse=lse)), flop=fl)
                line(f"cross-attn bwd L={L} d={d_}", res, N, timeit(lambda: ops.attn_bwd(qkv, None, out, dout, dqkv, heads, True, sc, lse=lse)),
                     flop=2.5 * fl)

            for fn in (conv, wgrad, film, gn) + ((attn,) if res <= 16 else ()):
                attempt(fn)


if __name__ == "__main__":
    main()
