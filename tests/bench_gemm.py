"""Microbenchmark of the tcgen05 kernel on the dominant shapes of the CFG-64 training step at batch 128 (N = 256 images):
CUDA-event time, achieved TFLOP/s and algorithmic bytes.  Also the command profiled with `ncu --set full` (profiles/).
    python tests/bench_gemm.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gemm_cases as gc  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402

CASES = [
    # name, builder, algorithmic bytes (inputs read once + output written once)
    ("conv3x3 128->128 @64x64 N=256 (+resid)", gc.conv_case("c", 256, 64, 64, [128], [9], 128, resid=True, out_scale=0.7071)),
    ("conv3x3 256->128 @64x64 N=256", gc.conv_case("c", 256, 64, 64, [256], [9], 128)),
    ("conv3x3 256->256 @32x32 N=256", gc.conv_case("c", 256, 32, 32, [256], [9], 256)),
    ("conv3x3 512->512 @8x8 N=256", gc.conv_case("c", 256, 8, 8, [512], [9], 512)),
    ("cond conv 144->1024 @64x64 N=256 (rowvec, silu out2)", gc.conv_case("c", 256, 64, 64, [144], [9], 1024, rowvec=True, out2=True)),
    ("FiLM GEMM [1M,1024]x[2048,1024]", gc.dense_case("d", 1048576, 1024, 2048)),
    ("wgrad conv 128x(9x128) over 1M pixels", gc.wgrad_case("w", 256, 64, 64, 128, 128, 9)),
    ("wgrad conv 256x(9x256) over 262k pixels", gc.wgrad_case("w", 256, 32, 32, 256, 256, 9)),
    ("wgrad FiLM 256x1024 over 1M pixels", gc.wgrad_case("w", 1, 1, 1048576, 256, 1024, 1)),
    ("attention-block dense [65536,256]x[256,256]", gc.dense_case("d", 65536, 256, 256)),
]


SMALL = [
    # the short-K / few-tile launches of the attention blocks: timed inside a CUDA graph (20 launches per replay) so the
    # Python launch cost does not hide the kernel's own fixed costs
    ("dense [65536,256]x[256,256]", gc.dense_case("d", 65536, 256, 256)),
    ("dense [65536,256]x[768,256]", gc.dense_case("d", 65536, 256, 768)),
    ("dense [16384,512]x[512,512]", gc.dense_case("d", 16384, 512, 512)),
    ("dense [8192,256]x[256,256]", gc.dense_case("d", 8192, 256, 256)),
    ("dense [1024,256]x[256,256]", gc.dense_case("d", 1024, 256, 256)),
    ("dense [128,256]x[256,256]", gc.dense_case("d", 128, 256, 256)),
    ("wgrad dense 256x256 over 65536 rows", gc.wgrad_case("w", 1, 1, 65536, 256, 256, 1)),
    ("wgrad dense 512x512 over 16384 rows", gc.wgrad_case("w", 1, 1, 16384, 512, 512, 1)),
]


def small():
    ops = CudaOps("cuda:0")
    for name, case in SMALL:
        d, outs = case.build(ops)
        flop = ops.gemm_flop(d)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for _ in range(3):
                ops.gemm(d)
            st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for _ in range(20):
                    ops.gemm(d)
            g.replay()
            st.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(5):
                g.replay()
            e1.record(st)
            st.synchronize()
        us = e0.elapsed_time(e1) / 100 * 1e3
        print(f"{name:45s} {us:8.2f} us/launch  {flop / us / 1e6:8.1f} TFLOP/s", flush=True)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "small":
        return small()
    ops = CudaOps("cuda:0")
    sel = int(sys.argv[1]) if len(sys.argv) > 1 else None
    for i, (name, case) in enumerate(CASES):
        if sel is not None and i != sel:
            continue
        d, outs = case.build(ops)
        flop = ops.gemm_flop(d)
        by = sum(t.numel() * t.element_size() for t in d.a) + d.b.numel() * d.b.element_size() + d.out.numel() * d.out.element_size()
        if d.resid is not None:
            by += d.resid.numel() * 2
        if d.out2 is not None:
            by += d.out2.numel() * 2
        for _ in range(2):
            ops.gemm(d)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 5
        e0.record()
        for _ in range(n):
            ops.gemm(d)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"{name:55s} {ms:8.3f} ms  {flop / ms / 1e9:8.1f} TFLOP/s  alg bytes {by / 1e6:9.1f} MB  ({by / ms / 1e6:7.1f} GB/s)", flush=True)
        del d, outs
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
