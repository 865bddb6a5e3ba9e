"""Microbenchmark of the per-step, batch-independent work over the parameters of CFG-64: weight repack (fp32 PyTorch
layout -> bf16 GEMM layouts), gradient arena zeroing, packed-gradient unpack, fused Adam.
    python tests/bench_param.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.trainer import Trainer  # noqa: E402


def timeit(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = "cuda:0"
    B = 2
    tr = Trainer(dict(H=64, W=64, ch=128), B, dev, use_graph=False)
    batch = tuple(t.to(dev) for t in synthetic.make_views(B, 64, seed=1))
    tr.step_resident(batch)
    tr.step_resident(batch)
    torch.cuda.synchronize()
    m, eng = tr.model, tr.engine
    n = m._flat.numel()
    P = dict(zip(m._names, m._plist))
    print(f"parameters: {n / 1e6:.1f} M")
    t = timeit(lambda: tr.opt.step())
    print(f"fused Adam            {t:7.3f} ms   {n * 28 / t / 1e6:7.0f} GB/s (28 B/param)")
    t = timeit(lambda: eng._pack_weights(P, True, repack=True))
    print(f"weight repack         {t:7.3f} ms")
    t = timeit(lambda: m._grad_flat.zero_())
    print(f"gradient arena zero   {t:7.3f} ms   {n * 4 / t / 1e6:7.0f} GB/s")
    for ph in range(3):
        t = timeit(lambda: (eng.ops.run_jobs(eng._unpack_jl[ph]), eng.ops.run_jobs(eng._copy_jl[ph])))
        print(f"gradient unpack ph{ph}   {t:7.3f} ms")
    t = timeit(lambda: eng._zero_phase("bwd"))
    print(f"packed-grad arena zero {t:7.3f} ms")


if __name__ == "__main__":
    main()
