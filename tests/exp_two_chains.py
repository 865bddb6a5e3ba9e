"""Experiment: does running the per-GPU batch as TWO concurrent half-batch chains (two streams) beat one chain?
Two independent trainers of batch B/2 replay their step graphs on two streams at once; compared with one trainer of batch B.
    python tests/exp_two_chains.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.trainer import Trainer  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    dev = torch.device("cuda", 0)
    kw = dict(H=64, W=64, ch=128)

    def make(b, seed):
        tr = Trainer(kw, b, dev, seed=seed)
        batch = tuple(t.to(dev) for t in synthetic.make_views(b, 64, seed=seed))
        for _ in range(3):
            tr.step_resident(batch)
        torch.cuda.synchronize()
        assert tr.graph is not None
        return tr

    one = make(B, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        one.graph[0].replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"one chain of batch {B}: {e0.elapsed_time(e1) / n:.3f} ms per step (graph only, no Adam)")
    del one
    torch.cuda.empty_cache()
    a, b = make(B // 2, 2), make(B // 2, 3)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    main_s = torch.cuda.current_stream()
    for trial in range(2):
        e0.record()
        for _ in range(n):
            s1.wait_stream(main_s)
            s2.wait_stream(main_s)
            with torch.cuda.stream(s1):
                a.graph[0].replay()
            with torch.cuda.stream(s2):
                b.graph[0].replay()
            main_s.wait_stream(s1)
            main_s.wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        print(f"two concurrent chains of batch {B // 2}: {e0.elapsed_time(e1) / n:.3f} ms per step pair")
    e0.record()
    for _ in range(n):
        a.graph[0].replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"one chain of batch {B // 2} alone: {e0.elapsed_time(e1) / n:.3f} ms")


if __name__ == "__main__":
    main()
