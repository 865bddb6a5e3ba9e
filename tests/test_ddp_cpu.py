"""CPU, world_size 2 over gloo: the data-parallel semantics of the training step (shard the batch, mean-all-reduce
the flat gradient arena) give the same gradient as one process on the whole batch.  The kernels are replaced by their
torch statements (tests/emu_ops.py); the plumbing under test is distributed_3d_diffusion_pytorch_b200.parallel."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _grads(B_slice, seed=7):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    from distributed_3d_diffusion_pytorch_b200.engine import Cfg, Engine
    from distributed_3d_diffusion_pytorch_b200.xunet import param_shapes
    from emu_ops import EmuOps
    kw = dict(H=8, W=8, ch=32)
    cfg = Cfg(**kw)
    g = torch.Generator().manual_seed(seed)
    sd = {k: torch.randn(s, generator=g) * 0.05 + (1.0 if k.endswith("gn.weight") else 0.0) for k, s in param_shapes(cfg).items()}
    B = 4
    img, R, T, K = synthetic.make_views(B, 8, seed=3)
    logsnr = torch.stack([torch.full((B,), 20.0), torch.linspace(-3, 3, B)], 1)
    noise = torch.randn(B, 3, 8, 8, generator=g)
    cm = torch.tensor([True, False, True, True])
    sl = B_slice
    eng = Engine(cfg, EmuOps(), sl.stop - sl.start)
    eps = eng.forward(sd, img[sl, 0], img[sl, 1], logsnr[sl], R[sl], T[sl], K[sl], cm[sl], training=False, need_bwd=True)
    n = noise[sl].numel()
    deps = 2 * (eps - noise[sl]) / n                      # d(mean squared error over the LOCAL shard)
    names = list(sd.keys())
    offs, tot = [], 0
    for k in names:
        offs.append(tot)
        tot += sd[k].numel()
    flat = torch.zeros(tot)
    Gr = {k: flat[o:o + sd[k].numel()].view(sd[k].shape) for k, o in zip(names, offs)}
    eng.backward(Gr, deps, grad_flat=flat)
    return flat


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(HERE))
    from distributed_3d_diffusion_pytorch_b200 import parallel
    full = (torch.arange(4),)
    lo = parallel.shard_batch(full, rank, world)[0]
    flat = _grads(slice(int(lo[0]), int(lo[-1]) + 1))
    w0 = torch.full((3,), float(rank))
    parallel.broadcast_parameters(w0)
    assert torch.equal(w0, torch.zeros(3))
    # the trainer's per-range update (trainer._update): a bf16 copy of each gradient range is SUM-reduced and Adam reads it in
    # place with grad_scale = 1 / world -- the same step as Adam on the fp32 mean gradient, up to the bf16 rounding of the payload
    from distributed_3d_diffusion_pytorch_b200.engine import Cfg, Engine
    from distributed_3d_diffusion_pytorch_b200.trainer import phase_ranges
    from distributed_3d_diffusion_pytorch_b200.xunet import param_shapes
    from emu_ops import EmuOps
    shapes = param_shapes(Cfg(H=8, W=8, ch=32))
    names, offs, tot = list(shapes.keys()), [], 0
    for k in names:
        offs.append(tot)
        tot += int(torch.Size(shapes[k]).numel())
    ranges = phase_ranges(names, offs, tot, Engine.PHASE_PREFIXES)
    local = flat.clone()
    p_a, m_a, v_a = torch.zeros(tot), torch.zeros(tot), torch.zeros(tot)
    for ph in (0, 1, 2):
        lo, hi = ranges[ph]
        gb = local[lo:hi].to(torch.bfloat16)
        parallel.allreduce_sum_(gb)
        EmuOps().adam(p_a[lo:hi], gb, m_a[lo:hi], v_a[lo:hi], 1e-3, 0.9, 0.99, 1e-8, 1, grad_scale=1.0 / world)
    parallel.allreduce_mean_(flat)
    p_b, m_b, v_b = torch.zeros(tot), torch.zeros(tot), torch.zeros(tot)
    EmuOps().adam(p_b, flat, m_b, v_b, 1e-3, 0.9, 0.99, 1e-8, 1)
    assert float((m_a - m_b).norm() / m_b.norm()) < 5e-3                # 2^-9 per rounded summand
    # first Adam step: the update is lr * g / (|g| + eps), i.e. the SIGN of the mean gradient wherever it is above eps: the
    # two can differ only where the ranks' gradients nearly cancel and the bf16 rounding decides the sign
    assert float(((p_a - p_b).abs() > 1e-5).float().mean()) < 5e-3
    assert float(p_b.abs().max()) > 5e-4 and float((v_a - v_b).norm() / v_b.norm()) < 1e-2
    if rank == 0:
        torch.save(flat, out)
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_data_parallel_matches_single_process(tmp_path):
    out = str(tmp_path / "g.pt")
    port = 29500 + os.getpid() % 500
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    g2 = torch.load(out)
    g1 = _grads(slice(0, 4))
    assert torch.allclose(g1, g2, atol=3e-4, rtol=1e-3), (g1 - g2).abs().max()   # fp32 noise: 1x1-pixel GroupNorm at the deepest level
