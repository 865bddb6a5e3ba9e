"""CPU: the three gradient ranges that the multi-GPU trainer all-reduces while the next backward phase computes
(trainer.phase_ranges / engine.PHASE_PREFIXES) tile the flat arena exactly once, and each phase of the engine's backward
leaves exactly its own range final -- a wrong range would leave a slice of the gradient un-reduced and the ranks would
drift apart without any error."""
import torch

from distributed_3d_diffusion_pytorch_b200 import synthetic
from distributed_3d_diffusion_pytorch_b200.engine import Cfg, Engine
from distributed_3d_diffusion_pytorch_b200.trainer import phase_ranges
from distributed_3d_diffusion_pytorch_b200.xunet import XUNet
from emu_ops import EmuOps


def test_phase_ranges_tile_the_arena_once():
    for kw in (dict(H=8, W=8, ch=32), dict(H=64, W=64, ch=128)):
        m = XUNet(**kw)
        numel = m.flat_parameters().numel()
        r = phase_ranges(m._names, m._offs, numel, Engine.PHASE_PREFIXES)
        assert r[2][0] == 0 and r[2][1] == r[1][0] and r[1][1] == r[0][0] and r[0][1] == numel
        assert all(lo < hi for lo, hi in r.values())
        for n, o, p in zip(m._names, m._offs, m._plist):
            ph = [i for i, pre in enumerate(Engine.PHASE_PREFIXES) if n.startswith(pre)]
            assert len(ph) == 1
            lo, hi = r[ph[0]]
            assert lo <= o and o + p.numel() <= hi, n


def test_each_backward_phase_finalises_exactly_its_range():
    kw, B = dict(H=8, W=8, ch=32), 2
    m = XUNet(**kw)
    g = torch.Generator().manual_seed(3)
    sd = {k: torch.randn(p.shape, generator=g) * 0.05 + (1.0 if k.endswith("gn.weight") else 0.0) for k, p in zip(m._names, m._plist)}
    img, R, T, K = synthetic.make_views(B, 8, seed=1)
    logsnr = torch.stack([torch.full((B,), 20.0), torch.tensor([-1.0, 2.0])], 1)
    cm = torch.tensor([True, False])
    deps = torch.randn(B, 3, 8, 8, generator=g) * 0.01
    numel = m.flat_parameters().numel()
    r = phase_ranges(m._names, m._offs, numel, Engine.PHASE_PREFIXES)

    def views(flat):
        return {n: flat[o:o + p.numel()].view(p.shape) for n, o, p in zip(m._names, m._offs, m._plist)}

    eng = Engine(Cfg(**kw), EmuOps(), B)
    eng.forward(sd, img[:, 0], img[:, 1], logsnr, R, T, K, cm, training=False, need_bwd=True)
    full = torch.zeros(numel)
    eng.backward(views(full), deps, grad_flat=full)
    assert all(full[lo:hi].abs().sum() > 0 for lo, hi in r.values())
    eng.forward(sd, img[:, 0], img[:, 1], logsnr, R, T, K, cm, training=False, need_bwd=True)
    flat = torch.zeros(numel)
    gv = views(flat)
    done = []
    for ph in (0, 1, 2):
        eng.backward(gv, deps, grad_flat=flat, phase=ph)
        done.append(ph)
        for q in (0, 1, 2):
            lo, hi = r[q]
            if q in done:          # final: equal to the whole backward's result, and later phases must not touch it again
                assert torch.equal(flat[lo:hi], full[lo:hi]), (ph, q)
            else:                  # not started: still the zeros phase 0 wrote
                assert flat[lo:hi].abs().sum() == 0, (ph, q)
    # phase 2 in its two parts ("2a": stem + the FiLM data gradient, the last reads of the level-wide FiLM operands; "2b": the
    # rest) gives the same gradients, and "2a" alone leaves range 2 unfinished (its unpack / copy jobs run at the end of "2b")
    eng.forward(sd, img[:, 0], img[:, 1], logsnr, R, T, K, cm, training=False, need_bwd=True)
    flat2 = torch.zeros(numel)
    gv2 = views(flat2)
    for ph in (0, 1, "2a"):
        eng.backward(gv2, deps, grad_flat=flat2, phase=ph)
    lo, hi = r[2]
    assert not torch.equal(flat2[lo:hi], full[lo:hi])
    assert torch.equal(flat2[hi:], full[hi:])
    eng.backward(gv2, deps, grad_flat=flat2, phase="2b")
    assert torch.equal(flat2, full)
