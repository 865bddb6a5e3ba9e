"""-m gpu: the graphed training step (trainer.py).  The update of each gradient range -- Adam and the repack of the bf16 GEMM
operands built from those parameters -- runs on a second stream beside the later backward phases, and the forward no longer
packs; these tests pin the two invariants that makes safe to rely on."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
KW = dict(H=16, W=16, ch=32, ch_mult=(1, 2), emb_ch=1024, num_res_blocks=1, attn_resolutions=(8,), attn_heads=2, dropout=0.0)


def _trainer(use_graph, seed=5, **kw):
    from distributed_3d_diffusion_pytorch_b200.trainer import Trainer
    return Trainer(dict(KW, **kw), 4, DEV, world=1, seed=seed, use_graph=use_graph, lr=1e-3)


def _batches(n, H=16):
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    return [tuple(t.to(DEV) for t in synthetic.make_views(4, H, seed=40 + i)) for i in range(n)]


def test_operands_after_graphed_steps_equal_a_fresh_pack():
    """every packed operand the next forward reads is the bf16 image of the CURRENT fp32 parameters: the per-range repacks
    (groups 0-2 behind their Adam step, the FiLM group after phase 2) together cover every job of the forward's own lists"""
    tr = _trainer(True)
    bs = _batches(2)
    for i in range(5):
        tr.step_resident(bs[i % 2])
    torch.cuda.synchronize()
    assert tr.graphed and tr._packed
    eng = tr.engine
    W = eng._pack_cache[True][1]
    got = {k: v.clone() for k, v in W.items()}
    film_b = {k: v.clone() for k, v in eng._bufs.items() if k.startswith("b.film")}
    P = dict(zip(tr.model._names, tr.model._plist))
    eng._pack_weights(P, True, repack=True)
    torch.cuda.synchronize()
    for k, v in W.items():
        assert torch.equal(got[k], v), k
    for k, v in film_b.items():
        assert torch.equal(v, eng._bufs[k]), k
    # ... and the job lists agree: same (source, destination) pairs in the forward lists and in the four groups
    first, rest = eng._pack_cache[True][2]
    groups = eng._pack_cache[True][3]
    assert sorted(first.signature() + rest.signature()) == sorted(sum((g.signature() for g in groups), ()))


def test_graphed_step_gradients_and_update_equal_the_plain_sequence():
    """One graphed step, taken apart: (1) its gradients -- computed while the second stream already updates and repacks the
    finished ranges -- equal those of a plain forward (which packs) + backward on the same parameters and the same draws, to
    within the noise of the fp32 atomics; (2) the parameters after the step are Adam(p, g, m, v) over the WHOLE arena, every
    element updated exactly once with its final gradient."""
    kw = dict(H=32, W=32, ch=64, ch_mult=(1, 2, 2), emb_ch=1024, num_res_blocks=2, attn_resolutions=(8, 16), attn_heads=2, dropout=0.0)
    a, b = _trainer(True, seed=9, **kw), _trainer(False, seed=9, **kw)
    bs = _batches(2, 32)
    for i in range(4):
        a.step_resident(bs[i % 2])
    torch.cuda.synchronize()
    p0 = a.model.flat_parameters().clone()
    m0, v0, t0 = a.opt.m.clone(), a.opt.v.clone(), a.opt.step_count
    a.step_resident(bs[0])
    torch.cuda.synchronize()
    g = a.model._grad_flat.clone()
    p1 = a.model.flat_parameters().clone()
    # (1) the plain sequence on trainer b: a's parameters before the step, a's draws
    b.model.flat_parameters().copy_(p0)
    b._set_batch(bs[0])
    for n in ("step_ctr", "logsnr", "noise", "mask", "xrand"):
        getattr(b, n).copy_(getattr(a, n))
    b._forward_loss()
    b._backward()
    torch.cuda.synchronize()
    gb = b.model._grad_flat
    assert float(a.loss) == pytest.approx(float(b.loss), rel=1e-3)
    cos = float(torch.dot(g, gb) / (g.norm() * gb.norm()))
    assert cos > 0.9999, cos
    assert float((g - gb).norm() / gb.norm()) < 1e-2
    for ph, (lo, hi) in a.ranges.items():            # ... range by range: a stale operand in one phase must not hide in the total
        c = float(torch.dot(g[lo:hi], gb[lo:hi]) / (g[lo:hi].norm() * gb[lo:hi].norm()))
        assert c > 0.9999, (ph, c)
    # (2) Adam over the whole arena (torch.optim.Adam's arithmetic, lr 1e-3, betas (0.9, 0.99), eps 1e-8)
    t = t0 + 1
    assert a.opt.step_count == t
    m1 = 0.9 * m0 + 0.1 * g
    v1 = 0.99 * v0 + 0.01 * g * g
    want = p0 - 1e-3 * (m1 / (1 - 0.9 ** t)) / ((v1 / (1 - 0.99 ** t)).sqrt() + 1e-8)
    assert float((p1 - want).abs().max()) < 2e-6 + 1e-5 * float(want.abs().max())
    assert torch.allclose(a.opt.m, m1, rtol=1e-5, atol=1e-9) and torch.allclose(a.opt.v, v1, rtol=1e-5, atol=1e-12)


def test_parameters_changed_from_outside_are_repacked_before_the_next_step():
    """load_state_dict between steps bumps model._param_version: the trainer's forward, which no longer packs, must not run on
    the operands of the old weights"""
    tr = _trainer(True)
    bs = _batches(2)
    for i in range(3):
        tr.step_resident(bs[i % 2])
    calls = []
    orig = tr.repack
    tr.repack = lambda: (calls.append(1), orig())[1]
    tr.step_resident(bs[1])
    assert not calls                                   # its own update keeps the operands current
    sd = {k: v.clone() * 1.5 for k, v in tr.model.state_dict().items()}
    tr.model.load_state_dict(sd)
    torch.cuda.synchronize()
    eng = tr.engine
    W = eng._pack_cache[True][1]
    k0 = next(k for k in W if k.endswith("conv1.weight"))
    before = W[k0].clone()
    tr.step_resident(bs[0])
    torch.cuda.synchronize()
    assert calls == [1]
    assert not torch.equal(before, W[k0])
