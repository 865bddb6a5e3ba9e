"""CPU: the engine's forward/backward ORCHESTRATION (kernel order, descriptors, layouts, gradient
routing) executed with the fp32 torch statement of every kernel (tests/emu_ops.py), against the oracle
(oracle/xunet_ref.py + autograd).  The CUDA kernels themselves are tested by `-m gpu`."""
import pytest
import torch

from conftest import load_golden
from distributed_3d_diffusion_pytorch_b200.engine import Cfg, Engine
from distributed_3d_diffusion_pytorch_b200.xunet import param_shapes
from emu_ops import EmuOps
from oracle import diffusion_ref as dr
from oracle import xunet_ref as xr


def _run(name, training_dropout=False, fuse_stats=False):
    G = load_golden(name)
    kw = G["kw"]
    cfg = Cfg(**kw)
    ocfg = xr.Cfg(**kw)
    assert list(param_shapes(cfg).items()) == list(xr.param_shapes(ocfg).items())
    sd = xr.init_state_dict(ocfg, seed=G["seed"])
    img, cm = G["img"], G["cond_mask"]
    B = img.shape[0]
    x_cond = torch.where(cm[:, None, None, None], img[:, 0], G["x_replace"])
    eng = Engine(cfg, EmuOps(), B)
    eng.fuse_stats = fuse_stats        # GroupNorm statistics from the producing GEMM's epilogue wherever it can take them
    eps = eng.forward(sd, x_cond, G["z_noisy"], G["batch_logsnr"], G["R"], G["T"], G["K"], cm, training=False, need_bwd=True)
    assert (eps - G["eps"]).abs().max() < 5e-5, (eps - G["eps"]).abs().max()
    n = G["noise"].numel()
    deps = 2 * (eps - G["noise"]) / n
    Gr = {k: torch.zeros_like(v) for k, v in sd.items()}
    eng.backward(Gr, deps)
    for k, nrm in G["grad_norms"].items():
        assert abs(Gr[k].norm().item() - nrm) <= 5e-4 * nrm + 2e-6, (k, Gr[k].norm().item(), nrm)
    for k, g in G["grads"].items():
        if g.norm() < 1e-5:
            continue
        cos = torch.nn.functional.cosine_similarity(Gr[k].flatten(), g.flatten(), dim=0)
        assert cos > 0.9999, (k, cos)
    # second step on the same engine: buffers are reused, results must not change
    eps2 = eng.forward(sd, x_cond, G["z_noisy"], G["batch_logsnr"], G["R"], G["T"], G["K"], cm, training=False, need_bwd=True)
    assert torch.equal(eps, eps2)
    Gr2 = {k: torch.zeros_like(v) for k, v in sd.items()}
    eng.backward(Gr2, deps)
    for k in Gr:
        assert torch.allclose(Gr[k], Gr2[k], atol=1e-7), k


def test_engine_orchestration_tiny16():
    _run("tiny16")


def test_engine_orchestration_tiny24():
    _run("tiny24")


def test_engine_orchestration_with_epilogue_statistics():
    from distributed_3d_diffusion_pytorch_b200 import engine as E
    taken, orig = [], E.Engine._gemm_stats

    def spy(self, d, sums, C, rows_per_image=0):
        ok = orig(self, d, sums, C, rows_per_image)
        taken.append(ok)
        return ok
    E.Engine._gemm_stats = spy
    try:
        _run("small16", fuse_stats=True)
    finally:
        E.Engine._gemm_stats = orig
    assert sum(taken) >= 20, (sum(taken), len(taken))          # conv1, conv2 / attention-linear producers with >= 4 channels per group


def test_cached_conditioning_forward_equals_uncached():
    """Sampler-side caching: the conditioning convs are evaluated once for the four distinct (camera, frame role,
    cond/uncond) images, FiLM for four images instead of 2b; per step only silu(c + logsnr_emb) runs.  Same eps as the plain
    forward of the batch-2b sampler step."""
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    kw = dict(H=16, W=16, ch=32)
    cfg, ocfg = Cfg(**kw), xr.Cfg(**kw)
    sd = xr.init_state_dict(ocfg, seed=9)
    b = 3
    img, R, T, K = synthetic.make_views(1, 16, seed=4)
    g = torch.Generator().manual_seed(0)
    z = torch.randn(b, 3, 16, 16, generator=g)
    x = torch.cat([img[:1, 0].repeat(b, 1, 1, 1), torch.randn(b, 3, 16, 16, generator=g)], 0)
    logsnr = torch.stack([torch.full((2 * b,), 20.0), torch.full((2 * b,), -1.25)], 1)
    Rb, Tb, Kb = R.repeat(2 * b, 1, 1, 1), T.repeat(2 * b, 1, 1), K.repeat(2 * b, 1, 1)
    mask = torch.cat([torch.ones(b, dtype=torch.bool), torch.zeros(b, dtype=torch.bool)])
    eng = Engine(cfg, EmuOps(), 2 * b)
    ref = eng.forward(sd, x, torch.cat([z, z]), logsnr, Rb, Tb, Kb, mask, training=False, need_bwd=False).clone()
    # the four distinct conditionings from a 2-pair pass: images 0..3 = (cond f0, target f1, uncond f0, uncond f1)
    engc = Engine(cfg, EmuOps(), 2)
    c = [t.clone() for t in engc.cond_features(sd, R.repeat(2, 1, 1, 1), T.repeat(2, 1, 1), K.repeat(2, 1, 1), torch.tensor([True, False]))]
    image = torch.tensor([2 * int(n >= 2 * b) + (n & 1) for n in range(4 * b)], dtype=torch.int32)
    rep = torch.tensor([0, 1, 2 * b, 2 * b + 1], dtype=torch.int32)
    eng2 = Engine(cfg, EmuOps(), 2 * b)
    junk = torch.full_like(Rb, float("nan"))
    out = eng2.forward(sd, x, torch.cat([z, z]), logsnr, junk, torch.full_like(Tb, float("nan")), Kb, mask, training=False, need_bwd=False,
                       cond=dict(c=c, image=image, rep=rep))
    assert torch.isfinite(out).all()
    assert (out - ref).abs().max() <= 1e-5 * ref.abs().max(), (out - ref).abs().max()
