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


def _run(name, training_dropout=False):
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
