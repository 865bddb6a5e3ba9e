"""Checkpoint compatibility with the reference (train.py:246-251,289,298; lightning/diff3d.py:40-45,107-108): the file is
{'model': state_dict (optionally `module.`-prefixed), 'optim': torch.optim.Adam state_dict, 'step'[, 'epoch']}.
CPU part: layout conversions in both directions.  GPU part: a run resumed here continues exactly like torch's Adam."""
import copy

import pytest
import torch

from distributed_3d_diffusion_pytorch_b200.diff3d import Diff3D, FusedAdam, load_checkpoint, save_checkpoint
from distributed_3d_diffusion_pytorch_b200.xunet import XUNet

KW = dict(H=8, W=8, ch=32)


def _torch_adam_run(params, grads_per_step, lr=1e-3):
    opt = torch.optim.Adam(params, lr=lr, betas=(0.9, 0.99))
    for grads in grads_per_step:
        for p, g in zip(params, grads):
            p.grad = g.clone()
        opt.step()
    return opt


def _random_grads(model, n_steps, seed):
    g = torch.Generator().manual_seed(seed)
    return [[torch.randn(p.shape, generator=g) * 0.1 for p in model._plist] for _ in range(n_steps)]


def test_torch_adam_state_dict_round_trip_cpu(tmp_path):
    torch.manual_seed(0)
    model = XUNet(**KW)
    ref_params = [torch.nn.Parameter(p.detach().clone()) for p in model._plist]
    grads = _random_grads(model, 2, seed=1)
    topt = _torch_adam_run(ref_params, grads)
    # reference-format file (saved from DataParallel: `module.` prefix), as train.py:289 writes it
    path = str(tmp_path / "latest.pt")
    torch.save({"optim": topt.state_dict(), "model": {"module." + n: p.detach() for n, p in zip(model._names, ref_params)}, "step": 2,
                "epoch": 0}, path)
    ours = XUNet(**KW)
    fopt = FusedAdam(ours, lr=123.0)
    assert load_checkpoint(path, ours, fopt) == 2
    assert fopt.step_count == 2 and fopt.param_groups[0]["lr"] == 1e-3 and tuple(fopt.param_groups[0]["betas"]) == (0.9, 0.99)
    for i, (n, p, o) in enumerate(zip(ours._names, ours._plist, ours._offs)):
        assert torch.equal(p.detach(), ref_params[i].detach()), n
        st = topt.state_dict()["state"][i]
        assert torch.equal(fopt.m[o:o + p.numel()].view(p.shape), st["exp_avg"]), n
        assert torch.equal(fopt.v[o:o + p.numel()].view(p.shape), st["exp_avg_sq"]), n
    # and back: what we save loads into torch's Adam and into the reference's loader unchanged
    path2 = str(tmp_path / "ours.pt")
    ck = save_checkpoint(path2, ours, fopt, step=2, epoch=0, module_prefix=True)
    assert set(ck.keys()) == {"optim", "model", "step", "epoch"} and all(k.startswith("module.") for k in ck["model"])
    ck = torch.load(path2, weights_only=False)
    topt2 = torch.optim.Adam([torch.nn.Parameter(p.detach().clone()) for p in ours._plist], lr=5.0, betas=(0.5, 0.5))
    topt2.load_state_dict(ck["optim"])
    sd2 = topt2.state_dict()
    assert sd2["param_groups"][0]["lr"] == 1e-3 and tuple(sd2["param_groups"][0]["betas"]) == (0.9, 0.99)
    for i in range(len(ours._plist)):
        assert torch.equal(sd2["state"][i]["exp_avg"], topt.state_dict()["state"][i]["exp_avg"])
        assert float(sd2["state"][i]["step"]) == 2.0
    # a fresh optimizer (no step taken) saves an empty state, like torch
    assert FusedAdam(XUNet(**KW)).state_dict()["state"] == {}


def test_unknown_optimizer_state_is_an_error_not_ignored():
    model = XUNet(**KW)
    opt = FusedAdam(model)
    with pytest.raises(ValueError):
        opt.load_state_dict({"momentum_buffer": 1})
    with pytest.raises(ValueError):        # SGD-like / wrong parameter count
        opt.load_state_dict({"state": {}, "param_groups": [{"lr": 1e-3, "params": [0, 1, 2]}]})
    # the old flat layout still loads
    opt.load_state_dict({"step": 4, "m": torch.ones_like(model.flat_parameters()), "v": torch.ones_like(model.flat_parameters()), "lr": 2e-4})
    assert opt.step_count == 4 and opt.m[0] == 1
    # Diff3D.configure_optimizers hands ckpt['optim'] to the optimizer (lightning/diff3d.py:107-108) and no longer drops it
    d = Diff3D.__new__(Diff3D)
    torch.nn.Module.__init__(d)
    d.xunet_denoiser, d.lr, d.use_scheduler, d.pretrained_optim = model, 1e-4, False, {"bogus": 1}
    with pytest.raises(ValueError):
        d.configure_optimizers()


@pytest.mark.gpu
def test_resumed_run_continues_like_torch_adam():
    """two torch-Adam steps on CPU -> checkpoint -> load here -> the third step (fused kernel) equals torch's third step"""
    dev = "cuda:0"
    torch.manual_seed(0)
    model = XUNet(**KW)
    ref_params = [torch.nn.Parameter(p.detach().clone()) for p in model._plist]
    grads = _random_grads(model, 3, seed=2)
    topt = _torch_adam_run(ref_params, grads[:2])
    sd_opt = copy.deepcopy(topt.state_dict())        # (state_dict() hands out the live moment tensors)
    sd_model = {n: p.detach().clone() for n, p in zip(model._names, ref_params)}
    _torch_adam_run_more = topt
    for p, g in zip(ref_params, grads[2]):
        p.grad = g.clone()
    _torch_adam_run_more.step()
    ours = XUNet(**KW).to(dev)
    ours.load_state_dict(sd_model)
    fopt = FusedAdam(ours, ema_decay=0.9)
    fopt.load_state_dict(sd_opt)
    before = ours.flat_parameters().clone()
    gv = ours._grad_views()
    for n, g in zip(ours._names, grads[2]):
        gv[n].copy_(g)
    fopt.step()
    torch.cuda.synchronize()
    for n, p, r in zip(ours._names, ours._plist, ref_params):
        assert (p.detach().cpu() - r.detach()).abs().max() <= 1e-6, n
    # EMA kept by the same kernel: ema = 0.9 * before + 0.1 * after
    assert torch.allclose(fopt.ema, 0.9 * before + 0.1 * ours.flat_parameters(), atol=1e-7)
    e = fopt.ema_state_dict()
    assert set(e.keys()) == set(ours._names) and e["conv.weight"].shape == ours._plist[ours._names.index("conv.weight")].shape
