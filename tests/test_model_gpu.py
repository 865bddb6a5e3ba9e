"""-m gpu: the full hot path (XUNet forward + backward, loss, sampler step, training step) on the
CUDA kernels against the oracle, at the tolerances north_star states:
XUNet output <= 2e-2 relative L2 (bf16), loss <= 1e-3 relative, gradient cosine >= 0.999, ray/posenc <= 1e-5."""
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu

OUT_TOL, LOSS_TOL, COS_TOL = 2e-2, 1e-3, 0.999


def _model(kw, seed, dev="cuda:0"):
    from distributed_3d_diffusion_pytorch_b200.xunet import XUNet
    from oracle import xunet_ref as xr
    m = XUNet(**kw).to(dev)
    m.load_state_dict(xr.init_state_dict(xr.Cfg(**kw), seed=seed))
    return m


def _fixture_run(name):
    from distributed_3d_diffusion_pytorch_b200.diff3d import diffusion_loss, q_sample
    G = load_golden(name)
    dev = "cuda:0"
    m = _model(G["kw"], G["seed"]).eval()
    img, cm = G["img"].to(dev), G["cond_mask"].to(dev)
    z_noisy = q_sample(img[:, 1], G["logsnr"].to(dev), G["noise"].to(dev))
    assert torch.allclose(z_noisy.cpu(), G["z_noisy"], atol=2e-6)
    x_cond = torch.where(cm[:, None, None, None], img[:, 0], G["x_replace"].to(dev))
    batch = {"x": x_cond, "z": z_noisy, "logsnr": G["batch_logsnr"].to(dev), "R": G["R"].to(dev), "t": G["T"].to(dev), "K": G["K"].to(dev)}
    eps = m(batch, cond_mask=cm)
    rel = ((eps.cpu() - G["eps"]).norm() / G["eps"].norm()).item()
    loss = diffusion_loss(G["noise"].to(dev), eps)
    loss_rel = abs(loss.item() - G["loss"].item()) / abs(G["loss"].item())
    loss.backward()
    grads = dict((n, p.grad.float().cpu()) for n, p in m.named_parameters())
    assert all(torch.isfinite(g).all() for g in grads.values())
    return G, rel, loss_rel, grads, [n for n, _ in m.named_parameters()]


def _cos(a, b):
    return torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()


@pytest.mark.parametrize("name", ["small16", "small24", "small32"])
def test_golden_fixture_parity(name):
    """against fixtures produced by the reference's own code, at north_star's bars: output 2e-2, loss 1e-3, gradient
    cosine 0.999 (on a seeded 262 144-element sample of the whole flat gradient stored in the fixture), plus every stored
    gradient tensor individually and the norm of the whole gradient"""
    G, rel, loss_rel, grads, names = _fixture_run(name)
    assert rel <= OUT_TOL, rel
    assert loss_rel <= LOSS_TOL, loss_rel
    flat = torch.cat([grads[n].flatten() for n in names])
    assert _cos(flat[G["grad_sample_idx"].long()], G["grad_sample"]) >= COS_TOL
    assert abs(flat.double().norm().item() - G["grad_total_norm"]) <= 2e-2 * G["grad_total_norm"]
    for k, g in G["grads"].items():
        if g.norm() < 1e-5:
            continue
        # single small tensors (biases, GroupNorm affine vectors of the 2x2-pixel level) in bf16; the north_star bar is the
        # cosine of the whole gradient, asserted above
        assert _cos(grads[k], g) >= 0.99, (k, _cos(grads[k], g))


@pytest.mark.parametrize("name", ["tiny16", "tiny24"])
def test_degenerate_fixture_structure_only_loosened_bars(name):
    """NOT a north_star-tolerance test.  ch=32 puts ONE channel in each GroupNorm group and the deepest level is 2x2 (3x3)
    pixels: statistics over 4-9 values amplify bf16 rounding (the same algorithm in torch with bf16 buffers gives 2.5e-2 /
    gradient cosine 0.79 on the worst tensor).  These files are the exact fp32 orchestration fixtures of
    tests/test_engine_cpu.py; on the GPU they only check that nothing is structurally wrong (3x looser output / loss bars,
    the first gradient computed in the backward pass).  The strict bars are asserted on small16 / small24 / small32 above."""
    G, rel, loss_rel, grads, _ = _fixture_run(name)
    assert rel <= 3 * OUT_TOL, rel
    assert loss_rel <= 3 * LOSS_TOL, loss_rel
    assert _cos(grads["lastconv.weight"], G["grads"]["lastconv.weight"]) >= 0.99


def test_pose_embedding_matches_reference_fixture():
    """K1 through the C ABI, fp32 output, against the reference-generated pose_emb (<= 1e-5 relative)"""
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    from oracle import xunet_ref as xr
    for name in ("tiny16", "tiny24"):
        G = load_golden(name)
        sd = xr.init_state_dict(xr.Cfg(**G["kw"]), seed=G["seed"])
        ops, dev = CudaOps("cuda:0"), "cuda:0"
        B, H = G["img"].shape[0], G["kw"]["H"]
        out = torch.zeros(2 * B, H, H, 144, device=dev)
        cp = "conditioningprocessor."
        ops.pose_embedding(G["K"].to(dev), G["R"].to(dev), G["T"].to(dev), G["cond_mask"].to(dev), sd[cp + "pos_emb"].to(dev),
                           sd[cp + "first_emb"].flatten().to(dev), sd[cp + "other_emb"].flatten().to(dev), out)
        ref = G["pose_emb"].reshape(2 * B, 144, H, H).permute(0, 2, 3, 1)
        rel = ((out.cpu() - ref).norm() / ref.norm()).item()
        assert rel <= 1e-5, rel
        assert (out.cpu() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item() + 2e-6


@pytest.mark.parametrize("kw,B,cm", [
    (dict(H=32, W=32, ch=64), 3, [False, True, True]),
    (dict(H=64, W=64, ch=128), 2, [True, False]),          # CFG-64, the configuration of the headline metric
    (dict(H=128, W=128, ch=256), 1, [True]),               # CFG-128, the paper's configuration (head_dim 256 at the deepest level)
])
def test_forward_backward_parity_vs_oracle(kw, B, cm):
    from distributed_3d_diffusion_pytorch_b200.selftest import parity_metrics
    m = parity_metrics(kw, B, seed=3, cond_mask=cm)
    assert m["finite"]
    print("parity", kw, m)
    assert m["out_rel_l2"] <= OUT_TOL, m
    assert m["loss_rel"] <= LOSS_TOL, m
    assert m["grad_cos"] >= COS_TOL, m
    assert m["worst_grad_cos"] >= 0.98, m


def test_sampler_step_matches_oracle():
    from distributed_3d_diffusion_pytorch_b200 import sampling, synthetic
    from oracle import diffusion_ref as dr, xunet_ref as xr
    dev = "cuda:0"
    kw = dict(H=16, W=16, ch=32)
    m = _model(kw, 4).eval()
    ocfg = xr.Cfg(**kw)
    sd = xr.init_state_dict(ocfg, seed=4)
    b = 8
    img, R, T, K = synthetic.make_views(1, 16, seed=9)
    g = torch.Generator().manual_seed(1)
    z, xu, noise = (torch.randn(b, 3, 16, 16, generator=g) for _ in range(3))
    x = img[0, 0][None].repeat(b, 1, 1, 1)
    Rb, Tb = R.repeat(b, 1, 1, 1), T.repeat(b, 1, 1)
    w = torch.arange(8)
    table = dr.logsnr_schedule_cosine(torch.linspace(1.0, 0.0, 257))
    for step in (3, 127, 255):
        l0, l1 = table[step], table[step + 1]
        ec, eu = sampling.predict_noise_pair(m, x, z, Rb, Tb, K[0], l0, x_uncond=xu)
        bo = dr.make_batch(x, l0.repeat(b), z, Rb, Tb, K.repeat(b, 1, 1))
        eco = xr.xunet_forward(sd, ocfg, bo, torch.ones(b, dtype=torch.bool))
        bo["x"] = xu
        euo = xr.xunet_forward(sd, ocfg, bo, torch.zeros(b, dtype=torch.bool))
        assert ((ec.cpu() - eco).norm() / eco.norm()).item() <= 2 * OUT_TOL
        assert ((eu.cpu() - euo).norm() / euo.norm()).item() <= 2 * OUT_TOL
        ref, _, _ = dr.sampler_update(z, ec.cpu(), eu.cpu(), w, l0, l1, noise)
        out = torch.empty(b, 3, 16, 16, device=dev)
        from distributed_3d_diffusion_pytorch_b200.diff3d import _ops
        _ops(torch.device(dev)).sampler_update(z.to(dev), ec.contiguous(), eu.contiguous(), w.float().to(dev), noise.to(dev), out, float(l0), float(l1))
        assert torch.allclose(out.cpu(), ref, atol=2e-5, rtol=1e-5), step
    # and the public loop runs end to end (4 steps) and is finite
    rec = [[x.numpy(), R[0, 0].numpy(), T[0, 0].numpy()]]
    o = sampling.sample(m, rec, R[0, 1], T[0, 1], K[0], w, timesteps=4)
    assert o.shape == (8, 3, 16, 16) and bool((o == o).all())


def test_graph_sampler_matches_functional_step():
    """the captured (CUDA graph) denoise step == the functional p_sample on the same draws"""
    from distributed_3d_diffusion_pytorch_b200 import sampling, synthetic
    from distributed_3d_diffusion_pytorch_b200.diff3d import logsnr_schedule_cosine
    dev = "cuda:0"
    m = _model(dict(H=16, W=16, ch=64), 6).eval()
    b = 8
    img, R, T, K = synthetic.make_views(1, 16, seed=2)
    table = logsnr_schedule_cosine(torch.linspace(1.0, 0.0, 257))
    w = torch.arange(8, dtype=torch.float32)
    cond = img[0, 0][None].repeat(b, 1, 1, 1)
    for use_graph in (False, True):
        gs = sampling.GraphSampler(m, b, use_graph=use_graph)
        gs.w.copy_(w)
        gs.set_views(cond, R[0, 0], T[0, 0], R[0, 1], T[0, 1], K[0])
        z0 = torch.randn(b, 3, 16, 16, generator=torch.Generator().manual_seed(3)).to(dev)
        gs.z.copy_(z0)
        z_ref = z0.clone()
        for i in (10, 11, 127, 200):          # 127 -> logsnr_next == 0: no noise added (Q11)
            torch.manual_seed(100 + i)
            gs.step(table[i], table[i + 1])
            torch.manual_seed(100 + i)
            xu = torch.empty(b, 3, 16, 16, device=dev).normal_()
            noise = torch.empty(b, 3, 16, 16, device=dev).normal_()
            Rb = torch.stack([R[0, 0], R[0, 1]], 0)[None].repeat(b, 1, 1, 1)
            Tb = torch.stack([T[0, 0], T[0, 1]], 0)[None].repeat(b, 1, 1)
            ec, eu = sampling.predict_noise_pair(m, cond, z_ref, Rb, Tb, K[0], table[i], x_uncond=xu)
            out = torch.empty_like(z_ref)
            from distributed_3d_diffusion_pytorch_b200.diff3d import _ops
            _ops(torch.device(dev)).sampler_update(z_ref, ec.contiguous(), eu.contiguous(), w.to(dev), noise, out, float(table[i]), float(table[i + 1]))
            z_ref = out
            assert torch.allclose(gs.z, z_ref, atol=2e-2, rtol=2e-2), (use_graph, i, (gs.z - z_ref).abs().max())


def test_training_step_with_dropout_and_fused_adam():
    """Diff3D.training_step + backward + fused Adam: runs with dropout on, loss goes down on a fixed batch"""
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    from distributed_3d_diffusion_pytorch_b200.diff3d import Diff3D
    from oracle import xunet_ref as xr
    torch.manual_seed(0)
    d = Diff3D(image_size=16, batch_size=4, n_samples=4, lr=2e-4)
    d.xunet_denoiser = _model(dict(H=16, W=16, ch=128), 2)
    d = d.to("cuda:0") if hasattr(d, "to") else d
    d.train()
    opt = d.optimizers()
    batch = synthetic.make_views(4, 16, seed=5, device="cuda:0")
    losses = []
    for it in range(12):
        torch.manual_seed(123)                    # same noise / logsnr / mask draw every iteration
        loss = d.training_step(batch, it)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert all(l == l for l in losses)
    assert losses[-1] < losses[0], losses


def test_no_cpu_fallback():
    from distributed_3d_diffusion_pytorch_b200 import _lib
    from distributed_3d_diffusion_pytorch_b200.xunet import XUNet
    m = XUNet(H=16, W=16, ch=32)                  # left on the CPU on purpose
    with pytest.raises(_lib.B3DError):
        m({"x": torch.zeros(1, 3, 16, 16), "z": torch.zeros(1, 3, 16, 16), "logsnr": torch.zeros(1, 2), "R": torch.eye(3).repeat(1, 2, 1, 1).double(),
           "t": torch.zeros(1, 2, 3).double(), "K": torch.eye(3)[None].double()}, cond_mask=torch.tensor([True]))


def test_cached_conditioning_matches_uncached_and_follows_parameter_updates():
    """GraphSampler(cache_conditioning=True) -- conditioning convs evaluated once per camera, FiLM for the 4 distinct
    conditionings -- computes the same arithmetic as the uncached sampler: single denoise steps from identical states agree up
    to the order of the fp32 atomics of the GroupNorm statistics, also when the conditioning entry / the target view change
    and the cache is hit; a parameter change between calls is picked up (the packed operands and the cache are keyed on the
    model's parameter version).  Whole trajectories are not compared: see test_glue_gpu.py on why they are chaotic."""
    import random
    from distributed_3d_diffusion_pytorch_b200 import sampling, synthetic
    dev = "cuda:0"
    m = _model(dict(H=32, W=32, ch=64), 6).eval()
    b = 4
    img, R, T, K = synthetic.make_views(3, 32, seed=2)
    w = torch.tensor([0.0, 1.0, 2.0, 3.0])
    rec = [[img[i, 0][None].repeat(b, 1, 1, 1).numpy(), R[i, 0].numpy(), T[i, 0].numpy()] for i in range(2)]
    rel = lambda a, c: float((a - c).norm() / c.norm())
    ga, gb = sampling.GraphSampler(m, b, cache_conditioning=False), sampling.GraphSampler(m, b, cache_conditioning=True)

    def one_step(gs, entry, view, seed, lam=-0.5):
        gs.w.copy_(w.to(dev))
        gs.set_views(rec[entry][0], rec[entry][1], rec[entry][2], R[2, view], T[2, view], K[0], cond_key=entry, target_key=view)
        torch.manual_seed(seed)
        gs.z.normal_()
        gs.step(torch.tensor(lam), torch.tensor(lam + 0.1))
        return gs.z.clone()

    # eager step, captured step, cache misses and hits, both conditioning entries, both target views
    for i, (entry, view) in enumerate([(0, 0), (1, 0), (0, 0), (1, 1), (0, 1), (1, 1)]):
        za, zb = one_step(ga, entry, view, 100 + i), one_step(gb, entry, view, 100 + i)
        assert rel(zb, za) <= 3e-3, (i, rel(zb, za))
    assert len(gb._entries) == 2 and all(e["c"] is not None for e in gb._entries.values())
    before = one_step(gb, 0, 0, 7)
    # parameters change: the next step must use them (not the stale packed operands / cached conditioning)
    with torch.no_grad():
        m.flat_parameters().mul_(1.3)
    m._param_version += 1
    after = one_step(gb, 0, 0, 7)
    fresh = one_step(sampling.GraphSampler(m, b, cache_conditioning=True), 0, 0, 7)
    assert rel(after, fresh) <= 3e-3, rel(after, fresh)
    assert rel(after, before) > 5 * rel(after, fresh) + 1e-3, (rel(after, before), rel(after, fresh))
    # and the public loop runs on the cached sampler over a growing record
    random.seed(5)
    out = gb.sample(rec[:1], R[2, 0], T[2, 0], K[0], w, timesteps=4)
    assert out.shape == (b, 3, 32, 32) and bool((out == out).all())
