"""-m gpu: every CUDA kernel against the plain-torch fp32 statement of the same op (tests/emu_ops.py),
called through the C ABI (ops.CudaOps -> libb3d.so)."""
import pytest
import torch

import gemm_cases
import op_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["fp32", "fp32_stats_of_stored_bf16"])
def backends(request):
    """Every case runs against TWO checkers: the pure fp32 statement of the op, and the one whose epilogue GroupNorm
    statistics are taken over the bf16-ROUNDED output (which is what the kernel is specified to sum: the statistics of the
    tensor it stores).  Both are held to the same tolerances."""
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    from emu_ops import EmuOps
    return CudaOps("cuda:0"), EmuOps("cpu", stats_dtype=torch.bfloat16 if request.param != "fp32" else None)


def _cmp(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert torch.isfinite(a).all()
    return (a - b).norm().item() / (b.norm().item() + 1e-12)


@pytest.mark.parametrize("case", gemm_cases.all_cases(), ids=lambda c: c.name)
def test_gemm(case, backends):
    cuda, emu = backends
    d_c, outs_c = case.build(cuda)
    d_e, outs_e = case.build(emu)
    cuda.gemm(d_c)
    torch.cuda.synchronize()
    emu.gemm(d_e)
    for a, b in zip(outs_c, outs_e):
        tol = 8e-3 if a.dtype == torch.bfloat16 else 2e-4
        if "stats" in case.name and a.dtype == torch.float32 and emu.stats_dtype is None:
            tol = 1e-3       # the kernel sums the bf16 values it STORES; the pure-fp32 checker sums unrounded ones (measured 2.4-3.3e-4)
        assert _cmp(a, b) <= tol, (case.name, _cmp(a, b))


@pytest.mark.parametrize("name", list(op_cases.all_cases().keys()))
def test_op(name, backends):
    cuda, emu = backends
    fn = op_cases.all_cases()[name]
    oc = fn(cuda)
    torch.cuda.synchronize()
    oe = fn(emu)
    for i, (a, b) in enumerate(zip(oc, oe)):
        tol = 8e-3 if a.dtype == torch.bfloat16 else 2e-4
        if name.startswith("pose_fp32"):
            tol = 1e-5           # north_star: ray / posenc outputs within 1e-5 relative
        assert _cmp(a, b) <= tol, (name, i)


def test_gn_image_chunked_launches_equal_unchunked():
    """b3d_gn_stats_apply / b3d_gn_bwd run their two passes over a chunk of images at a time when a chunk budget is set (so
    the second pass reads L2, not HBM): same results as one launch pair over all images"""
    from distributed_3d_diffusion_pytorch_b200 import _lib
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    ops = CudaOps("cuda:0")
    g = torch.Generator().manual_seed(5)
    N, HW, C = 10, 1024, 256
    mk = lambda *shape, dt=torch.bfloat16: torch.randn(*shape, generator=g).to("cuda:0", dt)
    x, dy = mk(N, HW, C), mk(N, HW, C)
    ss = mk(N * HW, 2 * C) * 0.5
    gamma, beta = 1 + 0.1 * mk(C, dt=torch.float32), 0.1 * mk(C, dt=torch.float32)
    res = {}
    _lib.set_gn_fuse_mb(0)    # (small tensors would otherwise take the cluster-fused single launch)
    for mb in (0, 1):         # 1 MB budget -> two images (1 MB of x) per chunk
        _lib.set_gn_chunk_mb(mb)
        try:
            for mode in (1, 2):
                sums = torch.zeros(N, 32, 2, device="cuda:0")
                out = torch.zeros(N, HW, C, device="cuda:0", dtype=torch.bfloat16)
                n0 = _lib.launch_count()
                ops.gn_stats_apply(x, None, sums, gamma, beta, mode, out, ss=ss if mode == 2 else None, F=C if mode == 2 else 0, drop_p=0.1 if mode == 2 else 0.0, seed=3)
                launches = _lib.launch_count() - n0
                dgamma, dbeta, ws = torch.zeros(C, device="cuda:0"), torch.zeros(C, device="cuda:0"), torch.zeros(N, 32, 2, device="cuda:0")
                dx = torch.zeros_like(out)
                dss = torch.zeros_like(ss) if mode == 2 else None
                ops.gn_bwd(x, None, sums, gamma, beta, mode, dy, dgamma, dbeta, ws, dx, ss=ss if mode == 2 else None, F=C if mode == 2 else 0, dss=dss,
                           drop_p=0.1 if mode == 2 else 0.0, seed=3)
                torch.cuda.synchronize()
                res[(mb, mode)] = (launches, sums, out, dx, dss, ws, dgamma, dbeta)
        finally:
            _lib.set_gn_chunk_mb(0)
    _lib.set_gn_fuse_mb(40)
    for mode in (1, 2):
        a, b = res[(0, mode)], res[(1, mode)]
        assert a[0] == 2 and b[0] > 2                       # really chunked
        # the same arithmetic with the fp32 partial sums combined in another order (blocks per image differ with the chunk)
        for i in (1, 5, 6, 7):
            assert torch.allclose(a[i], b[i], rtol=1e-4, atol=1e-3), (mode, i)
        for i in (2, 3, 4):
            if a[i] is not None:
                assert _cmp(b[i], a[i]) <= 1e-3, (mode, i)


@pytest.mark.parametrize("N,HW,C0,C1", [(4, 4096, 128, 0), (6, 1024, 256, 128), (3, 64, 512, 512), (2, 36, 64, 32), (5, 16, 32, 0)])
def test_gn_cluster_fused_launch_equals_two_launches(N, HW, C0, C1):
    """small tensors: b3d_gn_stats_apply / b3d_gn_bwd as ONE launch (thread-block cluster per image, partial sums exchanged
    through distributed shared memory) against the two-launch path, all modes, with and without a second source"""
    from distributed_3d_diffusion_pytorch_b200 import _lib
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    ops, dev = CudaOps("cuda:0"), "cuda:0"
    g = torch.Generator().manual_seed(N * HW + C0)
    C = C0 + C1
    mk = lambda *shape, dt=torch.bfloat16: torch.randn(*shape, generator=g).to(dev, dt)
    x0, x1 = mk(N, HW, C0) + 0.3, (mk(N, HW, C1) * 2 - 0.5 if C1 else None)
    dy, add0, add1a = mk(N, HW, C), mk(N, HW, C), mk(N, HW, C0)
    ss = mk(N * HW, 2 * C) * 0.5
    gamma, beta = 1 + 0.1 * mk(C, dt=torch.float32), 0.1 * mk(C, dt=torch.float32)
    res = {}
    try:
        for fuse in (0, 40):
            _lib.set_gn_fuse_mb(fuse)
            _lib.set_gn_fuse_bwd_mb(fuse)
            for mode in (0, 1, 2):
                film = dict(ss=ss, F=C, drop_p=0.1, seed=5) if mode == 2 else {}
                sums = torch.zeros(N, 32, 2, device=dev)
                out = torch.zeros(N, HW, C, device=dev, dtype=torch.bfloat16)
                n0 = _lib.launch_count()
                ops.gn_stats_apply(x0, x1, sums, gamma, beta, mode, out, **film)
                l_fwd = _lib.launch_count() - n0
                dgamma, dbeta, ws = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(N, 32, 2, device=dev)
                dx0 = torch.zeros(N, HW, C0, device=dev, dtype=torch.bfloat16)
                dx1 = torch.zeros(N, HW, C1, device=dev, dtype=torch.bfloat16) if C1 else None
                dss = torch.zeros_like(ss) if mode == 2 else None
                dss_sum = torch.zeros(2 * C, device=dev) if mode == 2 else None
                dx_sum = torch.zeros(C, device=dev)
                n0 = _lib.launch_count()
                ops.gn_bwd(x0, x1, sums, gamma, beta, mode, dy, dgamma, dbeta, ws, dx0, dx1, dss=dss, add0=add0, add1a=add1a, add0_scale=0.5,
                           dss_sum=dss_sum, dx_sum=dx_sum, **film)
                l_bwd = _lib.launch_count() - n0
                torch.cuda.synchronize()
                res[(fuse, mode)] = (l_fwd, l_bwd, [sums, out, dx0, dx1, dss, dgamma, dbeta, dss_sum, dx_sum])
    finally:
        _lib.set_gn_fuse_mb(40)
        _lib.set_gn_fuse_bwd_mb(0)
    for mode in (0, 1, 2):
        (lf0, lb0, a), (lf1, lb1, b) = res[(0, mode)], res[(40, mode)]
        assert (lf0, lb0) == (2, 2) and (lf1, lb1) == (1, 1)
        for i, (u, v) in enumerate(zip(a, b)):
            if u is not None:
                tol = 2e-3 if u.dtype == torch.bfloat16 else 2e-4        # same arithmetic, partial sums combined in another order
                assert _cmp(v, u) <= tol, (mode, i, _cmp(v, u))


@pytest.mark.parametrize("idx", [0, 2, 3, 4, 5, 7, 9, 10, 11])
def test_conv_halo_patch_form_equals_tile_load_form(idx):
    """stride-1 3x3 convolutions: the halo-patch form (one input patch per 64-channel block in shared memory, nine taps through
    shifted UMMA descriptors; packed rows = mode 2, 2048-byte row pitch = mode 1) against nine shifted tile loads (mode 0) on the
    same descriptor -- single CTA, two accumulators, CTA pairs, several sources with 9 and 1 taps, partial channel blocks"""
    import check_halo
    from distributed_3d_diffusion_pytorch_b200 import _lib
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    ops = CudaOps("cuda:0")
    name, case = check_halo.CASES[idx]
    d, outs = case.build(ops)
    res = {}
    try:
        for mode in (0, 1, 2):
            _lib.set_conv_halo(mode)
            for o in outs:
                o.zero_()
            ops.gemm(d)
            torch.cuda.synchronize()
            res[mode] = [o.clone() for o in outs]
    finally:
        _lib.set_conv_halo(2)
    for mode in (1, 2):
        for a, b in zip(res[mode], res[0]):
            assert _cmp(a, b) <= 1e-3, (name, mode, _cmp(a, b))        # same products, summed channel block by channel block instead of tap by tap
