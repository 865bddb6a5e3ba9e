"""CPU: the tensor-level op semantics (tests/emu_ops.py + gemm.py descriptor builders + weight packers)
against torch's own conv2d / autograd.  This pins the meaning of the C-ABI descriptors that the CUDA
kernels are then tested against on the GPU."""
import math

import pytest
import torch
import torch.nn.functional as F

from distributed_3d_diffusion_pytorch_b200 import gemm as G
from emu_ops import EmuOps

emu = EmuOps()


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("stride,H,Ci,Co", [(1, 8, 24, 40), (2, 8, 144, 32), (4, 16, 16, 8), (8, 16, 8, 16), (1, 6, 8, 3)])
def test_conv_fwd_dgrad_wgrad(stride, H, Ci, Co):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, Ci, H, H, generator=g, requires_grad=True)
    w = torch.randn(Co, Ci, 3, 3, generator=g, requires_grad=True)
    b = torch.randn(Co, generator=g)
    y = F.conv2d(x, w, b, stride=stride, padding=1)
    dy = torch.randn(y.shape, generator=g)
    y.backward(dy)
    Ho = y.shape[-1]
    Cpad = G.ceil64(Ci)
    wp = torch.zeros(Co, 9 * Cpad)
    emu.pack_weight(w.detach(), 0, wp, Cpad=Cpad)
    out = torch.zeros(2, Ho, Ho, Co)
    emu.gemm(G.conv_desc([nhwc(x.detach())], [9], wp, Co, out, stride, bias=b))
    assert torch.allclose(out, nhwc(y.detach()), atol=1e-4)
    # wgrad
    gp = torch.zeros(Co, 9 * Cpad)
    emu.gemm(G.wgrad_desc(nhwc(dy), nhwc(x.detach()), 9, gp, 0, Cpad, stride))
    dw = torch.zeros(Co, Ci, 3, 3)
    emu.unpack_wgrad(gp, 0, Cpad, dw)
    assert torch.allclose(dw, w.grad, atol=1e-3)
    if stride == 1:
        Copad = G.ceil64(Co)
        wt = torch.zeros(Ci, 9 * Copad)
        emu.pack_weight(w.detach(), 1, wt, Cpad=Copad)
        dx = torch.zeros(2, H, H, Ci)
        emu.gemm(G.conv_desc([nhwc(dy)], [9], wt, Ci, dx, 1))
        assert torch.allclose(dx, nhwc(x.grad), atol=1e-4)
    else:
        # strided adjoint through per-tap products + scatter (sum over the batch of 2 "frames")
        wr = torch.zeros(9 * Ci, Co)
        emu.pack_weight(w.detach(), 2, wr)
        Gm = torch.zeros(2 * Ho * Ho, 9 * Ci)
        emu.gemm(G.dense_desc(nhwc(dy).reshape(-1, Co), wr, Gm, 9 * Ci))
        dP = torch.zeros(2, H, H, Ci)
        emu.conv_tap_scatter(Gm, Ci, Ho, Ho, stride, dP, False)
        assert torch.allclose(dP, nhwc(x.grad), atol=1e-4)


def test_concat_sources_and_1x1():
    g = torch.Generator().manual_seed(1)
    xa, xb = torch.randn(2, 8, 8, 40, generator=g), torch.randn(2, 8, 8, 24, generator=g)
    w = torch.randn(16, 64, 1, 1, generator=g)
    ref = F.conv2d(torch.cat([xa, xb], -1).permute(0, 3, 1, 2), w)
    wp = torch.zeros(16, 128)
    emu.pack_weight(w, 0, wp, ci0=0, cn=40, Cpad=64, k_off=0)
    emu.pack_weight(w, 0, wp, ci0=40, cn=24, Cpad=64, k_off=64)
    out = torch.zeros(2, 8, 8, 16)
    emu.gemm(G.conv_desc([xa, xb], [1, 1], wp, 16, out))
    assert torch.allclose(out, nhwc(ref), atol=1e-4)


@pytest.mark.parametrize("cross", [False, True])
def test_attention_matmuls_match_mha(cross):
    g = torch.Generator().manual_seed(2)
    N, Hh, L, d = 4, 2, 12, 8
    C = Hh * d
    qkv = torch.randn(N, L, 3 * C, generator=g, requires_grad=True)
    sp = lambda t, i: t.view(N, L, 3, Hh, d)[:, :, i].permute(0, 2, 1, 3)
    q, k, v = sp(qkv, 0), sp(qkv, 1), sp(qkv, 2)
    idx = torch.arange(N) ^ int(cross)
    S = (q @ k[idx].transpose(-1, -2)) / math.sqrt(d)
    P = torch.softmax(S, -1)
    O = (P @ v[idx]).permute(0, 2, 1, 3).reshape(N, L, C)
    dO = torch.randn(N, L, C, generator=g)
    O.backward(dO)
    qd = qkv.detach()
    q, k, v = sp(qd, 0), sp(qd, 1), sp(qd, 2)
    Lp = 16
    s = torch.zeros(N, Hh, L, Lp)
    emu.gemm(G.attn_qk_desc(q, k, s[..., :L], cross, 1 / math.sqrt(d)))
    assert torch.allclose(s[..., :L], S.detach(), atol=1e-5)
    p = torch.zeros(N, Hh, L, Lp)
    emu.softmax(s[..., :L], p[..., :L])
    o = torch.zeros(N, L, C)
    ov = o.view(N, L, Hh, d).permute(0, 2, 1, 3)
    emu.gemm(G.attn_pv_desc(p[..., :L], v, ov, cross))
    assert torch.allclose(o, O.detach(), atol=1e-5)
    dov = dO.view(N, L, Hh, d).permute(0, 2, 1, 3)
    dp = torch.zeros(N, Hh, L, Lp)
    emu.gemm(G.attn_dp_desc(dov, v, dp[..., :L], cross))
    ds = torch.zeros(N, Hh, L, Lp)
    emu.softmax_bwd(p[..., :L], dp[..., :L], ds[..., :L])
    dqkv = torch.zeros(N, L, 3 * C)
    emu.gemm(G.attn_dv_desc(p[..., :L], dov, sp(dqkv, 2), cross))
    emu.gemm(G.attn_dq_desc(ds[..., :L], k, sp(dqkv, 0), cross, 1 / math.sqrt(d)))
    emu.gemm(G.attn_dk_desc(ds[..., :L], q, sp(dqkv, 1), cross, 1 / math.sqrt(d)))
    assert torch.allclose(dqkv, qkv.grad, atol=1e-5)


def test_groupnorm_matches_torch():
    g = torch.Generator().manual_seed(3)
    N, HW, C0, C1 = 2, 16, 64, 32
    x0 = torch.randn(N, HW, C0, generator=g, requires_grad=True)
    x1 = torch.randn(N, HW, C1, generator=g, requires_grad=True)
    gam = (1 + 0.1 * torch.randn(96, generator=g)).requires_grad_(True)
    bet = (0.1 * torch.randn(96, generator=g)).requires_grad_(True)
    ss = torch.randn(N * HW, 192, generator=g, requires_grad=True)
    x = torch.cat([x0, x1], -1)
    y = F.group_norm(x.permute(0, 2, 1), 32, gam, bet, 1e-5).permute(0, 2, 1)
    out_ref = y * (1 + ss[:, :96].reshape(N, HW, 96)) + ss[:, 96:].reshape(N, HW, 96)
    dy = torch.randn(N, HW, 96, generator=g)
    out_ref.backward(dy)
    sums = torch.zeros(N, 32, 2)
    emu.gn_stats(x0.detach(), x1.detach(), sums)
    out = torch.zeros(N, HW, 96)
    emu.gn_apply(x0.detach(), x1.detach(), sums, gam.detach(), bet.detach(), 2, out, ss=ss.detach(), F=96)
    assert torch.allclose(out, out_ref.detach(), atol=1e-5)
    dg, db, ws = torch.zeros(96), torch.zeros(96), torch.zeros(N, 32, 2)
    dx0, dx1, dss = torch.zeros(N, HW, C0), torch.zeros(N, HW, C1), torch.zeros(N * HW, 192)
    emu.gn_bwd(x0.detach(), x1.detach(), sums, gam.detach(), bet.detach(), 2, dy, dg, db, ws, dx0, dx1, ss=ss.detach(), F=96, dss=dss)
    assert torch.allclose(dx0, x0.grad, atol=1e-4) and torch.allclose(dx1, x1.grad, atol=1e-4)
    assert torch.allclose(dg, gam.grad, atol=1e-4) and torch.allclose(db, bet.grad, atol=1e-4)
    assert torch.allclose(dss, ss.grad, atol=1e-5)
