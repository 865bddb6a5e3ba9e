"""Case table for the HBM-bound kernels: each case returns (callable(ops) -> list of output tensors).
The same callable runs on CudaOps and on EmuOps; outputs are compared by the tests."""
import math

import torch

from distributed_3d_diffusion_pytorch_b200 import synthetic


def _mk(ops, t, dtype=None):
    return t.to(device=ops.device, dtype=dtype or ops.act_dtype).contiguous()


def _r(shape, g, scale=1.0):
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).float()


def pose_case(B, H, W, fp32, seed=0):
    def run(ops):
        g = torch.Generator().manual_seed(seed)
        _, R, T, K = synthetic.make_views(B, H, seed=seed)
        mask = (torch.arange(B) % 3 != 1)
        pos = torch.randn(144, H, W, generator=g) / 12
        fe, oe = torch.randn(144, generator=g) / 12, torch.randn(144, generator=g) / 12
        dev = ops.device
        out = ops.zeros((2 * B, H, W, 144), torch.float32 if fp32 else None)
        ops.pose_embedding(K.to(dev), R.to(dev), T.to(dev), mask.to(dev), pos.to(dev), fe.to(dev), oe.to(dev), out)
        return [out]
    return run


def gn_case(N, HW, C0, C1, mode, drop_p=0.0, seed=0, step=None):
    def run(ops):
        g = torch.Generator().manual_seed(seed)
        # device-side step counter mixed into the dropout seed (what a captured training graph uses)
        sp = _mk(ops, torch.tensor([step]), torch.int64) if step is not None else None
        C = C0 + C1
        x0 = _mk(ops, (_r((N, HW, C0), g) + 0.3).to(torch.bfloat16).float())
        x1 = _mk(ops, (_r((N, HW, C1), g) * 2 - 0.5).to(torch.bfloat16).float()) if C1 else None
        gamma = _mk(ops, 1 + 0.1 * torch.randn(C, generator=g), torch.float32)
        beta = _mk(ops, 0.1 * torch.randn(C, generator=g), torch.float32)
        sums = ops.zeros((N, 32, 2), torch.float32)
        ops.gn_stats(x0, x1, sums)
        F = C
        ssw = ops.zeros((N * HW, 2 * F + 64))
        ssw[:, 32:32 + 2 * F] = _mk(ops, _r((N * HW, 2 * F), g, 0.5))
        ss = ssw[:, 32:32 + 2 * F] if mode == 2 else None
        out = ops.zeros((N, HW, C))
        ops.gn_apply(x0, x1, sums, gamma, beta, mode, out, ss=ss, F=F, drop_p=drop_p, seed=77, seed_ptr=sp)
        dy = _mk(ops, _r((N, HW, C), g))
        dgamma, dbeta = ops.zeros((C,), torch.float32), ops.zeros((C,), torch.float32)
        ws = ops.zeros((N, 32, 2), torch.float32)
        dx0 = ops.zeros((N, HW, C0))
        dx1 = ops.zeros((N, HW, C1)) if C1 else None
        dss = ops.zeros((N * HW, 2 * F)) if mode == 2 else None
        add0 = _mk(ops, _r((N, HW, C), g))
        add1a = _mk(ops, _r((N, HW, C0), g))
        ops.gn_bwd(x0, x1, sums, gamma, beta, mode, dy, dgamma, dbeta, ws, dx0, dx1, ss=ss, F=F, dss=dss, drop_p=drop_p, seed=77,
                   add0=add0, add1a=add1a, seed_ptr=sp, add0_scale=0.7071067811865476 if (seed % 2) else 1.0)
        outs = [sums, out, dgamma, dbeta, ws, dx0]
        if C1:
            outs.append(dx1)
        if mode == 2:
            outs.append(dss)
        return outs
    return run


def misc_case(seed=0):
    def run(ops):
        g = torch.Generator().manual_seed(seed)
        outs = []
        dev = ops.device
        # ddpm embedding
        ls = torch.tensor([-25.0, -3.3, 0.0, 0.7, 12.0, 30.0])
        e = ops.zeros((6, 1024))
        ops.ddpm_embedding(ls.to(dev), e)
        outs.append(e)
        # pack / unpack
        B, H, W = 3, 8, 8
        x, z = torch.randn(B, 3, H, W, generator=g), torch.randn(B, 3, H, W, generator=g)
        pk = ops.zeros((2 * B, H, W, 8))
        ops.pack_input(x.to(dev), z.to(dev), pk)
        outs.append(pk)
        dx, dz = ops.zeros((B, 3, H, W), torch.float32), ops.zeros((B, 3, H, W), torch.float32)
        ops.unpack_input_grad(_mk(ops, _r((2 * B, H, W, 8), g)), dx, dz)
        outs += [dx, dz]
        y = _mk(ops, _r((2 * B, H, W, 16), g))
        eps = ops.zeros((B, 3, H, W), torch.float32)
        ops.unpack_output(y, eps)
        outs.append(eps)
        dy = ops.zeros((2 * B, H, W, 16))
        ops.pack_output_grad(eps, dy)
        outs.append(dy)
        # weight packing
        w = torch.randn(40, 24, 3, 3, generator=g)
        o0 = ops.zeros((40, 64 + 9 * 64))
        ops.pack_weight(w.to(dev), 0, o0, ci0=8, cn=16, Cpad=64, k_off=64)
        o1 = ops.zeros((24, 9 * 64))
        ops.pack_weight(w.to(dev), 1, o1, Cpad=64)
        o2 = ops.zeros((9 * 24, 40))
        ops.pack_weight(w.to(dev), 2, o2)
        outs += [o0, o1, o2]
        gpk = _mk(ops, torch.randn(40, 64 + 9 * 64, generator=g), torch.float32)
        dw = ops.zeros((40, 24, 3, 3), torch.float32)
        ops.unpack_wgrad(gpk, 64, 64, dw, ci0=8, cn=16)
        ops.unpack_wgrad(gpk, 64, 64, dw, ci0=8, cn=16, accumulate=True)
        outs.append(dw)
        # the same conversions through recorded job lists (one launch)
        from distributed_3d_diffusion_pytorch_b200.ops import JobList
        jl = JobList("pack")
        wd = w.to(dev)
        q0, q1, q2 = ops.zeros((40, 64 + 9 * 64)), ops.zeros((24, 9 * 64 + 64)), ops.zeros((9 * 24, 40))
        ops.record_pack_weight(jl, wd, 0, q0, ci0=8, cn=16, Cpad=64, k_off=64)
        ops.record_pack_weight(jl, wd, 1, q1, Cpad=64, k_off=64)
        ops.record_pack_weight(jl, wd, 2, q2)
        bsrc, bdst = torch.randn(50, generator=g).to(dev), ops.zeros((64,), torch.float32)
        ops.record_copy(jl, bsrc, bdst[8:58])
        ops.run_jobs(jl)
        outs += [q0, q1, q2, bdst]
        ju = JobList("unpack")
        dw2 = ops.zeros((40, 24, 3, 3), torch.float32)
        ops.record_unpack_wgrad(ju, gpk, 64, 64, dw2, ci0=8, cn=16)
        ops.record_unpack_wgrad(ju, gpk, 64, 64, dw2, ci0=0, cn=8, accumulate=True)
        ops.run_jobs(ju)
        outs.append(dw2)
        sc_in = _mk(ops, _r((300, 96), g))
        sc_out, sc_sum = ops.zeros((300, 96)), ops.zeros((96,), torch.float32)
        ops.scale_colsum(sc_in, sc_out, 0.70710678, sc_sum)
        outs += [sc_out, sc_sum]
        # resampling
        t = _mk(ops, _r((2, 8, 8, 32), g))
        p = ops.zeros((2, 4, 4, 32)); ops.avgpool2(t, p); outs.append(p)
        u = ops.zeros((2, 16, 16, 32)); ops.upsample2(t, u, 0.5); outs.append(u)
        # elementwise
        a, b = _mk(ops, _r((64, 128), g)), _mk(ops, _r((64, 128), g))
        sb = ops.zeros((64, 128)); ops.silu_bwd(a, b, sb); outs.append(sb)
        # silu backward fused with per-group and total column sums (in place, ragged row count per block)
        fd, fx = _mk(ops, _r((6 * 50, 1024), g)), _mk(ops, _r((6 * 50, 1024), g, 2.0))
        fgo, fto = ops.zeros((6, 1024), torch.float32), ops.zeros((1024,), torch.float32)
        ops.silu_bwd_colsum(fd, fx, fd, 50, fgo, fto)
        outs += [fd, fgo, fto]
        fgo2, fto2 = ops.zeros((6, 1024), torch.float32), ops.zeros((1024,), torch.float32)
        ops.silu_bwd_colsum(fx, None, None, 50, fgo2, fto2)          # pure two-level column sums (of an input: bf16-exact on both sides)
        outs += [fgo2, fto2]
        ad = ops.zeros((64, 128)); ops.add_scale(a, b, ad, 0.7071); outs.append(ad)
        ad2 = ops.zeros((64, 128)); ops.add_scale(a, None, ad2, 0.5); outs.append(ad2)
        cs = ops.zeros((4, 96), torch.float32)
        big = _mk(ops, _r((4 * 300, 128), g))
        ops.colsum(big[:, 16:112], 300, cs); outs.append(cs)
        cs1 = ops.zeros((96,), torch.float32)
        ops.colsum(big[:, 16:112], 1200, cs1, scale=0.7071067811865476); outs.append(cs1)
        bf = _mk(ops, _r((10, 4, 4, 64), g))
        bo = ops.zeros((2, 4, 4, 64)); ops.batch_frame_sum(bf, bo); outs.append(bo)
        for s_, h_ in ((1, 8), (2, 4), (4, 2), (8, 1)):
            Gm = _mk(ops, torch.randn(2 * h_ * h_, 9 * 144 + 16, generator=g), torch.float32)
            dP = ops.zeros((2, 8, 8, 144), torch.float32)
            ops.conv_tap_scatter(Gm, 144, h_, h_, s_, dP, False)
            ops.conv_tap_scatter(Gm, 144, h_, h_, s_, dP, True)
            outs.append(dP)
        dPb = _mk(ops, torch.randn(2, 8, 8, 144, generator=g), torch.float32)
        dpos, dfe, doe = ops.zeros((144, 8, 8), torch.float32), ops.zeros((144,), torch.float32), ops.zeros((144,), torch.float32)
        ops.pose_embedding_bwd(dPb, dpos, dfe, doe)
        outs += [dpos, dfe, doe]
        # softmax (fp32 and bf16 logits, padded row strides)
        for dt in (torch.float32, None):
            S = ops.zeros((2, 4, 36, 40), dt)
            S[..., :36] = _mk(ops, _r((2, 4, 36, 36), g, 2.0), dt)
            P = ops.zeros((2, 4, 36, 40)); ops.softmax(S[..., :36], P[..., :36]); outs.append(P)
            dP2 = ops.zeros((2, 4, 36, 40), dt); dP2[..., :36] = _mk(ops, _r((2, 4, 36, 36), g), dt)
            dS = ops.zeros((2, 4, 36, 48)); ops.softmax_bwd(P[..., :36], dP2[..., :36], dS[..., :36]); outs.append(dS)
        # diffusion glue
        Bq = 5
        zz, nz, xx, xr = [torch.randn(Bq, 3, 8, 8, generator=g).to(dev) for _ in range(4)]
        lsn = torch.tensor([-18.0, -2.0, 0.1, 3.0, 19.5]).to(dev)
        mk = torch.tensor([1, 0, 1, 1, 0], dtype=torch.uint8).to(dev)
        zo, xo = ops.zeros((Bq, 3, 8, 8), torch.float32), ops.zeros((Bq, 3, 8, 8), torch.float32)
        ops.q_sample(zz, nz, lsn, zo, xx, xr, mk, xo)
        outs += [zo, xo]
        scr = ops.loss_scratch()
        for lt in (0, 1, 2):
            lo = ops.zeros((1,), torch.float32)
            dpd = ops.zeros((Bq, 3, 8, 8), torch.float32)
            ops.loss(nz, zz * 1.5, dpd, lt, scr, lo)
            ops.loss(nz, zz * 1.5, dpd, lt, scr, lo)      # second call: ticket must have been reset
            outs += [lo, dpd]
        wv = torch.arange(Bq, dtype=torch.float32).to(dev)
        for l0, l1 in ((-20.0, -19.5), (0.3, 0.0), (0.0, 0.4), (19.0, 20.0)):
            so = ops.zeros((Bq, 3, 8, 8), torch.float32)
            ops.sampler_update(zz, nz, xx, wv, xr, so, l0, l1)
            outs.append(so)
            so2 = ops.zeros((Bq, 3, 8, 8), torch.float32)
            ops.sampler_update_dev(zz, nz, xx, wv, xr, so2, torch.tensor([l0, l1], dtype=torch.float32).to(dev))
            outs.append(so2)
        Wm, vr, vc = torch.randn(70, 300, generator=g).to(dev), torch.randn(70, generator=g).to(dev), torch.randn(300, generator=g).to(dev)
        y0, y1 = ops.zeros((70,), torch.float32), ops.zeros((300,), torch.float32)
        ops.matvec(Wm, vc, y0, add=vr)
        ops.matvec(Wm, vr, y1, transpose=True)
        ops.rank1_add(Wm, vr, vc)
        outs += [y0, y1, Wm]
        pa, ga = torch.randn(1000, generator=g).to(dev), torch.randn(1000, generator=g).to(dev)
        ma, va = ops.zeros((1000,), torch.float32), ops.zeros((1000,), torch.float32)
        for st in (1, 2, 3):
            ops.adam(pa, ga, ma, va, 1e-2, 0.9, 0.99, 1e-8, st)
        outs += [pa, ma, va]
        # the same step from a bf16 gradient buffer (data-parallel trainer: the all-reduced communication buffer, read in place)
        pb, gb16 = torch.randn(1003, generator=g).to(dev), torch.randn(1003, generator=g).to(dev).to(torch.bfloat16)
        mb, vb = ops.zeros((1003,), torch.float32), ops.zeros((1003,), torch.float32)
        for st in (1, 2):
            ops.adam(pb, gb16, mb, vb, 1e-2, 0.9, 0.99, 1e-8, st)
        outs += [pb, mb, vb]
        return outs
    return run


def attn_case(N, L, heads, d, cross, with_p, seed=0):
    def run(ops):
        g = torch.Generator().manual_seed(seed)
        C = heads * d
        qkv = _mk(ops, _r((N, L, 3 * C), g))
        out = ops.zeros((N, L, C))
        Lp = (L + 7) // 8 * 8
        probs = ops.zeros((N, heads, L, Lp)) if with_p is True else None
        if with_p == "lse":          # flash-style: the forward keeps 4 bytes per query row, the backward recomputes P
            lse = ops.zeros((N, heads, L), torch.float32)
            ops.attn_fwd(qkv, heads, cross, 1.0 / math.sqrt(d), out, lse=lse)
            dout = _mk(ops, _r((N, L, C), g))
            dqkv = ops.zeros((N, L, 3 * C))
            ops.attn_bwd(qkv, None, out, dout, dqkv, heads, cross, 1.0 / math.sqrt(d), lse=lse)
            return [out, lse, dqkv]
        ops.attn_fwd(qkv, heads, cross, 1.0 / math.sqrt(d), out, probs[..., :L] if with_p else None)
        outs = [out] + ([probs] if with_p else [])
        if with_p and ops.attn_bwd_supported(L, d):
            # fused backward from the probabilities / output the forward just produced
            dout = _mk(ops, _r((N, L, C), g))
            dqkv = ops.zeros((N, L, 3 * C))
            ops.attn_bwd(qkv, probs[..., :L], out, dout, dqkv, heads, cross, 1.0 / math.sqrt(d))
            outs.append(dqkv)
        return outs
    return run


def all_cases():
    c = {
        "attn_fused_L256_d64_self": attn_case(4, 256, 4, 64, False, True),
        "attn_fused_L64_d128_cross": attn_case(4, 64, 4, 128, True, True, seed=1),
        "attn_fused_L256_d256_cross_nop": attn_case(2, 256, 4, 256, True, False, seed=2),
        "attn_fused_L16_d16_self": attn_case(2, 16, 4, 16, False, True, seed=3),
        "attn_fused_L36_d32_cross": attn_case(2, 36, 2, 32, True, True, seed=4),
        "attn_fused_L200_d64_many": attn_case(64, 200, 4, 64, True, True, seed=5),
        "attn_lse_L256_d64_self": attn_case(4, 256, 4, 64, False, "lse", seed=6),
        "attn_lse_L64_d128_cross": attn_case(4, 64, 4, 128, True, "lse", seed=7),
        "attn_lse_L36_d32_cross": attn_case(2, 36, 2, 32, True, "lse", seed=8),
        "attn_lse_L200_d64_many": attn_case(64, 200, 4, 64, True, "lse", seed=9),
        "attn_lse_L16_d16_self": attn_case(2, 16, 4, 16, False, "lse", seed=10),
        "pose_fp32_64": pose_case(3, 64, 64, True),
        "pose_fp32_24": pose_case(2, 24, 24, True, seed=3),
        "pose_bf16_16": pose_case(2, 16, 16, False, seed=4),
        "misc": misc_case(),
    }
    for i, (N, HW, C0, C1, mode, dp) in enumerate([
        (2, 256, 128, 0, 1, 0.0), (4, 64, 512, 256, 1, 0.0), (2, 4096, 128, 0, 2, 0.0), (2, 1024, 256, 128, 0, 0.0),
        (2, 16, 32, 0, 1, 0.0), (2, 36, 64, 32, 2, 0.0), (3, 64, 256, 0, 2, 0.1), (2, 64, 1024, 1024, 1, 0.0),
    ]):
        c[f"gn_{N}x{HW}x{C0}+{C1}_m{mode}_p{dp}"] = gn_case(N, HW, C0, C1, mode, dp, seed=i)
    c["gn_3x64x256+0_m2_p0.1_step5"] = gn_case(3, 64, 256, 0, 2, 0.1, seed=9, step=5)
    return c
