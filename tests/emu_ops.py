"""TEST INFRASTRUCTURE ONLY.  Plain-torch fp32 statement of what every b3d entry point computes, with
the same tensor-level signatures as distributed_3d_diffusion_pytorch_b200.ops.CudaOps.

Two uses: (1) `-m gpu` tests compare each CUDA kernel with the matching method here (the "plain
PyTorch fp32 reference of the same op"); (2) `-m "not gpu"` tests run the engine's fwd/bwd
orchestration on CPU against the oracle, so layout / descriptor / ordering bugs are caught without a
GPU.  The product package never imports this file."""
import math

import torch
import torch.nn.functional as F


def _silu_grad(x):
    s = torch.sigmoid(x)
    return s * (1 + x * (1 - s))


def _ceil64(c):
    return (c + 63) // 64 * 64


SEED_STEP_STRIDE = 0x100000001B3      # B3D_SEED_STEP_STRIDE in csrc/b3d_common.cuh


def drop_keep(seed, idx, drop_p):
    """bit-exact mirror of the dropout keep decision (b3d_drop_pair / b3d_drop_thr16 in csrc/b3d_common.cuh, drop_mask8
    in csrc/groupnorm.cu) on an int64 tensor of flat element indices: True where the element is kept."""
    import numpy as np
    e = idx.numpy().astype(np.uint64)
    with np.errstate(over="ignore"):
        key = np.uint64((seed * 0x9E3779B97F4A7C15) & ((1 << 64) - 1)) + (e >> np.uint64(1))
        x = (key & np.uint64(0xFFFFFFFF)).astype(np.uint32) ^ ((key >> np.uint64(32)).astype(np.uint32) * np.uint32(0x9E3779B9))
        x ^= x >> np.uint32(16); x *= np.uint32(0x7feb352d)
        x ^= x >> np.uint32(15); x *= np.uint32(0x846ca68b)
        x ^= x >> np.uint32(16)
    field = np.where((e & np.uint64(1)) == 0, x & np.uint32(0xFFFF), x >> np.uint32(16)).astype(np.int64)
    thr16 = int(np.float32(drop_p) * np.float32(65536.0) + np.float32(0.5))
    return torch.from_numpy(field >= thr16)


class EmuOps:
    act_dtype = torch.float32
    name = "emu"

    def __init__(self, device="cpu", act_dtype=torch.float32, stats_dtype=None):
        self.device = torch.device(device)
        self.act_dtype = act_dtype
        # epilogue GroupNorm statistics are taken over the STORED output: per-kernel GPU comparisons round it like the GPU does
        self.stats_dtype = stats_dtype

    def side(self):
        import contextlib
        return contextlib.nullcontext()

    def join_side(self):
        pass

    def empty(self, shape, dtype=None):
        return torch.full(shape, float("nan"), dtype=dtype or self.act_dtype, device=self.device)

    def zeros(self, shape, dtype=None):
        return torch.zeros(shape, dtype=dtype or self.act_dtype, device=self.device)

    # ------------------------------------------------------------------ GEMM interpreter
    def gemm(self, d):
        s, pad = d.conv_stride, d.pad
        f32 = lambda t: t.float()
        if d.a_major == 0 and d.b_major == 0:
            N, H, W = d.N, d.H, d.W
            cols = []
            for t, taps in zip(d.a, d.a_taps):
                t = f32(t)                                   # [n, y, x, c]
                n_, hi, wi, c = t.shape
                cp = _ceil64(c)
                for tp in range(taps):
                    dx, dy = ((tp % 3) - pad, (tp // 3) - pad) if taps == 9 else (0, 0)
                    blk = torch.zeros(N, H, W, cp)
                    ys = torch.arange(H) * s + dy
                    xs = torch.arange(W) * s + dx
                    vy = (ys >= 0) & (ys < hi)
                    vx = (xs >= 0) & (xs < wi)
                    sub = t[:N][:, ys[vy]][:, :, xs[vx]]
                    tmp = torch.zeros(N, int(vy.sum()), W, c)
                    tmp[:, :, vx] = sub
                    blk[:, vy, :, :c] = tmp
                    cols.append(blk)
            A = torch.cat(cols, -1)                          # [N,H,W,Ktot]
            Bm = f32(d.b)                                    # [bn, by, cols, K]
            K = min(A.shape[-1], _ceil64(Bm.shape[-1]))
            Bp = torch.zeros(*Bm.shape[:-1], A.shape[-1])
            Bp[..., :Bm.shape[-1]] = Bm
            if d.batched:
                idx = torch.arange(N) ^ d.b_batch_xor
                D = torch.einsum("nyxk,nyck->nyxc", A, Bp[idx][:, :, :d.n_cols])
            else:
                D = torch.einsum("nyxk,ck->nyxc", A, Bp[0, 0, :d.n_cols])
            img = torch.arange(N)
        elif d.a_major == 0 and d.b_major == 1:
            assert d.batched
            A = f32(d.a[0])                                  # [n, y, x, k]
            Bm = f32(d.b)                                    # [n, y, k, cols]
            idx = torch.arange(d.N) ^ d.b_batch_xor
            D = torch.einsum("nyxk,nykc->nyxc", A, Bm[idx][..., :d.n_cols])
            img = torch.arange(d.N)
        else:
            A = f32(d.a[0])                                  # [n, y, x, rows]
            Bm = f32(d.b)                                    # [n, y', x', cols]
            if d.batched:
                idx = torch.arange(d.N) ^ d.b_batch_xor
                D = torch.einsum("nyxr,nyxc->nyrc", A[..., :d.m_rows], Bm[idx][..., :d.n_cols])
            else:
                N, H, W = d.N, d.H, d.W
                outs = []
                for tp in range(d.b_taps):
                    dx, dy = ((tp % 3) - pad, (tp // 3) - pad) if d.b_taps == 9 else (0, 0)
                    hi, wi = Bm.shape[1], Bm.shape[2]
                    ys = torch.arange(H) * s + dy
                    xs = torch.arange(W) * s + dx
                    vy = (ys >= 0) & (ys < hi)
                    vx = (xs >= 0) & (xs < wi)
                    sh = torch.zeros(N, H, W, Bm.shape[-1])
                    tmp = torch.zeros(N, int(vy.sum()), W, Bm.shape[-1])
                    tmp[:, :, vx] = Bm[:N][:, ys[vy]][:, :, xs[vx]]
                    sh[:, vy] = tmp
                    outs.append(torch.einsum("nyxr,nyxc->rc", A[:N, :H, :W, :d.m_rows], sh[..., :d.n_cols]))
                D = outs
            img = None
        self._epilogue(d, D, img)
        if getattr(d, "gn_sums", None) is not None:
            # statistics of the STORED output (after the rounding of the activation dtype), per image and group of gn_cpg channels
            o = d.out[..., :d.n_cols]
            o = (o.to(self.stats_dtype) if self.stats_dtype is not None else o).float()
            if d.gn_rows_per_image > 0:
                o = o.reshape(-1, d.gn_rows_per_image, d.n_cols)
            else:
                o = o.reshape(o.shape[0], -1, d.n_cols)
            g = o.reshape(o.shape[0], o.shape[1], 32, d.gn_cpg)
            d.gn_sums[..., 0] += g.sum((1, 3))
            d.gn_sums[..., 1] += (g * g).sum((1, 3))

    def gemm_stats_supported(self, d):
        if d.a_major != 0 or d.b_major != 0 or d.batched or d.gn_cpg not in (4, 8, 16, 32):      # (the emu's activations are fp32)
            return False
        if d.n_cols != 32 * d.gn_cpg or (d.resid is not None and d.out2 is not None):
            return False
        bw = min(d.W, 128)
        bh = max(1, min(128 // bw, d.H))
        bnn = max(1, min(128 // (bw * bh), d.N)) if bh == d.H else 1
        if bw * bh * bnn != 128 or d.W % bw or d.H % bh or d.N % bnn:
            return False
        rps = 64 // d.gn_cpg
        rpi = d.gn_rows_per_image if d.gn_rows_per_image > 0 else bw * bh
        if d.gn_rows_per_image > 0 and not (d.H == 1 and d.N == 1):
            return False
        rpib = min(rpi, 128)
        if rpib % rps or 128 % rpib or (rpi > 128 and rpi % 128):
            return False
        spi = rpib // rps
        return spi & (spi - 1) == 0

    def _epilogue(self, d, D, img):
        def finish(v, dst, dst2, resid, rowvec):
            v = v * d.alpha
            if d.bias is not None:
                v = v + d.bias[:d.n_cols].float()
            if rowvec is not None:
                v = v + rowvec
            if resid is not None:
                v = v * _silu_grad(resid.float()) if d.act == 2 else v + resid.float()
            v = v * d.out_scale
            if d.act == 1 and dst2 is None:
                v = F.silu(v)
            if d.out_atomic:
                dst.add_(v.to(dst.dtype))
            else:
                dst.copy_(v.to(dst.dtype))
            if dst2 is not None:
                dst2.copy_(F.silu(v).to(dst2.dtype))

        o = d.out
        if isinstance(D, list):           # wgrad taps: out is [1,1,rows,>=cols] plus tap stride
            for tp, v in enumerate(D):
                flat = o.as_strided((d.m_rows, d.n_cols), (o.stride(2), 1), o.storage_offset() + tp * d.out_tap_stride)
                finish(v, flat, None, None, None)
            return
        n_idx = torch.arange(D.shape[0]) ^ d.out_batch_xor
        rows = D.shape[2]
        dst = o[:, :, :rows, :d.n_cols]
        rv = d.rowvec[:, :d.n_cols].float()[:, None, None, :] if d.rowvec is not None else None
        resid = d.resid[:, :, :rows, :d.n_cols][n_idx] if d.resid is not None else None
        tmp = torch.empty_like(D)
        tmp2 = torch.empty_like(D) if d.out2 is not None else None
        if d.out_atomic:
            tmp.copy_(dst[n_idx].float())
        finish(D, tmp, tmp2, resid, rv)
        dst[n_idx] = tmp.to(dst.dtype)
        if d.out2 is not None:
            d.out2[:, :, :rows, :d.n_cols][n_idx] = tmp2.to(d.out2.dtype)

    def attn_fwd(self, qkv, heads, cross, scale, out, probs=None, lse=None):
        N, L, C3 = qkv.shape
        C = C3 // 3
        d = C // heads
        q, k, v = (qkv.float().view(N, L, 3, heads, d)[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        idx = torch.arange(N) ^ int(cross)
        S = (q @ k[idx].transpose(-1, -2)) * scale
        P = torch.softmax(S, -1)
        if lse is not None:          # log2-domain log-sum-exp of the scaled scores
            lse.copy_(torch.logsumexp(S, -1) * 1.4426950408889634)
        if probs is not None:
            probs[..., :L] = P.to(probs.dtype)
            P = probs[..., :L].float()
        out.copy_((P @ v[idx]).permute(0, 2, 1, 3).reshape(N, L, C).to(out.dtype))

    def attn_bwd_supported(self, L, d):
        Lb2 = (L + 127) // 128 * 128
        return 1 <= L <= 256 and 16 <= d <= 128 and d % 16 == 0 and Lb2 + 2 * (Lb2 // 128) * d <= 512 and d <= Lb2

    def attn_bwd(self, qkv, probs, o, dout, dqkv, heads, cross, scale, lse=None):
        N, L, C3 = qkv.shape
        C = C3 // 3
        d = C // heads
        q, k, v = (qkv.float().view(N, L, 3, heads, d)[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        idx = torch.arange(N) ^ int(cross)
        if probs is not None:
            P = probs[..., :L].float()
        else:                        # recomputed from the row statistics, rounded like the tile the kernel builds
            P = torch.exp2((q @ k[idx].transpose(-1, -2)) * (scale * 1.4426950408889634) - lse.float()[..., None]).to(self.act_dtype).float()
        dO = dout.float().view(N, L, heads, d).permute(0, 2, 1, 3)
        O = o.float().view(N, L, heads, d).permute(0, 2, 1, 3)
        dP = dO @ v[idx].transpose(-1, -2)
        D = (dO * O).sum(-1, keepdim=True)
        dS = (P * (dP - D)).to(self.act_dtype).float()
        dq = scale * (dS @ k[idx])
        dk = torch.zeros_like(dq)
        dv = torch.zeros_like(dq)
        dk[idx] = scale * (dS.transpose(-1, -2) @ q)
        dv[idx] = P.transpose(-1, -2) @ dO
        out = torch.stack([t.permute(0, 2, 1, 3).reshape(N, L, C) for t in (dq, dk, dv)], 2).reshape(N, L, 3 * C)
        dqkv.copy_(out.to(dqkv.dtype))

    # ------------------------------------------------------------------ conditioning
    def pose_embedding(self, K, R, t, mask, pos_emb, first_emb, other_emb, out):
        B = K.shape[0]
        N, H, W, ld = out.shape
        Kinv = torch.linalg.inv(K.double())
        M = R.double() @ Kinv[:, None]
        v, u = torch.meshgrid(torch.arange(H, dtype=torch.float64), torch.arange(W, dtype=torch.float64), indexing="ij")
        p = torch.stack([u + 0.5, v + 0.5, torch.ones_like(u)], -1)
        dr = torch.einsum("bfij,hwj->bfhwi", M, p)
        dr = (dr / dr.norm(dim=-1, keepdim=True)).float()
        pos = t.double()[:, :, None, None, :].expand(B, 2, H, W, 3).float()

        def nerf(x, D):
            sc = torch.tensor([2.0 ** i for i in range(D)])
            xb = (x[..., None, :] * sc[:, None]).reshape(*x.shape[:-1], -1)
            return torch.cat([x, torch.sin(torch.cat([xb, xb + torch.pi / 2.0], -1))], -1)

        pe = torch.cat([nerf(pos, 15), nerf(dr, 8)], -1)
        pe = torch.where(mask.bool()[:, None, None, None, None], pe, torch.zeros_like(pe))
        if pos_emb is not None:
            pe = pe + pos_emb.permute(1, 2, 0)[None, None]
        if first_emb is not None:
            pe = pe + torch.stack([first_emb.flatten(), other_emb.flatten()], 0)[None, :, None, None, :]
        out[..., :144] = pe.reshape(N, H, W, 144).to(out.dtype)

    def pose_embedding_bwd(self, dP, dpos, dfirst, dother):
        g = dP[..., :144]
        if dpos is not None:
            dpos.copy_((g[0] + g[1]).permute(2, 0, 1))
        if dfirst is not None:
            dfirst.view(-1).copy_(g[0].sum((0, 1)))
            dother.view(-1).copy_(g[1].sum((0, 1)))

    def ddpm_embedding(self, logsnr, out):
        n, e = out.shape
        tt = logsnr.flatten().float().clamp(-20, 20) * 1000.0
        half = e // 2
        f = torch.exp(torch.arange(half, dtype=torch.float32) * -(math.log(10000.0) / (half - 1)))
        a = tt[:, None] * f
        out.copy_(torch.cat([a.sin(), a.cos()], -1).to(out.dtype))

    def pack_input(self, x, z, out):
        B = x.shape[0]
        h = torch.stack([x, z], 1).reshape(2 * B, 3, *x.shape[2:]).permute(0, 2, 3, 1)
        out.zero_()
        out[..., :3] = h.to(out.dtype)

    def unpack_input_grad(self, g, dx, dz):
        B = dx.shape[0]
        gg = g[..., :3].float().permute(0, 3, 1, 2).reshape(B, 2, 3, *dx.shape[2:])
        dx.copy_(gg[:, 0])
        if dz is not None:
            dz.copy_(gg[:, 1])

    def unpack_output(self, y, eps):
        B = eps.shape[0]
        eps.copy_(y[..., :3].float().permute(0, 3, 1, 2).reshape(B, 2, 3, *eps.shape[2:])[:, 1])

    def pack_output_grad(self, deps, dy):
        B = deps.shape[0]
        dy.zero_()
        v = dy.view(B, 2, *dy.shape[1:])
        v[:, 1, :, :, :3] = deps.permute(0, 2, 3, 1).to(dy.dtype)

    # ------------------------------------------------------------------ parameter layouts
    def pack_weight(self, w, mode, out, ci0=0, cn=None, Cpad=None, k_off=0):
        Co, Ci = w.shape[0], w.shape[1]
        taps = w.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        w3 = w.reshape(Co, Ci, taps).float()
        if mode == 0:
            blk = torch.zeros(Co, taps, Cpad)
            blk[:, :, :cn] = w3[:, ci0:ci0 + cn].permute(0, 2, 1)
            out[:Co, k_off:k_off + taps * Cpad] = blk.reshape(Co, -1).to(out.dtype)
        elif mode == 1:
            blk = torch.zeros(Ci, taps, Cpad)
            blk[:, :, :Co] = w3.flip(2).permute(1, 2, 0)
            out[:Ci, k_off:k_off + taps * Cpad] = blk.reshape(Ci, -1).to(out.dtype)
        else:
            out[:taps * Ci, :Co] = w3.permute(2, 1, 0).reshape(taps * Ci, Co).to(out.dtype)

    def unpack_wgrad(self, g, k_off, Cpad, dw, ci0=0, cn=None, accumulate=False):
        Co, Ci = dw.shape[0], dw.shape[1]
        taps = dw.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        blk = g[:Co, k_off:k_off + taps * Cpad].reshape(Co, taps, Cpad)[:, :, :cn].permute(0, 2, 1)
        dst = dw.view(Co, Ci, taps)[:, ci0:ci0 + cn]
        if accumulate:
            dst.add_(blk)
        else:
            dst.copy_(blk)

    # recorded variants (one-launch job lists on the GPU): here the jobs are closures replayed by run_jobs
    def record_pack_weight(self, jl, w, mode, out, ci0=0, cn=None, Cpad=None, k_off=0):
        jl.jobs.append(lambda: self.pack_weight(w, mode, out, ci0, cn, Cpad, k_off))

    def record_copy(self, jl, src, dst):
        jl.jobs.append(lambda: dst.copy_(src))

    def record_unpack_wgrad(self, jl, g, k_off, Cpad, dw, ci0=0, cn=None, accumulate=False):
        jl.jobs.append(lambda: self.unpack_wgrad(g, k_off, Cpad, dw, ci0, cn, accumulate))

    def run_jobs(self, jl):
        for j in jl.jobs:
            j()

    def scale_colsum(self, a, out, scale, colsum):
        v = a.float() * scale
        out.copy_(v.to(out.dtype))
        colsum += v.reshape(-1, a.shape[-1]).sum(0)

    def cast(self, src, out):
        out.copy_(src.to(out.dtype))

    def rank1_add(self, A, u, v):
        A.add_(torch.outer(u.reshape(-1), v.reshape(-1)))

    def matvec(self, W, v, out, transpose=False, add=None):
        y = (W.t() if transpose else W).double() @ v.double()
        out.copy_((y + (add.double() if add is not None else 0.0)).float())

    # ------------------------------------------------------------------ resampling / elementwise
    def avgpool2(self, x, out, scale=0.25):
        N, H, W, C = x.shape
        out.copy_((x.float().reshape(N, H // 2, 2, W // 2, 2, C).sum((2, 4)) * scale).to(out.dtype))

    def upsample2(self, x, out, scale=1.0):
        out.copy_((x.float().repeat_interleave(2, 1).repeat_interleave(2, 2) * scale).to(out.dtype))

    def silu_bwd(self, dy, x, dx):
        dx.copy_((dy.float() * _silu_grad(x.float())).to(dx.dtype))

    def silu_bwd_colsum(self, dy, x, dx, rows_per_group, group_out, total_out):
        C = dy.shape[-1]
        v = (dy.float() * _silu_grad(x.float()) if x is not None else dy.float()).reshape(-1, rows_per_group, C)
        if dx is not None:
            dx.copy_(v.reshape(dx.shape).to(dx.dtype))
        if group_out is not None:
            group_out[:, :C].add_(v.sum(1))
        if total_out is not None:
            total_out[:C].add_(v.sum((0, 1)))

    def add_scale(self, a, b, out, scale):
        v = a.float() + (b.float() if b is not None else 0)
        out.copy_((v * scale).to(out.dtype))

    def colsum(self, x2d, rows_per_group, out, scale=1.0):
        rows, C = x2d.shape
        out.view(-1, out.shape[-1])[:, :C].add_(x2d.float().reshape(rows // rows_per_group, rows_per_group, C).sum(1) * scale)

    def batch_frame_sum(self, x, out):
        B = x.shape[0] // 2
        out.copy_(x.float().reshape(B, 2, *x.shape[1:]).sum(0).to(out.dtype))

    def conv_tap_scatter(self, G, C, h, w, s, dP, accumulate):
        _, H, W, _ = dP.shape
        g = G[:, :9 * C].reshape(2, h, w, 9, C)
        acc = torch.zeros(2, H + 2 * s + 2, W + 2 * s + 2, C)
        for ky in range(3):
            for kx in range(3):
                # input pixel (s*y + ky - 1, s*x + kx - 1) receives g[:, y, x, tap]
                ys = torch.arange(h) * s + ky - 1 + 1
                xs = torch.arange(w) * s + kx - 1 + 1
                acc[:, ys[:, None], xs[None, :]] += g[:, :, :, ky * 3 + kx]
        r = acc[:, 1:1 + H, 1:1 + W]
        if accumulate:
            dP[..., :C] += r
        else:
            dP[..., :C] = r

    def softmax(self, S, P):
        P.copy_(torch.softmax(S.float(), -1).to(P.dtype))

    def softmax_bwd(self, P, dP, dS):
        p, g = P.float(), dP.float()
        dS.copy_((p * (g - (p * g).sum(-1, keepdim=True))).to(dS.dtype))

    # ------------------------------------------------------------------ GroupNorm family
    @staticmethod
    def _cat(x0, x1):
        return torch.cat([x0.float(), x1.float()], -1) if x1 is not None else x0.float()

    def gn_stats(self, x0, x1, sums):
        x = self._cat(x0, x1)
        N, HW, C = x.shape
        g = x.reshape(N, HW, 32, C // 32)
        sums[:, :, 0] += g.sum((1, 3))
        sums[:, :, 1] += (g * g).sum((1, 3))

    @staticmethod
    def _norm(x, sums):
        N, HW, C = x.shape
        cnt = HW * (C // 32)
        mean = sums[:, :, 0] / cnt
        var = (sums[:, :, 1] / cnt - mean * mean).clamp_min(0)
        rstd = torch.rsqrt(var + 1e-5)
        mean_c = mean.repeat_interleave(C // 32, 1)[:, None, :]
        rstd_c = rstd.repeat_interleave(C // 32, 1)[:, None, :]
        return (x - mean_c) * rstd_c, rstd_c

    def _mask(self, shape, drop_p, seed):
        if drop_p <= 0:
            return torch.ones(shape)
        idx = torch.arange(int(torch.tensor(shape).prod())).reshape(shape)
        keep = drop_keep(seed, idx, drop_p)
        return torch.where(keep, torch.full(shape, 1.0 / (1.0 - drop_p)), torch.zeros(shape))

    def gn_apply(self, x0, x1, sums, gamma, beta, mode, out, ss=None, F=0, drop_p=0.0, seed=0, seed_ptr=None, ss_image=None):
        seed = (seed + (int(seed_ptr.item()) * SEED_STEP_STRIDE if seed_ptr is not None else 0)) & ((1 << 64) - 1)
        x = self._cat(x0, x1)
        xh, _ = self._norm(x, sums)
        y = xh * gamma.float() + beta.float()
        if mode == 1:
            y = torch.nn.functional.silu(y)
        elif mode == 2:
            N, HW, C = x.shape
            sc = ss[:, :F].float().reshape(-1, HW, F)
            sh = ss[:, F:2 * F].float().reshape(-1, HW, F)
            if ss_image is not None:          # image n takes the ss rows of image ss_image[n]
                sc, sh = sc[ss_image.long()], sh[ss_image.long()]
            y = (y * (1 + sc) + sh) * self._mask(y.shape, drop_p, seed)
        out.copy_(y.to(out.dtype))

    def cond_silu(self, c, lemb, lemb_row, out):
        out.copy_(F.silu(c.float() + lemb.float()[lemb_row.long()][:, None, None, :]).to(out.dtype))

    def gn_stats_apply(self, x0, x1, sums, gamma, beta, mode, out, **kw):
        self.gn_stats(x0, x1, sums)
        self.gn_apply(x0, x1, sums, gamma, beta, mode, out, **kw)

    def gn_bwd(self, x0, x1, sums, gamma, beta, mode, dy, dgamma, dbeta, ws, dx0, dx1=None, ss=None, F=0, dss=None, drop_p=0.0, seed=0,
               add0=None, add1a=None, add1b=None, seed_ptr=None, dss_sum=None, dx_sum=None, add0_scale=1.0):
        seed = (seed + (int(seed_ptr.item()) * SEED_STEP_STRIDE if seed_ptr is not None else 0)) & ((1 << 64) - 1)
        x = self._cat(x0, x1)
        N, HW, C = x.shape
        xh, rstd_c = self._norm(x, sums)
        z = xh * gamma.float() + beta.float()
        d = dy.float()
        if mode == 1:
            d = d * _silu_grad(z)
        elif mode == 2:
            d = d * self._mask(d.shape, drop_p, seed)
            dss[:, :F] = (d * z).reshape(N * HW, F).to(dss.dtype)
            dss[:, F:2 * F] = d.reshape(N * HW, F).to(dss.dtype)
            if dss_sum is not None:
                dss_sum[:F] += (d * z).sum((0, 1))
                dss_sum[F:2 * F] += d.sum((0, 1))
            d = d * (1 + ss[:, :F].float().reshape(N, HW, F))
        dgamma += (d * xh).sum((0, 1))
        dbeta += d.sum((0, 1))
        dg = d * gamma.float()
        cpg = C // 32
        cnt = HW * cpg
        s1 = dg.reshape(N, HW, 32, cpg).sum((1, 3))
        s2 = (dg * xh).reshape(N, HW, 32, cpg).sum((1, 3))
        ws[:, :, 0] += s1
        ws[:, :, 1] += s2
        s1c = (s1 / cnt).repeat_interleave(cpg, 1)[:, None, :]
        s2c = (s2 / cnt).repeat_interleave(cpg, 1)[:, None, :]
        dx = rstd_c * (dg - s1c - xh * s2c)
        if dx_sum is not None:
            dx_sum += dx.sum((0, 1))
        if add0 is not None:
            dx = dx + add0.float() * add0_scale
        C0 = x0.shape[-1]
        if add1a is not None:
            dx[..., :C0] += add1a.float()
        if add1b is not None:
            dx[..., C0:] += add1b.float()
        dx0.copy_(dx[..., :C0].to(dx0.dtype))
        if dx1 is not None:
            dx1.copy_(dx[..., C0:].to(dx1.dtype))

    # ------------------------------------------------------------------ diffusion glue
    def q_sample(self, z, noise, logsnr, z_out, x=None, xrand=None, mask=None, x_out=None):
        a = logsnr.sigmoid().sqrt()[:, None, None, None]
        s = (-logsnr).sigmoid().sqrt()[:, None, None, None]
        z_out.copy_(a * z + s * noise)
        if x_out is not None:
            x_out.copy_(torch.where(mask.bool()[:, None, None, None], x, xrand))

    def loss_scratch(self):
        return torch.zeros(1025)

    def loss(self, noise, pred, dpred, loss_type, scratch, loss_out, grad_scale=1.0):
        p = pred.detach().clone().requires_grad_(True)
        fn = [F.mse_loss, F.l1_loss, F.smooth_l1_loss][loss_type]
        l = fn(noise, p)
        l.backward()
        loss_out.copy_(l.detach())
        if dpred is not None:
            dpred.copy_(p.grad * grad_scale)

    def sampler_update(self, z, eps_c, eps_u, w, noise, out, logsnr, logsnr_next):
        logsnr, logsnr_next = torch.tensor(float(logsnr)), torch.tensor(float(logsnr_next))
        ww = w.float()[:, None, None, None]
        c = -torch.special.expm1(logsnr - logsnr_next)
        alpha, sigma, alpha_n = logsnr.sigmoid().sqrt(), (-logsnr).sigmoid().sqrt(), logsnr_next.sigmoid().sqrt()
        e = (1 + ww) * eps_c - ww * eps_u
        x0 = ((z - sigma * e) / alpha).clamp(-1, 1)
        m = alpha_n * (z * (1 - c) / alpha + c * x0)
        if logsnr_next != 0:
            m = m + ((-logsnr_next).sigmoid() * c).sqrt() * noise
        out.copy_(m)

    def sampler_update_dev(self, z, eps_c, eps_u, w, noise, out, sched):
        self.sampler_update(z, eps_c, eps_u, w, noise, out, float(sched[0]), float(sched[1]))

    def adam(self, p, g, m, v, lr, beta1, beta2, eps, step, grad_scale=1.0):
        gg = g.float() * grad_scale
        m.mul_(beta1).add_(gg, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(gg, gg, value=1 - beta2)
        p.sub_(lr * (m / (1 - beta1 ** step)) / ((v / (1 - beta2 ** step)).sqrt() + eps))
