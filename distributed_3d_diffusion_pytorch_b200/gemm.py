"""Builders of GemmDesc (ops.py / include/b3d.h) for the GEMM-shaped ops of the model.
All tensors are NHWC views [n, y, x, c] with stride(c) == 1."""
import math

from .ops import GemmDesc

NUM_SMS = 148


def ceil64(c):
    return (c + 63) // 64 * 64


def packed_k(chans_taps):
    """K extent of a packed weight whose A operand is the concat of (channels, taps) sources."""
    return sum(t * ceil64(c) for c, t in chans_taps)


def conv_desc(srcs, taps, wpack, n_cols, out, stride=1, **epi):
    """3x3 (taps=9, pad 1) / 1x1 (taps=1) convolution of the implicit concat of `srcs`.
    wpack: bf16 [>=n_cols, K] with K = packed_k(...).  out: [N,Ho,Wo,>=n_cols]."""
    N, Ho, Wo, _ = out.shape
    return GemmDesc(a=list(srcs), a_taps=list(taps), b=wpack[None, None], out=out, n_cols=n_cols, conv_stride=stride,
                    pad=1 if 9 in taps else 0, W=Wo, H=Ho, N=N, **epi)


def dense_desc(x2d, wpack, out2d, n_cols, **epi):
    """out2d[M, n_cols] = x2d[M, K] @ wpack[n_cols, K]^T."""
    M = x2d.shape[0]
    for k in ("resid", "out2"):
        if epi.get(k) is not None:
            epi[k] = epi[k][None, None]
    return GemmDesc(a=[x2d[None, None]], a_taps=[1], b=wpack[None, None], out=out2d[None, None], n_cols=n_cols, W=M, H=1, N=1, **epi)


def pick_split_k(tiles, kblocks):
    if tiles >= NUM_SMS:
        return 1
    # enough CTAs to fill the GPU about twice, but every split keeps >= 8 K blocks (its prologue, its 128 x N fp32
    # atomics and its launch slot are pure overhead for small problems)
    s = min(max(1, (2 * NUM_SMS) // max(tiles, 1)), max(1, kblocks // 8))
    return max(1, s)


def wgrad_desc(dy, x, taps, out_packed, k_off, Cpad, stride=1, split_k=None, alpha=1.0):
    """out_packed[co, k_off + tap*Cpad + ci] += sum_pixels dy[p, co] * x[p*stride + tap - pad, ci]  (fp32, atomic)."""
    N, H, W, Co = dy.shape
    Ci = x.shape[-1]
    bn = min(256, (Ci + 15) // 16 * 16) if Ci <= 256 else 0
    # narrow unpadded slabs: b3d_gemm puts 256 // Ci taps side by side in one tile (csrc/gemm_tc.cu, "tpg")
    tpg = 256 // Ci if (taps == 9 and Ci % 64 == 0 and Ci <= 128 and Cpad == Ci) else 1
    tiles = ((Co + 127) // 128) * max(1, math.ceil(Ci / 256)) * math.ceil(taps / tpg)
    kblocks = max(1, (N * H * W) // 64)
    if split_k is None:
        split_k = pick_split_k(tiles, kblocks)
    o = out_packed[:, k_off:][None, None]
    return GemmDesc(a=[dy], b=x, out=o, n_cols=Ci, m_rows=Co, a_major=1, b_major=1, conv_stride=stride, pad=1 if taps == 9 else 0,
                    b_taps=taps, W=W, H=H, N=N, split_k=split_k, out_tap_stride=Cpad, out_atomic=True, bn_tile=bn, alpha=alpha)


def wgrad_dense_desc(dy2d, x2d, out_packed, k_off=0, split_k=None, alpha=1.0):
    return wgrad_desc(dy2d[None, None], x2d[None, None], 1, out_packed, k_off, ceil64(x2d.shape[1]), 1, split_k, alpha)


# ---- attention (nn.MultiheadAttention internals); q/k/v/o: [N, heads, L, d] views, s/p: [N, heads, Lq, Lk]
def attn_qk_desc(q, k, s_out, cross, alpha):
    N, Hh, L, d = q.shape
    return GemmDesc(a=[q], b=k, out=s_out, n_cols=k.shape[2], W=L, H=Hh, N=N, batched=True, b_batch_xor=int(cross), alpha=alpha)


def attn_pv_desc(p, v, o_out, cross):
    N, Hh, L, Lk = p.shape
    return GemmDesc(a=[p], b=v, out=o_out, n_cols=v.shape[3], W=L, H=Hh, N=N, a_major=0, b_major=1, batched=True, b_batch_xor=int(cross))


def attn_dp_desc(do, v, dp_out, cross):
    N, Hh, L, d = do.shape
    return GemmDesc(a=[do], b=v, out=dp_out, n_cols=v.shape[2], W=L, H=Hh, N=N, batched=True, b_batch_xor=int(cross))


def attn_dv_desc(p, do, dv_out, cross):
    """dv[n ^ cross] = p[n]^T @ do[n]   (rows = Lk)"""
    N, Hh, Lq, Lk = p.shape
    return GemmDesc(a=[p], b=do, out=dv_out, n_cols=do.shape[3], m_rows=Lk, a_major=1, b_major=1, W=Lq, H=Hh, N=N, batched=True,
                    out_batch_xor=int(cross))


def attn_dq_desc(ds, k, dq_out, cross, alpha):
    N, Hh, Lq, Lk = ds.shape
    return GemmDesc(a=[ds], b=k, out=dq_out, n_cols=k.shape[3], W=Lq, H=Hh, N=N, a_major=0, b_major=1, batched=True,
                    b_batch_xor=int(cross), alpha=alpha)


def attn_dk_desc(ds, q, dk_out, cross, alpha):
    N, Hh, Lq, Lk = ds.shape
    return GemmDesc(a=[ds], b=q, out=dk_out, n_cols=q.shape[3], m_rows=Lk, a_major=1, b_major=1, W=Lq, H=Hh, N=N, batched=True,
                    out_batch_xor=int(cross), alpha=alpha)
