"""Shared case table for the tensor-core kernel: each case builds a GemmDesc on a given backend's
tensors so the same description runs on CudaOps (bf16) and on EmuOps (fp32 reference)."""
import math

import torch

from distributed_3d_diffusion_pytorch_b200 import gemm as G


def _rand(shape, gen, scale=1.0):
    return (torch.randn(shape, generator=gen) * scale).to(torch.bfloat16).float()


class Case:
    def __init__(self, name, build):
        self.name, self.build = name, build


def _mk(ops, t, dtype=None):
    """move a reference fp32 tensor to the backend (bf16 on CUDA, fp32 on emu)"""
    return t.to(device=ops.device, dtype=dtype or ops.act_dtype).contiguous()


def _with_stats(ops, d, outs, gn_cpg, rows_per_image, n_images):
    """ask for the GroupNorm statistics of the output in the epilogue (cases are chosen so that it is supported)"""
    d.gn_cpg, d.gn_rows_per_image = gn_cpg, rows_per_image
    assert ops.gemm_stats_supported(d), "test case must be eligible for epilogue statistics"
    d.gn_sums = ops.zeros((n_images, 32, 2), torch.float32)
    return d, outs + [d.gn_sums]


def conv_case(name, N, H, W, Cs, taps, Co, stride=1, bias=True, resid=False, rowvec=False, act=0, out2=False, alpha=1.0, out_scale=1.0,
              fp32_out=False, bn=0, seed=0, gn_cpg=0):
    def build(ops):
        g = torch.Generator().manual_seed(seed)
        Ho, Wo = (H + stride - 1) // stride if 9 in taps else H // stride, (W + stride - 1) // stride if 9 in taps else W // stride
        srcs = [_mk(ops, _rand((N, H, W, c), g)) for c in Cs]
        K = G.packed_k(list(zip(Cs, taps)))
        w = torch.zeros(Co, K)
        off = 0
        for c, tp in zip(Cs, taps):
            for t_ in range(tp):
                w[:, off:off + c] = _rand((Co, c), g, 1.0 / math.sqrt(sum(cc * tt for cc, tt in zip(Cs, taps))))
                off += G.ceil64(c)
        wp = _mk(ops, w)
        ldo = (Co + 7) // 8 * 8
        out = ops.zeros((N, Ho, Wo, ldo), torch.float32 if fp32_out else None)
        epi = dict(alpha=alpha, out_scale=out_scale, act=act, bn_tile=bn)
        if bias:
            epi["bias"] = _mk(ops, torch.randn(Co, generator=g), torch.float32)
        if resid:
            epi["resid"] = _mk(ops, _rand((N, Ho, Wo, ldo), g))
        if rowvec:
            epi["rowvec"] = _mk(ops, torch.randn(N, Co, generator=g), torch.float32)
        o2 = ops.zeros((N, Ho, Wo, ldo)) if out2 else None
        if out2:
            epi["out2"] = o2
        d = G.conv_desc(srcs, taps, wp, Co, out, stride, **epi)
        if gn_cpg:
            return _with_stats(ops, d, [out] + ([o2] if out2 else []), gn_cpg, 0, N)
        return d, [out] + ([o2] if out2 else [])
    return Case(name, build)


def dense_case(name, M, K, Ncols, seed=0, gn_cpg=0, rows_per_image=0, **kw):
    def build(ops):
        g = torch.Generator().manual_seed(seed)
        x = _mk(ops, _rand((M, K), g))
        w = torch.zeros(Ncols, G.ceil64(K))
        w[:, :K] = _rand((Ncols, K), g, 1.0 / math.sqrt(K))
        wp = _mk(ops, w)
        ldo = (Ncols + 7) // 8 * 8
        out = ops.zeros((M, ldo))
        epi = dict(kw)
        epi["bias"] = _mk(ops, torch.randn(Ncols, generator=g), torch.float32)
        d = G.dense_desc(x, wp, out, Ncols, **epi)
        if gn_cpg:
            return _with_stats(ops, d, [out], gn_cpg, rows_per_image, M // rows_per_image)
        return d, [out]
    return Case(name, build)


def wgrad_case(name, N, H, W, Co, Ci, taps, stride=1, split_k=None, seed=0):
    def build(ops):
        g = torch.Generator().manual_seed(seed)
        Hi, Wi = H * stride, W * stride
        dy = _mk(ops, _rand((N, H, W, Co), g, 0.25))
        x = _mk(ops, _rand((N, Hi, Wi, Ci), g, 0.25))
        Cpad = G.ceil64(Ci)
        out = ops.zeros((Co, taps * Cpad + 64), torch.float32)
        return G.wgrad_desc(dy, x, taps, out, 64, Cpad, stride, split_k), [out]
    return Case(name, build)


def attn_cases(name, N, Hh, L, d, cross, seed=0):
    """the six matmuls of one attention call on a packed qkv tensor [N, L, 3C]"""
    C = Hh * d
    Lp = (L + 7) // 8 * 8
    alpha = 1.0 / math.sqrt(d)

    def views(ops, g):
        qkv = _mk(ops, _rand((N, L, 3 * C), g))
        sp = lambda i: qkv[:, :, i * C:(i + 1) * C].reshape(N, L, Hh, d).permute(0, 2, 1, 3) if False else \
            qkv.view(N, L, 3, Hh, d)[:, :, i].permute(0, 2, 1, 3)
        return qkv, sp(0), sp(1), sp(2)

    def qk(ops):
        g = torch.Generator().manual_seed(seed)
        _, q, k, v = views(ops, g)
        s = ops.zeros((N, Hh, L, Lp))
        return G.attn_qk_desc(q, k, s[..., :L], cross, alpha), [s]

    def pv(ops):
        g = torch.Generator().manual_seed(seed + 1)
        _, q, k, v = views(ops, g)
        p = ops.zeros((N, Hh, L, Lp))
        p[..., :L] = _mk(ops, torch.softmax(torch.randn((N, Hh, L, L), generator=g), -1).to(torch.bfloat16).float())
        o = ops.zeros((N, L, C))
        return G.attn_pv_desc(p[..., :L], v, o.view(N, L, Hh, d).permute(0, 2, 1, 3), cross), [o]

    def dp(ops):
        g = torch.Generator().manual_seed(seed + 2)
        _, q, k, v = views(ops, g)
        do = _mk(ops, _rand((N, L, C), g))
        dpo = ops.zeros((N, Hh, L, Lp))
        return G.attn_dp_desc(do.view(N, L, Hh, d).permute(0, 2, 1, 3), v, dpo[..., :L], cross), [dpo]

    def dv(ops):
        g = torch.Generator().manual_seed(seed + 3)
        p = ops.zeros((N, Hh, L, Lp))
        p[..., :L] = _mk(ops, torch.softmax(torch.randn((N, Hh, L, L), generator=g), -1).to(torch.bfloat16).float())
        do = _mk(ops, _rand((N, L, C), g))
        dqkv = ops.zeros((N, L, 3 * C))
        dvv = dqkv.view(N, L, 3, Hh, d)[:, :, 2].permute(0, 2, 1, 3)
        return G.attn_dv_desc(p[..., :L], do.view(N, L, Hh, d).permute(0, 2, 1, 3), dvv, cross), [dqkv]

    def dq(ops):
        g = torch.Generator().manual_seed(seed + 4)
        _, q, k, v = views(ops, g)
        ds = ops.zeros((N, Hh, L, Lp))
        ds[..., :L] = _mk(ops, _rand((N, Hh, L, L), g, 0.1))
        dqkv = ops.zeros((N, L, 3 * C))
        dqv = dqkv.view(N, L, 3, Hh, d)[:, :, 0].permute(0, 2, 1, 3)
        return G.attn_dq_desc(ds[..., :L], k, dqv, cross, alpha), [dqkv]

    def dk(ops):
        g = torch.Generator().manual_seed(seed + 5)
        _, q, k, v = views(ops, g)
        ds = ops.zeros((N, Hh, L, Lp))
        ds[..., :L] = _mk(ops, _rand((N, Hh, L, L), g, 0.1))
        dqkv = ops.zeros((N, L, 3 * C))
        dkv = dqkv.view(N, L, 3, Hh, d)[:, :, 1].permute(0, 2, 1, 3)
        return G.attn_dk_desc(ds[..., :L], q, dkv, cross, alpha), [dqkv]

    return [Case(f"{name}.{n}", f) for n, f in (("qk", qk), ("pv", pv), ("dp", dp), ("dv", dv), ("dq", dq), ("dk", dk))]


def all_cases():
    c = [
        dense_case("dense_128x64x128", 128, 64, 128),
        dense_case("dense_256x128x256", 256, 128, 256),
        dense_case("dense_300x1024x512", 300, 1024, 512),
        dense_case("dense_8x1024x1024_silu", 8, 1024, 1024, act=1),
        dense_case("dense_1000x144x48", 1000, 144, 48),
        dense_case("dense_4096x1024x2048", 4096, 1024, 2048),
        dense_case("dense_512x256x384_bn192", 512, 256, 384),
        # 256-row tiles (two accumulators per B tile): need >= 2 tiles per SM, hence the larger pixel counts
        conv_case("conv_m2_rows_64x64_c64_128_resid", 20, 64, 64, [64], [9], 128, resid=True, out_scale=0.70710678, seed=11),
        conv_case("conv_m2_rows_30x32_c32_96_rowvec_out2", 84, 30, 32, [32], [9], 96, rowvec=True, out2=True, seed=12),
        conv_case("conv_m2_images_8x8_c64_64", 1210, 8, 8, [64], [9], 64, seed=13),
        conv_case("conv_m2_1x1_two_src_64x64", 19, 64, 64, [64, 32], [1, 1], 128, seed=14),
        dense_case("dense_m2_76000x192x128", 76000, 192, 128, seed=15),
        dense_case("dense_m2_80001x64x72_silu", 80001, 64, 72, act=1, seed=16),
        # CTA pairs (cta_group::2, M = 256 across two CTAs): full-width tiles (BN = 256), >= 16 K blocks, >= 2 tiles per pair
        conv_case("conv_pair_rows_32x32_c128_256_resid", 160, 32, 32, [128], [9], 256, resid=True, out_scale=0.70710678, seed=21),
        conv_case("conv_pair_images_8x8_c128_512_rowvec", 1300, 8, 8, [128], [9], 512, rowvec=True, seed=22),
        conv_case("conv_pair_two_src_1x1_20x20", 210, 20, 20, [640, 416], [1, 1], 256, act=1, seed=23),
        dense_case("dense_pair_80001x1024x256", 80001, 1024, 256, seed=24),
        dense_case("dense_pair_40000x1088x768", 40000, 1088, 768, seed=25),
        # CTA pairs whose CTAs keep two accumulators each (narrow tiles, 512 rows per pair tile)
        conv_case("conv_pairquad_rows_64x64_c128_128_resid", 40, 64, 64, [128], [9], 128, resid=True, out_scale=0.70710678, seed=31),
        conv_case("conv_pairquad_images_8x8_c128_64_rowvec", 2410, 8, 8, [128], [9], 64, rowvec=True, seed=32),
        conv_case("conv_pairquad_rows_30x32_c256_96", 86, 30, 32, [256], [9], 96, act=1, seed=33),
        dense_case("dense_pairquad_80001x1024x96", 80001, 1024, 96, seed=34),
        # GroupNorm statistics of the output taken in the epilogue (plain, 256-row, pair and pair-quad tiles; 1..2 images per box)
        conv_case("conv_stats_64x64_c64_128_cpg4", 2, 64, 64, [64], [9], 128, gn_cpg=4, seed=41),
        conv_case("conv_stats_8x8_c64_256_cpg8_resid", 6, 8, 8, [64], [9], 256, gn_cpg=8, resid=True, out_scale=0.70710678, seed=42),
        conv_case("conv_stats_16x16_c64_512_cpg16", 3, 16, 16, [64], [9], 512, gn_cpg=16, seed=43),
        conv_case("conv_stats_8x8_c32_1024_cpg32", 4, 8, 8, [32], [9], 1024, gn_cpg=32, seed=44),
        conv_case("conv_stats_m2_64x64_c64_128", 20, 64, 64, [64], [9], 128, gn_cpg=4, seed=45),
        conv_case("conv_stats_pair_32x32_c128_256", 160, 32, 32, [128], [9], 256, gn_cpg=8, seed=46),
        conv_case("conv_stats_pairquad_8x8_c128_128", 2432, 8, 8, [128], [9], 128, gn_cpg=4, seed=47),
        dense_case("dense_stats_2048x256x256_L64", 2048, 256, 256, gn_cpg=8, rows_per_image=64, seed=48),
        dense_case("dense_stats_81920x1024x256_L256_pair", 81920, 1024, 256, gn_cpg=8, rows_per_image=256, seed=49),
        conv_case("conv_16x16_c64_128", 2, 16, 16, [64], [9], 128),
        conv_case("conv_64x64_c128_128_resid", 2, 64, 64, [128], [9], 128, resid=True, out_scale=0.70710678),
        conv_case("conv_8x8_c512_512", 4, 8, 8, [512], [9], 512),
        conv_case("conv_24x24_c32_64", 1, 24, 24, [32], [9], 64),
        conv_case("conv_6x6_c64_64", 3, 6, 6, [64], [9], 64),
        conv_case("conv_3x3img_c128", 2, 3, 3, [128], [9], 128),
        conv_case("conv_stem_c8_128", 2, 32, 32, [8], [9], 128),
        conv_case("conv_last_c128_3", 2, 32, 32, [128], [9], 3),
        conv_case("conv_s2_c144_1024_rowvec_out2", 2, 32, 32, [144], [9], 256, stride=2, rowvec=True, out2=True),
        conv_case("conv_s4_c144", 2, 32, 32, [144], [9], 128, stride=4),
        conv_case("conv_s8_c144", 2, 64, 64, [144], [9], 128, stride=8),
        conv_case("conv_cat_256+128_1x1", 2, 16, 16, [256, 128], [1, 1], 128),
        conv_case("conv_fused_skip_3src", 2, 16, 16, [128, 64, 64], [9, 1, 1], 128, resid=False),
        conv_case("conv_partial_blocks_3src", 2, 16, 16, [88, 40, 72], [9, 1, 1], 128, rowvec=True),     # last 64-channel blocks with 2 / 3 / 1 real K steps
        conv_case("conv_c144_s1_pairwide", 4, 64, 64, [144], [9], 256),
        conv_case("conv_silu_grad_of_resid", 2, 32, 32, [128], [1], 256, resid=True, act=2, bias=False),      # value * silu'(resid): the FiLM data gradient
        conv_case("conv_silu_grad_of_resid_wide", 8, 64, 64, [512], [1], 1024, resid=True, act=2, bias=False),
        conv_case("conv_fp32out_alpha", 1, 16, 16, [64], [9], 64, fp32_out=True, alpha=0.5),
        wgrad_case("wgrad_dense_like", 1, 1, 4096, 256, 128, 1),
        wgrad_case("wgrad_16x16_c128_64", 2, 16, 16, 128, 64, 9),
        wgrad_case("wgrad_64x64_c128_128", 2, 64, 64, 128, 128, 9),
        wgrad_case("wgrad_8x8_c256_512_split1", 4, 8, 8, 256, 512, 9, split_k=1),
        wgrad_case("wgrad_24x24_c64_32", 1, 24, 24, 64, 32, 9),
        wgrad_case("wgrad_6x6_c64_64", 3, 6, 6, 64, 64, 9),
        wgrad_case("wgrad_s2_c144", 2, 16, 16, 128, 144, 9, stride=2),
        wgrad_case("wgrad_s8_c144", 2, 8, 8, 128, 144, 9, stride=8),
        wgrad_case("wgrad_last_co16", 2, 32, 32, 16, 128, 9),
        wgrad_case("wgrad_stem_ci8", 2, 32, 32, 128, 8, 9),
    ]
    c += attn_cases("attn_self_L256_d64", 4, 4, 256, 64, False)
    c += attn_cases("attn_cross_L64_d128", 4, 4, 64, 128, True)
    c += attn_cases("attn_cross_L16_d16", 2, 4, 16, 16, True)
    c += attn_cases("attn_self_L36_d32", 2, 4, 36, 32, False)
    return c
