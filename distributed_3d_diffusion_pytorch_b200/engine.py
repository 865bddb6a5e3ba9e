"""Execution engine of the X-UNet hot path: the forward and the hand-written backward of
`XUNet.forward` (reference xunet.py:477-536) as explicit sequences of b3d kernel launches.

* activations: bf16 NHWC, images n = 2*b + frame; every buffer is allocated once per (batch size) and
  reused, so a step allocates nothing and can be captured in a CUDA graph;
* parameters arrive as fp32 tensors under the reference's state_dict names; they are repacked to
  bf16 K-major GEMM operands at the start of each forward (they change every optimizer step);
* gradients leave as fp32 tensors under the same names (PyTorch layouts);
* FiLM (xunet.py:74-87) is restructured: the per-pixel Linear(1024 -> 2F) of ALL ResnetBlocks of a
  resolution level is ONE GEMM against the level's silu(emb) (forward), and ONE data-gradient GEMM
  (backward); GroupNorm/FiLM/dropout/SiLU are applied by the fused GN kernels.

`ops` is the kernel backend (ops.CudaOps); tests substitute a torch statement of the same ops to
check this orchestration on CPU."""
import contextlib
import math
import os

import torch

from . import gemm as G
from .ops import JobList

RSQRT2 = 1.0 / math.sqrt(2.0)


class Cfg:
    """XUNet hyper-parameters (class attributes of the reference, xunet.py:356-366)."""
    DEFAULTS = dict(H=128, W=128, ch=256, ch_mult=(1, 2, 2, 4), emb_ch=1024, num_res_blocks=3, attn_resolutions=(2, 3, 4),
                    attn_heads=4, dropout=0.1, use_pos_emb=True, use_ref_pose_emb=True)

    def __init__(self, **kw):
        for k, v in self.DEFAULTS.items():
            setattr(self, k, kw.get(k, v))
        self.num_resolutions = len(self.ch_mult)
        self.dim_in = [self.ch] + [self.ch * m for m in self.ch_mult[:-1]]
        self.dim_out = [self.ch * m for m in self.ch_mult]


class ResSpec:
    def __init__(self, name, lvl, cin, f, resample, cat):
        self.name, self.lvl, self.cin, self.f, self.resample, self.cat = name, lvl, cin, f, resample, cat
        self.film_off = 0


class AttnSpec:
    def __init__(self, name, lvl, c, cross):
        self.name, self.lvl, self.c, self.cross = name, lvl, c, cross


def build_plan(cfg):
    """Execution order of xunet.py:497-531 as a list of (kind, spec).  kinds: 'res', 'attn', 'push', 'pop'."""
    plan, nres, nb = [], cfg.num_resolutions, cfg.num_res_blocks
    film_cols = [0] * nres

    def res(name, lvl, cin, f, resample=None, cat=False):
        s = ResSpec(name, lvl, cin, f, resample, cat)
        s.film_off = film_cols[lvl]
        film_cols[lvl] += 2 * f
        plan.append(("res", s))

    def xblock(pfx, lvl, cin, f, use_attn, cat=False):
        res(pfx + ".resnetblock", lvl, cin, f, None, cat)
        if use_attn:
            plan.append(("attn", AttnSpec(pfx + ".attnblock_self", lvl, f, False)))
            plan.append(("attn", AttnSpec(pfx + ".attnblock_cross", lvl, f, True)))

    plan.append(("push", None))
    for lvl in range(nres):
        for i in range(nb):
            xblock(f"xunetblocks.{lvl}.{i}", lvl, cfg.dim_in[lvl] if i == 0 else cfg.dim_out[lvl], cfg.dim_out[lvl], lvl in cfg.attn_resolutions)
            plan.append(("push", None))
        if lvl != nres - 1:
            res(f"xunetblocks.{lvl}.{nb}", lvl, cfg.dim_out[lvl], cfg.dim_out[lvl], "down")
            plan.append(("push", None))
    xblock("middle", nres - 1, cfg.dim_out[-1], cfg.dim_out[-1], nres in cfg.attn_resolutions)
    for lvl in reversed(range(nres)):
        for i in range(nb + 1):
            if i == 0:
                ph = cfg.dim_out[lvl + 1] if lvl + 1 < nres else cfg.dim_out[lvl]
                pe = cfg.dim_out[lvl]
            elif i == nb:
                ph, pe = cfg.dim_out[lvl], cfg.dim_in[lvl]
            else:
                ph, pe = cfg.dim_out[lvl], cfg.dim_out[lvl]
            plan.append(("pop", None))
            xblock(f"upsample.{lvl}.{i}", lvl, ph + pe, cfg.dim_out[lvl], lvl in cfg.attn_resolutions, cat=True)
        if lvl != 0:
            res(f"upsample.{lvl}.{nb + 1}", lvl, cfg.dim_out[lvl], cfg.dim_out[lvl], "up")
    return plan, film_cols


class Act:
    """an activation tensor [N, h, w, C] with its (lazily created) gradient buffer"""

    def __init__(self, t, key):
        self.t, self.key, self.has_grad = t, key, False
        self.sums = None     # GroupNorm (sum, sumsq) [N,32,2] of this tensor at C/32 channels per group, when its producer's epilogue took them


class Engine:
    def __init__(self, cfg, ops, B):
        self.cfg, self.ops, self.B, self.N = cfg, ops, B, 2 * B
        self.plan, self.film_cols = build_plan(cfg)
        self._bufs = {}
        self._zero = {"fwd": [None, 0, {}], "bwd": [None, 0, {}]}
        self._zero_cap = {"fwd": 1 << 22, "bwd": 1 << 22}
        self.saved = None
        self._pack_cache = {}
        self._unpack_jl, self._copy_jl, self._unpack_sig, self._unpack_rec = None, None, None, False
        # weight-gradient launches never block the data-gradient chain (see side_tmp): their inputs live in block-private buffers
        # and the side streams join at the end of a phase only.  Costs memory (+5 GB at batch 128 of CFG-64, 63.5 GB in all;
        # measured 117.86 -> 117.22 ms per step there, more for small batches), so it stops where the private buffers would
        # outgrow that: 2^27 elements per level-0 activation tensor
        self.defer_side = 2 * B * cfg.H * cfg.W * cfg.ch <= 64 * 64 * 256 * 128
        if os.environ.get("B3D_DEFER_SIDE") is not None:      # measurement switch
            self.defer_side = os.environ["B3D_DEFER_SIDE"] == "1"
        # ... and then the side work is itself a serial chain of small launches as long as the main one: spread its regions
        # (one `with ops.side():` each, independent of each other) over several streams
        if self.defer_side and hasattr(self.ops, "n_side"):
            self.ops.n_side = max(self.ops.n_side, int(os.environ.get("B3D_SIDE_STREAMS", "3")))      # measured: 1 / 2 / 3 / 4 streams
            # = 20.80 / 20.36 / 20.29 / 20.34 ms per step at 16 pairs, 116.80 / 116.4 / 115.8 / - at 128
        # GroupNorm statistics of conv1's output from its own epilogue (measured in-run: +0.5% at batch 128, -0.4% at batch 16,
        # where the extra epilogue work costs more than the small gn_stats launch: large batches only; B3D_NO_FUSE_STATS=1 = off)
        self.fuse_stats = os.environ.get("B3D_NO_FUSE_STATS") != "1" and 2 * B * cfg.H * cfg.W > 64 * 64 * 64
        # attention backward: recompute P from the saved row log-sum-exp (no L x L tensor in HBM, 134 MB less per 16x16 block at
        # batch 128) instead of reading back the P the forward wrote; B3D_ATTN_SAVE_P=1 selects the latter (measurement switch)
        self.attn_recompute = os.environ.get("B3D_ATTN_SAVE_P") != "1"
        self.seed_ptr = None      # optional int64 device scalar added to every dropout seed (see ops.gn_apply)
        for lvl in range(cfg.num_resolutions):
            assert cfg.H % (2 ** lvl) == 0 and cfg.W % (2 ** lvl) == 0

    # ------------------------------------------------------------------ buffers
    def buf(self, key, shape, dtype=None):
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape):
            t = self.ops.empty(tuple(shape), dtype)
            self._bufs[key] = t
        return t

    def tmp(self, name, shape, dtype=None):
        """scratch reused by every block with the same shape (stream order makes the reuse safe)"""
        return self.buf(f"tmp.{name}:{tuple(shape)}", shape, dtype)

    def side_tmp(self, block, name, shape, dtype=None):
        """scratch that side-stream (weight-gradient) launches read.  Shared by every block of the same shape when the block
        joins the side stream at its end; private to the block when the join is deferred to the end of the phase (small
        batches: the main stream never waits for a weight gradient, at the price of one buffer per block)."""
        if self.defer_side:
            return self.buf(f"{block}.side.{name}", shape, dtype)
        return self.tmp(name, shape, dtype)

    def _block_join(self):
        if not self.defer_side:
            self.ops.join_side()

    def _gemm_stats(self, d, sums, C, rows_per_image=0):
        """launch d; if its epilogue can also take the GroupNorm statistics of the output (32 groups over C channels), let it
        (sums, zeroed per forward, is accumulated) and return True -- the consumer then skips its gn_stats pass"""
        d.gn_cpg, d.gn_rows_per_image = C // 32, rows_per_image
        ok = self.fuse_stats and C % 32 == 0 and self.ops.gemm_stats_supported(d)
        if ok:
            d.gn_sums = sums
        else:
            d.gn_cpg = d.gn_rows_per_image = 0
        self.ops.gemm(d)
        return ok

    def _producer_stats(self, d, act, C, rows_per_image=0):
        """launch the GEMM that produces `act`; when its epilogue can, it also leaves the GroupNorm statistics of the output in
        act.sums, so that a single-source GroupNorm reading `act` skips its statistics pass"""
        sums = self.zbuf("fwd", act.key + ".osum", (self.N, 32, 2))
        if self._gemm_stats(d, sums, C, rows_per_image):
            act.sums = sums

    def zbuf(self, phase, key, shape):
        """fp32 scratch that must be zero at the start of the phase: carved from one arena, one memset per phase"""
        arena, used, table = self._zero[phase]
        if arena is None:
            arena = self.ops.zeros((self._zero_cap[phase],), torch.float32)
            self._zero[phase][0] = arena
        if key not in table:
            n = int(math.prod(shape))
            n_al = (n + 63) // 64 * 64
            assert used + n_al <= self._zero_cap[phase], "zero arena exhausted"
            table[key] = (used, n)
            self._zero[phase][1] = used + n_al
        off, n = table[key]
        return arena[off:off + n].view(shape)

    def _zero_phase(self, phase):
        arena, used, _ = self._zero[phase]
        if arena is not None and used:
            arena[:used].zero_()

    def lvl_hw(self, lvl):
        return self.cfg.H >> lvl, self.cfg.W >> lvl

    # ------------------------------------------------------------------ weights
    def _pack_weights(self, P, training, repack=True, overlap=False):
        """fp32 PyTorch-layout parameters -> bf16 K-major operands (per forward: parameters change every step).
        The ~450 conversions are recorded once (all addresses are static) and then run as TWO kernel launches per step: the
        operands the forward needs first (logsnr MLP, level-0 conditioning conv and FiLM, stem, level-0 down blocks), and all
        the others.  overlap=True puts the second, large launch (HBM-bound, ~1 GB) on the side stream, beside the tensor-bound
        level-0 work; the caller joins the side stream before the first use of any other operand."""
        sig = (training, len(P), tuple(P[k].data_ptr() for k in list(P.keys())[:8]))
        cached = self._pack_cache.get(training)
        if cached is None or cached[0] != sig:
            cached = (sig,) + self._record_pack(P, training) + (P,)
            self._pack_cache[training] = cached
            repack = True
        if repack:
            first, rest = cached[2]
            self.ops.run_jobs(first)
            with (self.ops.side() if overlap else contextlib.nullcontext()):
                self.ops.run_jobs(rest)
                for specs in cached[4]:
                    self._compose(P, cached[1], specs, training)
        return cached[1]

    @staticmethod
    def _needed_first(name):
        cp = "conditioningprocessor."
        return name.startswith((cp + "logsnr_emb_emb", cp + "convs.0.", "conv.", "xunetblocks.0.", "film0"))

    def _record_pack(self, P, training):
        """Every job is recorded twice: in the (first, rest) lists a forward runs, and in the list of its parameter's group --
        0 / 1 / 2 = the backward phase that finalises the parameter's gradient (PHASE_PREFIXES), 3 = the level-wide FiLM
        operands, which phase 2's data-gradient GEMM still reads.  A trainer repacks group g right behind the Adam step of
        that range, beside the later backward phases (pack_group), and runs its forwards with repack=False."""
        ops, cfg, W = self.ops, self.cfg, {}
        jl_first, jl_rest = JobList("pack"), JobList("pack")
        groups = [JobList("pack") for _ in range(4)]
        compose = [[] for _ in range(4)]       # attention blocks whose composed output projection follows group g's repack
        pick = lambda name: jl_first if self._needed_first(name) else jl_rest      # (backward-only operands always go to jl_rest)

        def group_of(name):
            if name.startswith("film"):
                return 3
            hits = [ph for ph, pre in enumerate(self.PHASE_PREFIXES) if name.startswith(pre)]
            assert len(hits) == 1, name
            return hits[0]

        def rec(name, jl, *a, **k):
            ops.record_pack_weight(jl, *a, **k)
            ops.record_pack_weight(groups[group_of(name)], *a, **k)

        def rec_copy(name, jl, src, dst):
            ops.record_copy(jl, src, dst)
            ops.record_copy(groups[group_of(name)], src, dst)

        def conv3(name, w, Co_alloc=None):
            Co, Ci = w.shape[0], w.shape[1]
            cp = G.ceil64(Ci)
            f = self.buf("w.f." + name, (Co_alloc or Co, 9 * cp))
            if Co_alloc:
                f.zero_()
            rec(name, pick(name), w, 0, f, Cpad=cp)
            W[name] = f
            if training:
                cop = G.ceil64(Co)
                t = self.buf("w.t." + name, (Ci, 9 * cop))
                rec(name, jl_rest, w, 1, t, Cpad=cop)
                W[name + ".T"] = t

        def lin(name, w, srcs=None, need_t=None):
            need_t = training if need_t is None else need_t
            Co, Ci = w.shape[0], w.shape[1]
            srcs = srcs or [Ci]
            f = self.buf("w.f." + name, (Co, sum(G.ceil64(c) for c in srcs)))
            ci0 = koff = 0
            for c in srcs:
                rec(name, pick(name), w, 0, f, ci0=ci0, cn=c, Cpad=G.ceil64(c), k_off=koff)
                ci0 += c
                koff += G.ceil64(c)
            W[name] = f
            if need_t:
                t = self.buf("w.t." + name, (Ci, G.ceil64(Co)))
                rec(name, jl_rest, w, 1, t, Cpad=G.ceil64(Co))
                W[name + ".T"] = t

        cp = "conditioningprocessor."
        lin(cp + "logsnr_emb_emb.0.weight", P[cp + "logsnr_emb_emb.0.weight"])
        lin(cp + "logsnr_emb_emb.2.weight", P[cp + "logsnr_emb_emb.2.weight"])
        for l in range(cfg.num_resolutions):
            w = P[cp + f"convs.{l}.weight"]
            f = self.buf(f"w.f.cc{l}", (cfg.emb_ch, 9 * 192))
            rec(cp, pick(cp + f"convs.{l}."), w, 0, f, Cpad=192)
            W[cp + f"convs.{l}.weight"] = f
            if training:
                t = self.buf(f"w.r.cc{l}", (9 * 144, cfg.emb_ch))
                rec(cp, jl_rest, w, 2, t)
                W[cp + f"convs.{l}.weight.R"] = t
        # stem: the packed input tensor has 8 channels (3 real)
        f = self.buf("w.f.stem", (cfg.ch, 9 * 64))
        rec("conv.", jl_first, P["conv.weight"], 0, f, Cpad=64)
        W["conv.weight"] = f
        conv3("lastconv.weight", P["lastconv.weight"], Co_alloc=16)
        lb = self.buf("lastconv.b16", (16,), torch.float32)
        lb.zero_()
        rec_copy("lastconv.", jl_rest, P["lastconv.bias"], lb[:3])
        # FiLM: one [sum 2F, 1024] operand per level (+ its transpose for the data gradient)
        for l in range(cfg.num_resolutions):
            W[f"film{l}"] = self.buf(f"w.f.film{l}", (self.film_cols[l], 1024))
            if f"b.film{l}" not in self._bufs:
                self._bufs[f"b.film{l}"] = self.ops.zeros((self.film_cols[l],), torch.float32)
            if training:
                W[f"film{l}.T"] = self.buf(f"w.t.film{l}", (1024, self.film_cols[l]))
        for kind, s in self.plan:
            if kind == "res":
                n = s.name
                conv3(n + ".conv1.weight", P[n + ".conv1.weight"])
                conv3(n + ".conv2.weight", P[n + ".conv2.weight"])
                if s.cin != s.f:
                    lin(n + ".dense.weight", P[n + ".dense.weight"], s.srcs_c if s.cat else None)
                    # conv2 and the 1x1 skip as ONE implicit GEMM (forward): [conv2 taps | dense per source] along K
                    srcs_c, cpF = (s.srcs_c if s.cat else [s.cin]), G.ceil64(s.f)
                    f = self.buf("w.f." + n + ".conv2+dense", (s.f, 9 * cpF + sum(G.ceil64(c) for c in srcs_c)))
                    rec(n, pick(n), P[n + ".conv2.weight"], 0, f, Cpad=cpF)
                    ci0, koff = 0, 9 * cpF
                    for c in srcs_c:
                        rec(n, pick(n), P[n + ".dense.weight"], 0, f, ci0=ci0, cn=c, Cpad=G.ceil64(c), k_off=koff)
                        ci0 += c
                        koff += G.ceil64(c)
                    W[n + ".conv2+dense"] = f
                fw = P[n + ".film.dense.weight"]
                fl = f"film{s.lvl}"
                rec(fl, pick(fl), fw, 0, W[fl][s.film_off:s.film_off + 2 * s.f], Cpad=1024)
                rec_copy(fl, pick(fl), P[n + ".film.dense.bias"], self._bufs[f"b.film{s.lvl}"][s.film_off:s.film_off + 2 * s.f])
                if training:
                    rec(fl, jl_rest, fw, 1, W[fl + ".T"], Cpad=2 * s.f, k_off=s.film_off)
            elif kind == "attn":
                a = s.name + ".attn_layer.attn"
                lin(a + ".in_proj_weight", P[a + ".in_proj_weight"])
                # out_proj and the block's `linear` follow each other with nothing in between (xunet.py:211-219): the engine
                # applies them as ONE layer, W_m = W_lin W_out, b_m = W_lin b_out + b_lin, composed from the packed operands
                # after every repack (_compose); the backward maps the gradient of W_m back onto both (see _attn_bwd)
                lin(a + ".out_proj.weight", P[a + ".out_proj.weight"], need_t=True)
                lin(s.name + ".linear.weight", P[s.name + ".linear.weight"], need_t=False)
                C = s.c
                assert C % 64 == 0
                W[s.name + ".merged"] = self.buf("w.f." + s.name + ".merged", (C, C))
                if training:
                    W[s.name + ".merged.T"] = self.buf("w.t." + s.name + ".merged", (C, C))
                self.buf(s.name + ".bm", (C,), torch.float32)
                compose[group_of(s.name)].append(s)
        return W, (jl_first, jl_rest), groups, compose

    def _compose(self, P, W, specs, training):
        """W_m = W_lin W_out (bf16, from the packed bf16 operands, fp32 accumulation), its transpose for the data gradient, and
        b_m = W_lin b_out + b_lin (fp32) for the attention blocks in `specs`: three small launches per block."""
        ops = self.ops
        for s in specs:
            n, a, C = s.name, s.name + ".attn_layer.attn", s.c
            wl, wo, wot = W[n + ".linear.weight"], W[a + ".out_proj.weight"], W[a + ".out_proj.weight.T"]
            ops.gemm(G.dense_desc(wl, wot, W[n + ".merged"], C))                   # [i, j] = sum_k W_lin[i, k] W_out[k, j]
            if training:
                ops.gemm(G.dense_desc(wot, wl, W[n + ".merged.T"], C))             # [j, i]
            ops.matvec(P[n + ".linear.weight"].view(C, C), P[a + ".out_proj.bias"], self._bufs[n + ".bm"], add=P[n + ".linear.bias"])

    def pack_group(self, g, training=True):
        """Repack the operands of parameter group g (0 / 1 / 2 = backward phase, 3 = FiLM) on the current stream; needs a
        forward that has recorded the jobs (see _record_pack)."""
        cached = self._pack_cache.get(training)
        if cached is None:
            from ._lib import B3DError
            raise B3DError("pack_group before the first forward of this engine")
        self.ops.run_jobs(cached[3][g])
        self._compose(cached[5], cached[1], cached[4][g], training)

    # ------------------------------------------------------------------ forward
    def cond_features(self, P, R, t, K, cond_mask, repack=True):
        """The step-invariant part of the conditioning path (xunet.py:311-349 without `+ logsnr_emb`): conv_l(pose_emb) + bias
        per level, fp32 [2B, h_l, w_l, emb_ch], for the pairs of this engine's batch.  It depends on K / R / t / cond_mask
        only, so a sampler evaluates it once per (camera, frame role) and reuses it for all denoising steps.
        The tensors are this engine's buffers: copy what must outlive the next call."""
        ops, cfg, N = self.ops, self.cfg, self.N
        self._assign_cat_sources()
        Wp = self._pack_weights(P, False, repack)
        cp = "conditioningprocessor."
        pe = self.buf("pose_emb", (N, cfg.H, cfg.W, 144))
        ops.pose_embedding(K, R, t, cond_mask, P.get(cp + "pos_emb") if cfg.use_pos_emb else None,
                           P[cp + "first_emb"].reshape(-1) if cfg.use_ref_pose_emb else None,
                           P[cp + "other_emb"].reshape(-1) if cfg.use_ref_pose_emb else None, pe)
        out = []
        for l in range(cfg.num_resolutions):
            h, w = self.lvl_hw(l)
            c = self.buf(f"cond_c{l}", (N, h, w, cfg.emb_ch), torch.float32)
            ops.gemm(G.conv_desc([pe], [9], Wp[cp + f"convs.{l}.weight"], cfg.emb_ch, c, 2 ** l, bias=P[cp + f"convs.{l}.bias"]))
            out.append(c)
        return out

    def forward(self, P, x, z, logsnr, R, t, K, cond_mask, training=False, need_bwd=None, seed=0, repack=True, cond=None):
        """P: dict name -> fp32 param.  x,z: fp32 [B,3,H,W]; logsnr [B,2]; R [B,2,3,3], t [B,2,3], K [B,3,3] fp64;
        cond_mask [B].  Returns eps_hat fp32 [B,3,H,W] (frame 1).  Keeps what backward needs in self.saved.
        cond (inference only): cached conditioning -- dict(c=[fp32 [D,h_l,w_l,E] per level] from cond_features, image=int32 [N]
        (image n uses conditioning image[n]), rep=int32 [D] (an image n that uses conditioning j: its logsnr embedding is
        taken)); R / t / K / cond_mask are then not read, and FiLM is evaluated for the D distinct conditionings only."""
        ops, cfg, B, N = self.ops, self.cfg, self.B, self.N
        H, Wd = cfg.H, cfg.W
        assert x.shape == (B, 3, H, Wd), (x.shape, (B, 3, H, Wd))
        self._assign_cat_sources()
        if self._zero["bwd"][0] is None:   # packed fp32 weight-gradient staging lives in the bwd arena
            self._zero_cap["bwd"] = int(sum(p.numel() for p in P.values()) * 1.4) + (1 << 23)
        self._zero_phase("fwd")
        need_bwd = training if need_bwd is None else need_bwd
        Wp = self._pack_weights(P, need_bwd, repack, overlap=True)      # repack=False: parameters unchanged since the last call (sampling)
        drop_p = float(cfg.dropout) if training else 0.0
        sv = self.saved = {"P": P, "W": Wp, "need_bwd": need_bwd, "drop_p": drop_p, "seed": seed, "acts": {}}
        cp = "conditioningprocessor."
        E = cfg.emb_ch

        # ---- logsnr embedding MLP (xunet.py:305-308)
        ls = self.buf("logsnr", (N,), torch.float32)
        ls.copy_(logsnr.reshape(-1).to(torch.float32))
        e0 = self.buf("e0", (N, E))
        ops.ddpm_embedding(ls, e0)
        h1 = self.buf("h1", (N, E))
        h1s = self.buf("h1s", (N, E))
        ops.gemm(G.dense_desc(e0, Wp[cp + "logsnr_emb_emb.0.weight"], h1, E, bias=P[cp + "logsnr_emb_emb.0.bias"], out2=h1s))
        lemb = self.buf("lemb", (N, E), torch.float32)
        ops.gemm(G.dense_desc(h1s, Wp[cp + "logsnr_emb_emb.2.weight"], lemb, E, bias=P[cp + "logsnr_emb_emb.2.bias"]))

        self._ss_image = None
        ss = []
        if cond is not None:
            assert not need_bwd, "cached conditioning is an inference path"
            self._ss_image = cond["image"]
            D = cond["rep"].numel()
            for l in range(cfg.num_resolutions):
                h, w = self.lvl_hw(l)
                sil = self.buf(f"emb_silu_c{l}", (D, h, w, E))
                s_l = self.buf(f"ss_c{l}", (D * h * w, self.film_cols[l]))
                with (ops.side() if l >= 1 else contextlib.nullcontext()):
                    ops.cond_silu(cond["c"][l], lemb, cond["rep"], sil)
                    ops.gemm(G.dense_desc(sil.view(D * h * w, E), Wp[f"film{l}"], s_l, self.film_cols[l], bias=self._bufs[f"b.film{l}"]))
                ss.append(s_l)

        # ---- pose embedding (K1) and the four strided conditioning convs (+ logsnr_emb, SiLU for FiLM)
        pe = self.buf("pose_emb", (N, H, Wd, 144)) if cond is None else None
        if cond is None:
            ops.pose_embedding(K, R, t, cond_mask, P.get(cp + "pos_emb") if cfg.use_pos_emb else None,
                               P[cp + "first_emb"].reshape(-1) if cfg.use_ref_pose_emb else None,
                               P[cp + "other_emb"].reshape(-1) if cfg.use_ref_pose_emb else None, pe)
        # Level 0 on the main stream (the trunk needs it first); the coarser levels on the side stream, beside the stem and
        # the level-0 blocks, joined where the trunk first reaches level 1.
        for l in range(cfg.num_resolutions if cond is None else 0):
            h, w = self.lvl_hw(l)
            pre = self.buf(f"emb_pre{l}", (N, h, w, E))
            sil = self.buf(f"emb_silu{l}", (N, h, w, E))
            s_l = self.buf(f"ss{l}", (N * h * w, self.film_cols[l]))
            with (ops.side() if l >= 1 else contextlib.nullcontext()):
                ops.gemm(G.conv_desc([pe], [9], Wp[cp + f"convs.{l}.weight"], E, pre, 2 ** l, bias=P[cp + f"convs.{l}.bias"],
                                     rowvec=lemb, out2=sil))
                ops.gemm(G.dense_desc(sil.view(N * h * w, E), Wp[f"film{l}"], s_l, self.film_cols[l], bias=self._bufs[f"b.film{l}"]))
            ss.append(s_l)

        # ---- stem (xunet.py:493-495)
        xin = self.buf("xin", (N, H, Wd, 8))
        ops.pack_input(x, z, xin)
        h = Act(self.buf("stem", (N, H, Wd, cfg.ch)), "stem")
        self._producer_stats(G.conv_desc([xin], [9], Wp["conv.weight"], cfg.ch, h.t, bias=P["conv.bias"]), h, cfg.ch)

        hs, skip, bi = [], None, 0
        for kind, s in self.plan:
            if kind == "push":
                hs.append(h)
            elif kind == "pop":
                skip = hs.pop()
            elif kind == "res":
                if s.lvl >= 1 or not self._needed_first(s.name):
                    ops.join_side()            # FiLM operands of the coarser levels, the second weight-pack launch (no-op once joined)
                h = self._res_fwd(s, h, skip if s.cat else None, ss[s.lvl], P, Wp, drop_p, seed * 1000 + bi)
                skip = None
                bi += 1
            else:
                ops.join_side()
                h = self._attn_fwd(s, h, P, Wp)
        assert not hs
        ops.join_side()

        # ---- head (xunet.py:535-536)
        hh, ww, C = h.t.shape[1], h.t.shape[2], h.t.shape[3]
        sums = self.zbuf("fwd", "lastgn.sums", (N, 32, 2))
        x3 = h.t.view(N, hh * ww, C)
        al = self.buf("lastgn.a", (N, hh * ww, C))
        ops.gn_stats_apply(x3, None, sums, P["lastgn.gn.weight"], P["lastgn.gn.bias"], 1, al)
        y = self.buf("lastconv.y", (N, H, Wd, 16))
        ops.gemm(G.conv_desc([al.view(N, H, Wd, C)], [9], Wp["lastconv.weight"], 3, y, bias=self._bufs["lastconv.b16"]))
        eps = self.buf("eps", (B, 3, H, Wd), torch.float32)
        ops.unpack_output(y, eps)
        sv["final"] = h
        sv["ss"] = ss
        return eps

    def _assign_cat_sources(self):
        """channel split (h, skip) of the up-path concat inputs (xunet.py:435-451)"""
        cfg, nres, nb = self.cfg, self.cfg.num_resolutions, self.cfg.num_res_blocks
        for kind, s in self.plan:
            if kind == "res" and s.cat:
                lvl = s.lvl
                i = int(s.name.split(".")[2])
                if i == 0:
                    ph = cfg.dim_out[lvl + 1] if lvl + 1 < nres else cfg.dim_out[lvl]
                elif i == nb:
                    ph = cfg.dim_out[lvl]
                else:
                    ph = cfg.dim_out[lvl]
                s.srcs_c = [ph, s.cin - ph]

    def _res_fwd(self, s, h_in, skip, ss_l, P, Wp, drop_p, seed):
        """ResnetBlock.forward, xunet.py:133-152."""
        ops, N, n = self.ops, self.N, s.name
        hh, ww = h_in.t.shape[1], h_in.t.shape[2]
        HW, F = hh * ww, s.f
        x0 = h_in.t.view(N, HW, -1)
        x1 = skip.t.view(N, HW, -1) if skip is not None else None
        a0 = self.buf(n + ".a0", (N, HW, s.cin))
        if skip is None and h_in.sums is not None:      # statistics already taken by the epilogue of the GEMM that produced h_in
            sums0 = h_in.sums
            ops.gn_apply(x0, None, sums0, P[n + ".groupnorm0.gn.weight"], P[n + ".groupnorm0.gn.bias"], 1, a0)
        else:
            sums0 = self.zbuf("fwd", n + ".s0", (N, 32, 2))
            ops.gn_stats_apply(x0, x1, sums0, P[n + ".groupnorm0.gn.weight"], P[n + ".groupnorm0.gn.bias"], 1, a0)
        c1 = self.buf(n + ".c1", (N, hh, ww, F))
        sums1 = self.zbuf("fwd", n + ".s1", (N, 32, 2))
        c13 = c1.view(N, HW, F)
        # conv1's epilogue also sums the GroupNorm statistics of its output when the tile geometry allows it
        fused = self._gemm_stats(G.conv_desc([a0.view(N, hh, ww, s.cin)], [9], Wp[n + ".conv1.weight"], F, c1, bias=P[n + ".conv1.bias"]),
                                 sums1, F)
        a1 = self.buf(n + ".a1", (N, HW, F))
        ssv = ss_l[:, s.film_off:s.film_off + 2 * F]
        (ops.gn_apply if fused else ops.gn_stats_apply)(c13, None, sums1, P[n + ".groupnorm1.gn.weight"], P[n + ".groupnorm1.gn.bias"], 2, a1, ss=ssv, F=F, drop_p=drop_p,
                     seed=seed, seed_ptr=self.seed_ptr, ss_image=self._ss_image)
        out = self.buf(n + ".out", (N, hh, ww, F))
        if s.cin != F:
            # the 1x1 `dense` skip (xunet.py:148-150) rides in conv2's K loop as extra 1-tap segments over h_in (and the skip
            # tensor): one launch, and its output never goes through HBM to come back as the residual.  Its bias enters as
            # a per-image vector with image stride 0.
            srcs = [a1.view(N, hh, ww, F), h_in.t] + ([skip.t] if skip is not None else [])
            d2 = G.conv_desc(srcs, [9] + [1] * (len(srcs) - 1), Wp[n + ".conv2+dense"], F, out, bias=P[n + ".conv2.bias"],
                             rowvec=P[n + ".dense.bias"].view(1, F).expand(N, F), out_scale=RSQRT2)
        else:
            assert skip is None
            d2 = G.conv_desc([a1.view(N, hh, ww, F)], [9], Wp[n + ".conv2.weight"], F, out, bias=P[n + ".conv2.bias"], resid=h_in.t,
                             out_scale=RSQRT2)
        o = Act(out, n + ".out")
        if s.resample is None:
            self._producer_stats(d2, o, F)        # (a resampled output is normalised after the resampling: no use for these)
        else:
            ops.gemm(d2)
        self.saved["acts"][n] = dict(h_in=h_in, skip=skip, sums0=sums0, a0=a0, c1=c1, sums1=sums1, a1=a1, out=o, seed=seed)
        if s.resample == "down":
            r = self.buf(n + ".rs", (N, hh // 2, ww // 2, F))
            ops.avgpool2(out, r, 0.25)
            o = Act(r, n + ".rs")
        elif s.resample == "up":
            r = self.buf(n + ".rs", (N, hh * 2, ww * 2, F))
            ops.upsample2(out, r, 1.0)
            o = Act(r, n + ".rs")
        self.saved["acts"][n]["final"] = o
        return o

    def _qkv_views(self, qkv, L, C):
        N, Hh = self.N, self.cfg.attn_heads
        d = C // Hh
        return [qkv.view(N, L, 3, Hh, d)[:, :, i].permute(0, 2, 1, 3) for i in range(3)]

    def _attn_fwd(self, s, h_in, P, Wp):
        """AttnBlock.forward, xunet.py:193-220 (nn.MultiheadAttention: in_proj -> softmax(QK^T/sqrt d)V -> out_proj)."""
        ops, N, n, C = self.ops, self.N, s.name, s.c
        hh, ww = h_in.t.shape[1], h_in.t.shape[2]
        L, Hh = hh * ww, self.cfg.attn_heads
        d = C // Hh
        a = n + ".attn_layer.attn"
        x3 = h_in.t.view(N, L, C)
        g = self.buf(n + ".g", (N, L, C))
        if h_in.sums is not None:
            sums = h_in.sums
            ops.gn_apply(x3, None, sums, P[n + ".groupnorm.gn.weight"], P[n + ".groupnorm.gn.bias"], 0, g)
        else:
            sums = self.zbuf("fwd", n + ".s", (N, 32, 2))
            ops.gn_stats_apply(x3, None, sums, P[n + ".groupnorm.gn.weight"], P[n + ".groupnorm.gn.bias"], 0, g)
        qkv = self.buf(n + ".qkv", (N, L, 3 * C))
        ops.gemm(G.dense_desc(g.view(N * L, C), Wp[a + ".in_proj_weight"], qkv.view(N * L, 3 * C), 3 * C, bias=P[a + ".in_proj_bias"]))
        q, k, v = self._qkv_views(qkv, L, C)
        Lp = (L + 7) // 8 * 8
        need_p = self.saved["need_bwd"]
        # flash-style where the fused backward covers the shape: the forward keeps one fp32 per query row and the backward
        # recomputes P; otherwise P itself is kept for the unfused backward
        lse = self.buf(n + ".lse", (N, Hh, L), torch.float32) if (need_p and self.attn_recompute and ops.attn_bwd_supported(L, d)) else None
        p = self.buf(n + ".p", (N, Hh, L, Lp))[..., :L] if ((need_p and lse is None) or L > 256) else None
        o = self.buf(n + ".o", (N, L, C))
        if L <= 256 and d % 16 == 0 and d <= 256:
            # fused: S and O stay in TMEM, P goes through shared memory (written out only for an unfused backward)
            ops.attn_fwd(qkv, Hh, s.cross, 1.0 / math.sqrt(d), o, probs=p if (need_p and lse is None) else None, lse=lse)
        else:
            logits = self.tmp("logits", (N, Hh, L, Lp), torch.float32)[..., :L]      # fp32: never rounded before the softmax
            ops.gemm(G.attn_qk_desc(q, k, logits, s.cross, 1.0 / math.sqrt(d)))
            ops.softmax(logits, p)
            ops.gemm(G.attn_pv_desc(p, v, o.view(N, L, Hh, d).permute(0, 2, 1, 3), s.cross))
        # out_proj and `linear` as the one composed layer (see _record_pack / _compose)
        out = self.buf(n + ".out", (N, hh, ww, C))
        oa = Act(out, n + ".out")
        self._producer_stats(G.dense_desc(o.view(N * L, C), Wp[n + ".merged"], out.view(N * L, C), C, bias=self._bufs[n + ".bm"],
                                          resid=h_in.t.view(N * L, C), out_scale=RSQRT2), oa, C, rows_per_image=L)
        self.saved["acts"][n] = dict(h_in=h_in, sums=sums, g=g, qkv=qkv, p=p, lse=lse, o=o, out=oa)
        return oa

    # ------------------------------------------------------------------ backward
    def _grad_of(self, act):
        return self.buf("d." + act.key, act.t.shape)

    # The backward pass runs in three phases whose parameter gradients are FINAL (unpacked, in PyTorch layout) when the
    # phase returns, in the reverse of the flat arena's order: 0 = head + up path (arena tail: upsample.*, lastgn, lastconv),
    # 1 = middle + down path (xunetblocks.*, middle.*), 2 = stem + conditioning (arena head).  A data-parallel trainer
    # all-reduces each range while the next phase computes (trainer.py).
    PHASE_PREFIXES = (("upsample.", "lastgn.", "lastconv."), ("xunetblocks.", "middle."), ("conditioningprocessor.", "conv."))

    def backward(self, Gr, deps, grad_flat=None, phase=None):
        """Gr: dict name -> fp32 grad tensor (PyTorch layout), overwritten.  deps: fp32 [B,3,H,W] = dL/d eps_hat.
        grad_flat: optional flat tensor that all of Gr are views of (zeroed with one memset).
        phase: None = whole backward; 0, 1, 2 = one phase (must be called in that order).  Phase 2 may be given as its two parts
        "2a" (the last reads of the level-wide FiLM operands: a trainer may repack those as soon as it has finished) and "2b"."""
        for ph in ((0, 1, 2) if phase is None else (phase,)):
            base, part = (2, {"2a": 0, "2b": 1}[ph]) if ph in ("2a", "2b") else (ph, None)
            self._cur_phase = base
            if base == 2:
                self._bwd_phase2(Gr, deps, grad_flat, part)
            else:
                (self._bwd_phase0, self._bwd_phase1)[base](Gr, deps, grad_flat)
            self.ops.join_side()
            if part != 0:
                self.ops.run_jobs(self._unpack_jl[base])     # packed fp32 weight gradients of this phase -> PyTorch layout, one launch
                self.ops.run_jobs(self._copy_jl[base])

    def _bwd_phase0(self, Gr, deps, grad_flat):
        ops, cfg, B, N = self.ops, self.cfg, self.B, self.N
        sv = self.saved
        assert sv is not None and sv["need_bwd"], "backward needs a forward run with need_bwd=True"
        P, Wp = sv["P"], sv["W"]
        H, Wd, E = cfg.H, cfg.W, cfg.emb_ch
        self._zero_phase("bwd")
        usig = (len(Gr), tuple(Gr[k].data_ptr() for k in list(Gr.keys())[:8]))
        self._unpack_rec = self._unpack_jl is None or self._unpack_sig != usig
        if self._unpack_rec:
            self._unpack_jl, self._copy_jl, self._unpack_sig = [JobList("unpack") for _ in range(3)], [JobList("pack") for _ in range(3)], usig
        if grad_flat is not None:
            grad_flat.zero_()
        else:
            for g in Gr.values():
                g.zero_()
        for a in self._all_acts():
            a.has_grad = False
        self._dss = [self.buf(f"dss{l}", (N * (H >> l) * (Wd >> l), self.film_cols[l])) for l in range(cfg.num_resolutions)]

        # ---- head
        h = sv["final"]
        C = h.t.shape[3]
        HW = H * Wd
        dy = self.buf("lastconv.dy", (N, H, Wd, 16))
        ops.pack_output_grad(deps, dy)
        tmpb = self.zbuf("bwd", "lastconv.db", (16,))
        ops.colsum(dy.view(N * HW, 16), N * HW, tmpb)
        Gr["lastconv.bias"].copy_(tmpb[:3])
        al = self._bufs["lastgn.a"]
        self._conv_wgrad("lastconv.weight", dy, al.view(N, H, Wd, C), Gr["lastconv.weight"], co_alloc=16)
        da = self.buf("lastgn.da", (N, H, Wd, C))
        ops.gemm(G.conv_desc([dy], [9], Wp["lastconv.weight.T"], C, da))
        dh = self._grad_of(h)
        ops.gn_bwd(h.t.view(N, HW, C), None, self.zbuf("fwd", "lastgn.sums", (N, 32, 2)), P["lastgn.gn.weight"], P["lastgn.gn.bias"], 1,
                   da.view(N, HW, C), Gr["lastgn.gn.weight"], Gr["lastgn.gn.bias"], self.zbuf("bwd", "lastgn.ws", (N, 32, 2)),
                   dh.view(N, HW, C))
        h.has_grad = True

        self._bwd_blocks(Gr, lambda name: name.startswith("upsample."))
        self._film_wgrad(Gr, lambda name: name.startswith("upsample."))

    def _bwd_blocks(self, Gr, select):
        sv = self.saved
        P, Wp, drop_p = sv["P"], sv["W"], sv["drop_p"]
        for kind, s in reversed(self.plan):
            if s is None or not select(s.name):
                continue
            if kind == "res":
                self._res_bwd(s, P, Wp, Gr, self._dss[s.lvl], drop_p)
            elif kind == "attn":
                self._attn_bwd(s, P, Wp, Gr)

    def _bwd_phase1(self, Gr, deps, grad_flat):
        self._bwd_blocks(Gr, lambda name: not name.startswith("upsample."))
        self._film_wgrad(Gr, lambda name: not name.startswith("upsample."))

    def _film_wgrad(self, Gr, select):
        """Weight gradients of FiLM's Linear(1024 -> 2F) (xunet.py:80,84) of all the selected blocks of a level as ONE GEMM
        dss[:, cols]^T silu(emb): the level's silu(emb) -- 2 GB at 64x64 for batch 128 -- is read once per (level, phase)
        instead of once per block (the per-block GEMMs were HBM-bound on exactly that operand).  The blocks of a phase own a
        contiguous column range of the level-wide dss (plan order: down path, then up path); the rows of the result are
        copied to the per-block parameter gradients by the phase's copy job list."""
        ops, N = self.ops, self.N
        for l in range(self.cfg.num_resolutions):
            specs = [s for kind, s in self.plan if kind == "res" and s.lvl == l and select(s.name)]
            if not specs:
                continue
            c0, c1 = min(s.film_off for s in specs), max(s.film_off + 2 * s.f for s in specs)
            assert sum(2 * s.f for s in specs) == c1 - c0, "FiLM columns of a phase are not contiguous"
            hh, ww = self.lvl_hw(l)
            M = N * hh * ww
            sil = self._bufs[f"emb_silu{l}"].view(M, 1024)
            gp = self.zbuf("bwd", f"gp.film{l}.{self._cur_phase}", (c1 - c0, 1024))
            with ops.side():
                ops.gemm(G.wgrad_dense_desc(self._dss[l][:, c0:c1], sil, gp))
            if self._unpack_rec:
                for s in specs:
                    ops.record_copy(self._copy_jl[self._cur_phase], gp[s.film_off - c0:s.film_off - c0 + 2 * s.f], Gr[s.name + ".film.dense.weight"])

    def _bwd_phase2(self, Gr, deps, grad_flat, part=None):
        ops, cfg, B, N = self.ops, self.cfg, self.B, self.N
        sv = self.saved
        P, Wp = sv["P"], sv["W"]
        H, Wd, E = cfg.H, cfg.W, cfg.emb_ch
        HW = H * Wd
        dss = self._dss
        cp = "conditioningprocessor."
        if part in (None, 0):
            # ---- stem
            stem = Act(self._bufs["stem"], "stem")
            dstem = self._grad_of(stem)
            xin = self._bufs["xin"]
            with ops.side():       # parameter gradients beside the data path (joined at the end of the part)
                ops.colsum(dstem.view(N * HW, cfg.ch), N * HW, Gr["conv.bias"])
                gp = self.zbuf("bwd", "stem.gp", (cfg.ch, 9 * 64))
                ops.gemm(G.wgrad_desc(dstem, xin, 9, gp, 0, 64))
                self._unpack(gp, 0, 64, Gr["conv.weight"])
            # ---- FiLM data gradient of every level: the gradient of the conditioning convs' outputs.  The GEMM multiplies by
            # silu'(pre) in its epilogue (pre arrives as the residual tile), so that gradient is written once
            for l in range(cfg.num_resolutions):
                hh, ww = self.lvl_hw(l)
                M = N * hh * ww
                de = self.buf(f"de{l}", (N, hh, ww, E))
                ops.gemm(G.dense_desc(dss[l], Wp[f"film{l}.T"], de.view(M, E), E, resid=self._bufs[f"emb_pre{l}"].view(M, E), act=2))
        if part == 0:
            return

        # ---- conditioning convs, learned embeddings, logsnr MLP
        dlemb = self.zbuf("bwd", "dlemb", (N, E))
        dP = self.buf("dP", (2, H, Wd, 144), torch.float32)
        pe = self._bufs["pose_emb"]
        for l in range(cfg.num_resolutions):
            hh, ww = self.lvl_hw(l)
            M = N * hh * ww
            de = self._bufs[f"de{l}"]
            # d(logsnr_emb) per image and the conv's bias gradient are column sums of de, one read
            ops.silu_bwd_colsum(de, None, None, hh * ww, dlemb, Gr[cp + f"convs.{l}.bias"])
            with ops.side():
                gp = self.zbuf("bwd", f"cc{l}.gp", (E, 9 * 192))
                ops.gemm(G.wgrad_desc(de, pe, 9, gp, 0, 192, 2 ** l))
                self._unpack(gp, 0, 192, Gr[cp + f"convs.{l}.weight"])
            if cfg.use_pos_emb or cfg.use_ref_pose_emb:
                des = self.buf(f"desum{l}", (2, hh, ww, E))
                ops.batch_frame_sum(de, des)
                Gm = self.buf(f"Gm{l}", (2 * hh * ww, 9 * 144), torch.float32)
                ops.gemm(G.dense_desc(des.view(2 * hh * ww, E), Wp[cp + f"convs.{l}.weight.R"], Gm, 9 * 144))
                ops.conv_tap_scatter(Gm, 144, hh, ww, 2 ** l, dP, l > 0)
        if cfg.use_pos_emb or cfg.use_ref_pose_emb:
            ops.pose_embedding_bwd(dP, Gr[cp + "pos_emb"] if cfg.use_pos_emb else None,
                                   Gr[cp + "first_emb"].view(-1) if cfg.use_ref_pose_emb else None,
                                   Gr[cp + "other_emb"].view(-1) if cfg.use_ref_pose_emb else None)
        dl = self.buf("dlemb16", (N, E))
        ops.cast(dlemb, dl)
        h1, h1s, e0 = self._bufs["h1"], self._bufs["h1s"], self._bufs["e0"]
        ops.colsum(dl, N, Gr[cp + "logsnr_emb_emb.2.bias"])
        ops.gemm(G.wgrad_dense_desc(dl, h1s, Gr[cp + "logsnr_emb_emb.2.weight"]))
        dh1 = self.buf("dh1", (N, E))
        ops.gemm(G.dense_desc(dl, Wp[cp + "logsnr_emb_emb.2.weight.T"], dh1, E))
        ops.silu_bwd(dh1, h1, dh1)
        ops.colsum(dh1, N, Gr[cp + "logsnr_emb_emb.0.bias"])
        ops.gemm(G.wgrad_dense_desc(dh1, e0, Gr[cp + "logsnr_emb_emb.0.weight"]))

    def _unpack(self, gp, k_off, cpad, dw, ci0=0, cn=None):
        if self._unpack_rec:
            self.ops.record_unpack_wgrad(self._unpack_jl[self._cur_phase], gp, k_off, cpad, dw, ci0=ci0, cn=cn)

    def _all_acts(self):
        sv = self.saved
        out = [sv["final"]]
        for d in sv["acts"].values():
            for k in ("h_in", "skip", "out", "final"):
                a = d.get(k)
                if a is not None:
                    out.append(a)
        return out

    def _conv_wgrad(self, wname, dy, x, gw, co_alloc=None, stride=1, alpha=1.0):
        """3x3 weight gradient: packed fp32 [Co, 9*Cpad] through split-K atomics, then back to [Co,Ci,3,3]."""
        Co, Ci = gw.shape[0], gw.shape[1]
        cp = G.ceil64(Ci)
        gp = self.zbuf("bwd", "gp." + wname, (co_alloc or Co, 9 * cp))
        self.ops.gemm(G.wgrad_desc(dy, x, 9, gp, 0, cp, stride, alpha=alpha))
        self._unpack(gp, 0, cp, gw)

    def _take_grad(self, act):
        """(buffer, existing) for writing a new gradient contribution of `act`"""
        g = self._grad_of(act)
        had = act.has_grad
        act.has_grad = True
        return g, (g if had else None)

    def _res_bwd(self, s, P, Wp, Gr, dss_l, drop_p):
        ops, N, n = self.ops, self.N, s.name
        sa = self.saved["acts"][n]
        h_in, skip, out = sa["h_in"], sa["skip"], sa["out"]
        hh, ww = out.t.shape[1], out.t.shape[2]
        HW, F, M = hh * ww, s.f, N * hh * ww
        # adjoint of the resampling (xunet.py:17-28)
        if s.resample == "down":
            dout = self._grad_of(out)
            ops.upsample2(self._grad_of(sa["final"]), dout, 0.25)
        elif s.resample == "up":
            dout = self._grad_of(out)
            ops.avgpool2(self._grad_of(sa["final"]), dout, 1.0)
        else:
            dout = self._grad_of(out)
        # The block's output was (h + h_in) / sqrt(2) (xunet.py:152): that factor is the alpha of every GEMM that reads dout and
        # the scale of the skip-path addend below -- a scaled copy of dout (one more pass over HBM, one more launch on the
        # data-gradient chain) is never made.  Bias gradients = column sums of dout, beside the chain.
        dout2 = dout.view(M, F)
        # weight gradients of conv2 and the 1x1 skip, d(conv2.bias): on the side stream, beside the data-gradient chain below
        # (they only read dout and saved activations)
        with ops.side():
            ops.colsum(dout2, M, Gr[n + ".conv2.bias"], scale=RSQRT2)
            self._conv_wgrad(n + ".conv2.weight", dout, sa["a1"].view(N, hh, ww, F), Gr[n + ".conv2.weight"], alpha=RSQRT2)
            if s.cin != F:
                gw = Gr[n + ".dense.weight"].view(F, s.cin)
                srcs = [h_in] + ([skip] if skip is not None else [])
                koff = 0
                if len(srcs) == 1:
                    ops.gemm(G.wgrad_dense_desc(dout2, h_in.t.view(M, s.cin), gw, alpha=RSQRT2))
                else:
                    gp = self.zbuf("bwd", "gp." + n + ".dense", (F, sum(G.ceil64(a.t.shape[3]) for a in srcs)))
                    ci0 = 0
                    for a in srcs:
                        c = a.t.shape[3]
                        ops.gemm(G.wgrad_desc(dout, a.t, 1, gp, koff, G.ceil64(c), alpha=RSQRT2))
                        self._unpack(gp, koff, G.ceil64(c), Gr[n + ".dense.weight"], ci0=ci0, cn=c)
                        koff += G.ceil64(c)
                        ci0 += c
        # conv2 data gradient
        da1 = self.tmp("da1", (N, hh, ww, F))
        ops.gemm(G.conv_desc([dout], [9], Wp[n + ".conv2.weight.T"], F, da1, alpha=RSQRT2))
        # 1x1 skip
        if s.cin != F:
            if self._unpack_rec:       # d(dense.bias) = the same column sums as d(conv2.bias): copied at the end of backward
                ops.record_copy(self._copy_jl[self._cur_phase], Gr[n + ".conv2.bias"], Gr[n + ".dense.bias"])
            dskip = self.tmp("dskip", (N, HW, s.cin))
            ops.gemm(G.dense_desc(dout2, Wp[n + ".dense.weight.T"], dskip.view(M, s.cin), s.cin, alpha=RSQRT2))
            add0, add0_scale = dskip, 1.0
        else:
            add0, add0_scale = dout.view(N, HW, F), RSQRT2
        # GN1 + FiLM (+dropout)
        ssv = self.saved["ss"][s.lvl][:, s.film_off:s.film_off + 2 * F]
        dssv = dss_l[:, s.film_off:s.film_off + 2 * F]
        dc1 = self.side_tmp(n, "dc1", (N, hh, ww, F))
        ops.gn_bwd(sa["c1"].view(N, HW, F), None, sa["sums1"], P[n + ".groupnorm1.gn.weight"], P[n + ".groupnorm1.gn.bias"], 2,
                   da1.view(N, HW, F), Gr[n + ".groupnorm1.gn.weight"], Gr[n + ".groupnorm1.gn.bias"],
                   self.zbuf("bwd", n + ".ws1", (N, 32, 2)), dc1.view(N, HW, F), ss=ssv, F=F, dss=dssv, drop_p=drop_p, seed=sa["seed"],
                   seed_ptr=self.seed_ptr, dss_sum=Gr[n + ".film.dense.bias"], dx_sum=Gr[n + ".conv1.bias"])
        # (the FiLM dense weight gradients, xunet.py:80,84, are taken per level and phase in _film_wgrad)
        with ops.side():
            # conv1 (its bias gradient and the FiLM bias gradient were reduced inside the GroupNorm backward above)
            self._conv_wgrad(n + ".conv1.weight", dc1, sa["a0"].view(N, hh, ww, s.cin), Gr[n + ".conv1.weight"])
        da0 = self.tmp("da0", (N, hh, ww, s.cin))
        ops.gemm(G.conv_desc([dc1], [9], Wp[n + ".conv1.weight.T"], s.cin, da0))
        # GN0 + SiLU, merged with the skip-path gradient and with earlier contributions to the inputs
        g0, e0 = self._take_grad(h_in)
        x0 = h_in.t.view(N, HW, -1)
        if skip is not None:
            g1, e1 = self._take_grad(skip)
            x1 = skip.t.view(N, HW, -1)
        else:
            g1 = e1 = x1 = None
        v3 = lambda t_: t_.view(N, HW, -1) if t_ is not None else None
        ops.gn_bwd(x0, x1, sa["sums0"], P[n + ".groupnorm0.gn.weight"], P[n + ".groupnorm0.gn.bias"], 1, da0.view(N, HW, s.cin),
                   Gr[n + ".groupnorm0.gn.weight"], Gr[n + ".groupnorm0.gn.bias"], self.zbuf("bwd", n + ".ws0", (N, 32, 2)), v3(g0), v3(g1),
                   add0=add0, add1a=v3(e0), add1b=v3(e1), add0_scale=add0_scale)
        self._block_join()

    def _attn_bwd(self, s, P, Wp, Gr):
        ops, N, n, C = self.ops, self.N, s.name, s.c
        sa = self.saved["acts"][n]
        h_in, out = sa["h_in"], sa["out"]
        hh, ww = out.t.shape[1], out.t.shape[2]
        L, Hh = hh * ww, self.cfg.attn_heads
        d, M = C // Hh, N * hh * ww
        a = n + ".attn_layer.attn"
        alpha = 1.0 / math.sqrt(d)
        dout = self._grad_of(out)
        gs = dout.view(M, C)      # the (h + h_in) / sqrt(2) of xunet.py:220 is folded into the consumers of dout (see _res_bwd)
        # the composed out_proj + `linear` layer y = o W_m^T + b_m, W_m = W_lin W_out: with Gm = dy^T o,
        #   d W_lin = Gm W_out^T + d b_lin b_out^T,  d W_out = W_lin^T Gm,  d b_lin = colsum(dy),  d b_out = W_lin^T d b_lin
        # (weight-space products of C x C matrices, beside the data-gradient chain; joined at the end of the block)
        with ops.side():
            ops.colsum(gs, M, Gr[n + ".linear.bias"], scale=RSQRT2)
            Gm = self.zbuf("bwd", n + ".Gm", (C, C))
            ops.gemm(G.wgrad_dense_desc(gs, sa["o"].view(M, C), Gm, alpha=RSQRT2))
            Gb = self.side_tmp(n, "Gb", (C, C))
            ops.cast(Gm, Gb)
            ops.gemm(G.dense_desc(Gb, Wp[a + ".out_proj.weight"], Gr[n + ".linear.weight"].view(C, C), C))
            ops.rank1_add(Gr[n + ".linear.weight"].view(C, C), Gr[n + ".linear.bias"], P[a + ".out_proj.bias"])
            ops.gemm(G.wgrad_dense_desc(Wp[n + ".linear.weight"], Gb, Gr[a + ".out_proj.weight"]))
            ops.matvec(P[n + ".linear.weight"].view(C, C), Gr[n + ".linear.bias"], Gr[a + ".out_proj.bias"], transpose=True)
        do = self.tmp("do", (N, L, C))
        ops.gemm(G.dense_desc(gs, Wp[n + ".merged.T"], do.view(M, C), C, alpha=RSQRT2))
        # attention core
        p = sa["p"]
        dqkv = self.side_tmp(n, "dqkv", (N, L, 3 * C))
        if ops.attn_bwd_supported(L, d):
            # fused: dP, dS and the dK / dV accumulation over the query tiles stay on the SM
            if sa["lse"] is not None:
                ops.attn_bwd(sa["qkv"], None, sa["o"], do, dqkv, Hh, s.cross, alpha, lse=sa["lse"])
            else:
                ops.attn_bwd(sa["qkv"], p, sa["o"], do, dqkv, Hh, s.cross, alpha)
        else:
            q, k, v = self._qkv_views(sa["qkv"], L, C)
            dov = do.view(N, L, Hh, d).permute(0, 2, 1, 3)
            Lp = (L + 7) // 8 * 8
            dpl = self.tmp("logits", (N, Hh, L, Lp), torch.float32)[..., :L]
            dp = self.tmp("ds", (N, Hh, L, Lp))[..., :L]
            ops.gemm(G.attn_dp_desc(dov, v, dpl, s.cross))
            ops.softmax_bwd(p, dpl, dp)
            dq, dk, dv = self._qkv_views(dqkv, L, C)
            ops.gemm(G.attn_dv_desc(p, dov, dv, s.cross))
            ops.gemm(G.attn_dq_desc(dp, k, dq, s.cross, alpha))
            ops.gemm(G.attn_dk_desc(dp, q, dk, s.cross, alpha))
        # in_proj
        dq2 = dqkv.view(M, 3 * C)
        with ops.side():
            ops.colsum(dq2, M, Gr[a + ".in_proj_bias"])
            ops.gemm(G.wgrad_dense_desc(dq2, sa["g"].view(M, C), Gr[a + ".in_proj_weight"]))
        dg = self.tmp("dg", (N, L, C))
        ops.gemm(G.dense_desc(dq2, Wp[a + ".in_proj_weight.T"], dg.view(M, C), C))
        # GroupNorm + residual branch
        g0, e0 = self._take_grad(h_in)
        ops.gn_bwd(h_in.t.view(N, L, C), None, sa["sums"], P[n + ".groupnorm.gn.weight"], P[n + ".groupnorm.gn.bias"], 0, dg,
                   Gr[n + ".groupnorm.gn.weight"], Gr[n + ".groupnorm.gn.bias"], self.zbuf("bwd", n + ".ws", (N, 32, 2)),
                   g0.view(N, L, C), None, add0=gs.view(N, L, C), add1a=e0.view(N, L, C) if e0 is not None else None, add0_scale=RSQRT2)
        self._block_join()
