"""Drop-in for the reference `xunet.XUNet` (xunet.py:355-536): same constructor kwargs, same
state_dict names/shapes (647 tensors at ch=128), same `forward(batch, *, cond_mask) -> [B,3,H,W]`.
Underneath, forward and backward run on the b3d sm_100a kernels through engine.Engine; there is
no eager/PyTorch compute path and no CPU fallback."""
import math
from typing import Tuple

import torch

from . import _lib
from .engine import Cfg, Engine, build_plan


def param_shapes(cfg):
    """name -> shape, in the reference's registration order (xunet.py:273-299,385,393-474)."""
    S = {}
    p = "conditioningprocessor."
    if cfg.use_pos_emb:
        S[p + "pos_emb"] = (144, cfg.H, cfg.W)
    if cfg.use_ref_pose_emb:
        S[p + "first_emb"] = (1, 1, 144, 1, 1)
        S[p + "other_emb"] = (1, 1, 144, 1, 1)
    for i in (0, 2):
        S[p + f"logsnr_emb_emb.{i}.weight"] = (cfg.emb_ch, cfg.emb_ch)
        S[p + f"logsnr_emb_emb.{i}.bias"] = (cfg.emb_ch,)
    for i in range(cfg.num_resolutions):
        S[p + f"convs.{i}.weight"] = (cfg.emb_ch, 144, 3, 3)
        S[p + f"convs.{i}.bias"] = (cfg.emb_ch,)
    S["conv.weight"] = (cfg.ch, 3, 3, 3)
    S["conv.bias"] = (cfg.ch,)
    plan, _ = build_plan(cfg)
    for kind, s in plan:
        n = None if s is None else s.name
        if kind == "res":
            for g, c in (("groupnorm0", s.cin), ("groupnorm1", s.f)):
                S[f"{n}.{g}.gn.weight"] = (c,)
                S[f"{n}.{g}.gn.bias"] = (c,)
            S[n + ".conv1.weight"] = (s.f, s.cin, 3, 3)
            S[n + ".conv1.bias"] = (s.f,)
            S[n + ".film.dense.weight"] = (2 * s.f, 1024)       # FiLM's emb_ch default (xunet.py:77,119)
            S[n + ".film.dense.bias"] = (2 * s.f,)
            S[n + ".conv2.weight"] = (s.f, s.f, 3, 3)
            S[n + ".conv2.bias"] = (s.f,)
            if s.cin != s.f:
                S[n + ".dense.weight"] = (s.f, s.cin, 1, 1)
                S[n + ".dense.bias"] = (s.f,)
        elif kind == "attn":
            c = s.c
            S[n + ".groupnorm.gn.weight"] = (c,)
            S[n + ".groupnorm.gn.bias"] = (c,)
            a = n + ".attn_layer.attn"
            S[a + ".in_proj_weight"] = (3 * c, c)
            S[a + ".in_proj_bias"] = (3 * c,)
            S[a + ".out_proj.weight"] = (c, c)
            S[a + ".out_proj.bias"] = (c,)
            S[n + ".linear.weight"] = (c, c, 1, 1)
            S[n + ".linear.bias"] = (c,)
    S["lastgn.gn.weight"] = (cfg.ch,)
    S["lastgn.gn.bias"] = (cfg.ch,)
    S["lastconv.weight"] = (3, cfg.ch, 3, 3)
    S["lastconv.bias"] = (3,)
    return S


def _init_like_reference(name, t):
    """torch's default initialisers for the layer types the reference instantiates, plus its explicit
    zero / normal initialisations (xunet.py:131,191,282-290,474)."""
    with torch.no_grad():
        if name.endswith(".gn.weight"):
            t.fill_(1.0)
        elif name.endswith(".gn.bias") or name.endswith("in_proj_bias") or name.endswith("out_proj.bias"):
            t.zero_()
        elif name.endswith("pos_emb") or name.endswith("first_emb") or name.endswith("other_emb"):
            t.normal_(std=1.0 / math.sqrt(144))
        elif name.endswith("conv2.weight") or name.endswith(".linear.weight") or name == "lastconv.weight":
            t.zero_()
        elif name.endswith("in_proj_weight"):
            torch.nn.init.xavier_uniform_(t)
        elif name.endswith("weight"):
            torch.nn.init.kaiming_uniform_(t, a=math.sqrt(5))
        else:  # bias of conv / linear: U(+-1/sqrt(fan_in))
            fan_in = getattr(t, "_b3d_fan_in", 1)
            b = 1.0 / math.sqrt(fan_in) if fan_in > 0 else 0.0
            t.uniform_(-b, b)


class _Node(torch.nn.Module):
    """parameter container; the tree reproduces the reference's attribute paths"""


class _XUNetFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, z, logsnr, R, t, K, cond_mask, seed, need_bwd, *params):
        eng = model._engine(x.shape[0])
        P = dict(zip(model._names, params))
        eps = eng.forward(P, x, z, logsnr, R, t, K, cond_mask, training=model.training, need_bwd=need_bwd, seed=seed)
        # the activations a backward needs live in the engine (one set per batch size), not in ctx: remember which forward
        # they belong to, so a backward that arrives after another forward of the same batch size fails instead of silently
        # using the other call's activations
        eng.generation = getattr(eng, "generation", 0) + 1
        ctx.model, ctx.eng, ctx.generation = model, eng, eng.generation
        return eps.clone()

    @staticmethod
    def backward(ctx, deps):
        model, eng = ctx.model, ctx.eng
        if ctx.generation != eng.generation:
            raise RuntimeError("XUNet backward: another forward with the same batch size ran since this one; the engine keeps ONE set of "
                               "saved activations per batch size (run backward before the next forward, or use different batch sizes). "
                               "Gradients are also overwritten, not accumulated, by every backward.")
        Gr = model._grad_views()
        eng.backward(Gr, deps.contiguous().float(), grad_flat=model._grad_flat)
        if model.fused_grads:
            # training loops of this package: .grad of every parameter IS its slice of the flat arena
            for n, p in zip(model._names, model._plist):
                p.grad = Gr[n]
            return (None,) * 10 + (None,) * len(model._names)
        return (None,) * 10 + tuple(Gr[n] for n in model._names)


class XUNet(torch.nn.Module):
    H: int = 128
    W: int = 128
    ch: int = 256
    ch_mult: Tuple[int] = (1, 2, 2, 4)
    emb_ch: int = 1024
    num_res_blocks: int = 3
    attn_resolutions: Tuple[int] = (2, 3, 4)
    attn_heads: int = 4
    dropout: float = 0.1
    use_pos_emb: bool = True
    use_ref_pose_emb: bool = True

    def __init__(self, **kwargs):
        self.__dict__.update(kwargs)        # kwargs override class attributes, as in the reference (xunet.py:368-369)
        super().__init__()
        nlev = len(self.ch_mult)
        assert self.H % (2 ** (nlev - 1)) == 0, f"Size of the image must me multiple of {2 ** (nlev - 1)}"
        assert self.W % (2 ** (nlev - 1)) == 0, f"Size of the image must me multiple of {2 ** (nlev - 1)}"
        assert self.emb_ch == 1024, "FiLM hard-codes emb_ch=1024 in the reference (xunet.py:77,119)"
        self.num_resolutions = nlev
        self.cfg = Cfg(H=self.H, W=self.W, ch=self.ch, ch_mult=tuple(self.ch_mult), emb_ch=self.emb_ch,
                       num_res_blocks=self.num_res_blocks, attn_resolutions=tuple(self.attn_resolutions), attn_heads=self.attn_heads,
                       dropout=self.dropout, use_pos_emb=self.use_pos_emb, use_ref_pose_emb=self.use_ref_pose_emb)
        self.dim_in, self.dim_out = self.cfg.dim_in, self.cfg.dim_out
        shapes = param_shapes(self.cfg)
        self._names = list(shapes.keys())
        for name, shp in shapes.items():
            node = self
            parts = name.split(".")
            for a in parts[:-1]:
                if not hasattr(node, a):
                    node.add_module(a, _Node())
                node = getattr(node, a)
            t = torch.empty(shp)
            if name.endswith("bias") and not name.endswith(".gn.bias"):
                w = shapes.get(name[:-4] + "weight")
                t._b3d_fan_in = int(math.prod(w[1:])) if w else 0
            _init_like_reference(name, t)
            node.register_parameter(parts[-1], torch.nn.Parameter(t))
        self.fused_grads = False
        self._engines = {}
        self._ops = None
        self._flat = self._grad_flat = None
        self._step_seed = 0
        self._param_version = 0     # bumped whenever the parameters change through this package (samplers key their caches on it)
        self._flatten()

    # ---- one flat fp32 arena for parameters (and one for gradients): allreduce / Adam run on a single tensor
    def _flatten(self):
        ps = [dict(self.named_parameters())[n] for n in self._names]
        dev = ps[0].device
        if self._flat is None or self._flat.device != dev or any(p.data.untyped_storage().data_ptr() != self._flat.untyped_storage().data_ptr() for p in ps):
            offs, tot = [], 0
            for p in ps:
                offs.append(tot)
                tot += (p.numel() + 63) // 64 * 64
            flat = torch.zeros(tot, dtype=torch.float32, device=dev)
            for p, o in zip(ps, offs):
                v = flat[o:o + p.numel()].view(p.shape)
                v.copy_(p.data)
                p.data = v
            self._flat, self._offs = flat, offs
            self._grad_flat = None
            self._engines = {}
            self._ops = None
            self._param_version = getattr(self, "_param_version", 0) + 1
        self._plist = ps

    def _apply(self, fn, recurse=True):
        r = super()._apply(fn, recurse)
        self._flatten()
        return r

    def flat_parameters(self):
        return self._flat

    def flat_gradients(self):
        self._grad_views()
        return self._grad_flat

    def _grad_views(self):
        if self._grad_flat is None:
            self._grad_flat = torch.zeros_like(self._flat)
            self._gviews = {n: self._grad_flat[o:o + p.numel()].view(p.shape) for n, p, o in zip(self._names, self._plist, self._offs)}
        return self._gviews

    def _engine(self, B):
        dev = self._flat.device
        if dev.type != "cuda":
            raise _lib.B3DError("XUNet runs on the b3d CUDA kernels only: move the module to a CUDA device (no CPU fallback)")
        if self._ops is None:
            from .ops import CudaOps
            self._ops = CudaOps(dev)
        if B not in self._engines:
            self._engines[B] = Engine(self.cfg, self._ops, B)
        return self._engines[B]

    def forward(self, batch, *, cond_mask):
        x = batch["x"]
        B, C, H, W = x.shape
        for key, temp in batch.items():
            assert temp.shape[0] == B, f"{key} should have batch size of {B}, not {temp.shape[0]}"
        assert B == cond_mask.shape[0]
        assert (H, W) == (self.H, self.W), ((H, W), (self.H, self.W))
        dev = self._flat.device
        f32 = lambda t: t.to(device=dev, dtype=torch.float32)
        f64 = lambda t: t.to(device=dev, dtype=torch.float64)
        self._step_seed += 1
        need_bwd = torch.is_grad_enabled() and any(p.requires_grad for p in self._plist)
        return _XUNetFn.apply(self, f32(x).contiguous(), f32(batch["z"]).contiguous(), f32(batch["logsnr"]), f64(batch["R"]),
                              f64(batch["t"]), f64(batch["K"]), cond_mask.to(dev), self._step_seed, need_bwd, *self._plist)

    def load_state_dict(self, state_dict, strict=True, **kw):
        """accepts checkpoints saved from nn.DataParallel (`module.` prefix, sampling.py:52-56)"""
        sd = {(k[7:] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        self._param_version += 1
        return super().load_state_dict(sd, strict=strict, **kw)
