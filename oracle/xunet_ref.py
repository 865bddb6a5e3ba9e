"""TEST INFRASTRUCTURE ONLY -- CPU/any-device fp32 restatement of the reference X-UNet forward as a
pure function of a state_dict.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
/ --impl reference legs may import this; the product package never does.

It is pinned against the reference's own code (executed under oracle/stubs in the build
container) by tests/golden/make_golden.py -> tests/golden/*.pt, and re-checked by
tests/test_oracle.py.  The one boundary that stays PARITY UNPINNED is the camera-ray arithmetic
(third-party `visu3d`, see oracle/rays.py).

Every function cites the reference lines it restates (paths relative to /root/reference).
Autograd works through it (plain torch ops), so it is also the gradient oracle."""
import math

import numpy as np
import torch
import torch.nn.functional as F

from oracle import rays as _rays


class Cfg:
    """XUNet class attributes, xunet.py:356-366, overridable like `XUNet(**kwargs)` (:368-369)."""

    def __init__(self, **kw):
        self.H = 128
        self.W = 128
        self.ch = 256
        self.ch_mult = (1, 2, 2, 4)
        self.emb_ch = 1024
        self.num_res_blocks = 3
        self.attn_resolutions = (2, 3, 4)
        self.attn_heads = 4
        self.dropout = 0.1
        self.use_pos_emb = True
        self.use_ref_pose_emb = True
        self.__dict__.update(kw)
        self.num_resolutions = len(self.ch_mult)
        self.dim_in = [self.ch] + [self.ch * m for m in self.ch_mult[:-1]]   # xunet.py:388
        self.dim_out = [self.ch * m for m in self.ch_mult]                  # xunet.py:389


def posenc_ddpm(t, emb_ch, max_time=1.0):
    """xunet.py:32-46 (without the in-place aliasing, Q5)."""
    t = t.float() * (1000.0 / max_time)
    half = emb_ch // 2
    f = torch.exp(torch.arange(half, dtype=torch.float32) * -(np.log(10000) / (half - 1))).to(t.device)
    e = t[..., None] * f
    return torch.cat([torch.sin(e), torch.cos(e)], -1)


def posenc_nerf(x, min_deg, max_deg):
    """xunet.py:49-59: [x, sin(x*2^d) (d-major, xyz-minor), sin(x*2^d + pi/2)]."""
    scales = torch.tensor([2.0 ** i for i in range(min_deg, max_deg)], dtype=torch.float32, device=x.device)
    xb = (x[..., None, :] * scales[:, None]).reshape(*x.shape[:-1], -1)
    emb = torch.sin(torch.cat([xb, xb + torch.pi / 2.0], -1))
    return torch.cat([x, emb], -1)


def pose_embedding(sd, cfg, R, t, K, cond_mask, H, W):
    """xunet.py:311-336 -> [B,2,144,H,W] fp32 (before the four strided convs)."""
    dev = sd["conv.weight"].device
    Kb = K.detach().cpu().numpy()[:, None]
    Kb = np.broadcast_to(Kb, R.shape)
    pos, dr = _rays.camera_rays(Kb, R.detach().cpu().numpy(), t.detach().cpu().numpy(), H, W)
    pe = torch.cat([posenc_nerf(torch.from_numpy(pos).to(dev), 0, 15),
                    posenc_nerf(torch.from_numpy(dr).to(dev), 0, 8)], -1)          # [B,2,H,W,144]
    m = cond_mask.to(dev).bool()[:, None, None, None, None]
    pe = torch.where(m, pe, torch.zeros_like(pe)).permute(0, 1, 4, 2, 3)
    if cfg.use_pos_emb:
        pe = pe + sd["conditioningprocessor.pos_emb"][None, None]
    if cfg.use_ref_pose_emb:
        pe = torch.cat([sd["conditioningprocessor.first_emb"], sd["conditioningprocessor.other_emb"]], 1) + pe
    return pe


def conditioning(sd, cfg, batch, cond_mask):
    """ConditioningProcessor.forward, xunet.py:301-352."""
    B, _, H, W = batch["x"].shape
    p = "conditioningprocessor."
    logsnr = torch.clip(batch["logsnr"].float(), -20, 20)
    e = posenc_ddpm(logsnr, cfg.emb_ch, 1.0)
    e = F.linear(e, sd[p + "logsnr_emb_emb.0.weight"], sd[p + "logsnr_emb_emb.0.bias"])
    logsnr_emb = F.linear(F.silu(e), sd[p + "logsnr_emb_emb.2.weight"], sd[p + "logsnr_emb_emb.2.bias"])
    pe = pose_embedding(sd, cfg, batch["R"], batch["t"], batch["K"], cond_mask, H, W)
    pe2 = pe.reshape(B * 2, 144, H, W)
    embs = []
    for i in range(cfg.num_resolutions):
        y = F.conv2d(pe2, sd[p + f"convs.{i}.weight"], sd[p + f"convs.{i}.bias"], stride=2 ** i, padding=1)
        embs.append(y.reshape(B, 2, *y.shape[1:]))
    return logsnr_emb, embs, pe


def group_norm(sd, pfx, h):
    """GroupNorm, xunet.py:61-71: 32 groups per frame, eps 1e-5."""
    B, Fr, C, H, W = h.shape
    return F.group_norm(h.reshape(B * Fr, C, H, W), 32, sd[pfx + ".gn.weight"], sd[pfx + ".gn.bias"], 1e-5).reshape(h.shape)


def film(sd, pfx, h, emb):
    """FiLM, xunet.py:74-87: per-pixel Linear(1024->2F) on silu(emb); h*(1+scale)+shift."""
    y = F.linear(F.silu(emb.permute(0, 1, 3, 4, 2)), sd[pfx + ".dense.weight"], sd[pfx + ".dense.bias"])
    y = y.permute(0, 1, 4, 2, 3)
    Fch = h.shape[2]
    return h * (1.0 + y[:, :, :Fch]) + y[:, :, Fch:]


def resnet_block(sd, pfx, h_in, emb, resample=None, dropout=0.0, training=False, taps=None):
    """ResnetBlock.forward, xunet.py:133-152 (no SiLU between FiLM and conv2, Q1)."""
    B, Fr, C, H, W = h_in.shape
    h = F.silu(group_norm(sd, pfx + ".groupnorm0", h_in))
    h = F.conv2d(h.reshape(B * Fr, C, H, W), sd[pfx + ".conv1.weight"], sd[pfx + ".conv1.bias"], padding=1)
    Fo = h.shape[1]
    h = group_norm(sd, pfx + ".groupnorm1", h.reshape(B, Fr, Fo, H, W))
    h = film(sd, pfx + ".film", h, emb)
    h = F.dropout(h, dropout, training)
    h = F.conv2d(h.reshape(B * Fr, Fo, H, W), sd[pfx + ".conv2.weight"], sd[pfx + ".conv2.bias"], padding=1)
    h = h.reshape(B, Fr, Fo, H, W)
    if C != Fo:
        h_in = F.conv2d(h_in.reshape(B * Fr, C, H, W), sd[pfx + ".dense.weight"], sd[pfx + ".dense.bias"]).reshape(B, Fr, Fo, H, W)
    out = (h + h_in) / np.sqrt(2)
    if resample == "down":      # avgpool_downsample, xunet.py:23-28
        out = F.avg_pool2d(out.reshape(B * Fr, Fo, H, W), 2, 2).reshape(B, Fr, Fo, H // 2, W // 2)
    elif resample == "up":      # nearest_neighbor_upsample, xunet.py:17-20
        out = out.repeat_interleave(2, -2).repeat_interleave(2, -1)
    if taps is not None:
        taps[pfx] = out
    return out


def mha(sd, pfx, q, kv, heads):
    """nn.MultiheadAttention(C, heads, batch_first=True)(q, kv, kv)[0], xunet.py:161,175.
    q,kv: [B,L,C]."""
    C = q.shape[-1]
    w, b = sd[pfx + ".in_proj_weight"], sd[pfx + ".in_proj_bias"]
    Q = F.linear(q, w[:C], b[:C])
    K_ = F.linear(kv, w[C:2 * C], b[C:2 * C])
    V = F.linear(kv, w[2 * C:], b[2 * C:])
    B, L, _ = Q.shape
    d = C // heads
    sp = lambda x: x.reshape(B, -1, heads, d).transpose(1, 2)
    Q, K_, V = sp(Q), sp(K_), sp(V)
    P = torch.softmax((Q @ K_.transpose(-1, -2)) / math.sqrt(d), -1)
    O = (P @ V).transpose(1, 2).reshape(B, L, C)
    return F.linear(O, sd[pfx + ".out_proj.weight"], sd[pfx + ".out_proj.bias"])


def attn_block(sd, pfx, h_in, attn_type, heads, taps=None):
    """AttnBlock.forward, xunet.py:193-220: self (q,kv)=(f,f); cross (0,1),(1,0), shared weights."""
    B, Fr, C, H, W = h_in.shape
    h = group_norm(sd, pfx + ".groupnorm", h_in)
    h0 = h[:, 0].reshape(B, C, H * W).transpose(1, 2)
    h1 = h[:, 1].reshape(B, C, H * W).transpose(1, 2)
    a = pfx + ".attn_layer.attn"
    if attn_type == "self":
        o0, o1 = mha(sd, a, h0, h0, heads), mha(sd, a, h1, h1, heads)
    elif attn_type == "cross":
        o0, o1 = mha(sd, a, h0, h1, heads), mha(sd, a, h1, h0, heads)
    else:
        raise NotImplementedError(attn_type)
    h = torch.stack([o0.transpose(1, 2), o1.transpose(1, 2)], 1).reshape(B * Fr, C, H, W)
    h = F.conv2d(h, sd[pfx + ".linear.weight"], sd[pfx + ".linear.bias"]).reshape(B, Fr, C, H, W)
    out = (h + h_in) / np.sqrt(2)
    if taps is not None:
        taps[pfx] = out
    return out


def xunet_block(sd, pfx, h, emb, use_attn, cfg, training, taps):
    """XUNetBlock.forward, xunet.py:246-256."""
    h = resnet_block(sd, pfx + ".resnetblock", h, emb, None, cfg.dropout, training, taps)
    if use_attn:
        h = attn_block(sd, pfx + ".attnblock_self", h, "self", cfg.attn_heads, taps)
        h = attn_block(sd, pfx + ".attnblock_cross", h, "cross", cfg.attn_heads, taps)
    return h


def xunet_forward(sd, cfg, batch, cond_mask, training=False, taps=None):
    """XUNet.forward, xunet.py:477-536.  `sd`: state_dict (no `module.` prefix) of fp32 tensors.
    `taps`: optional dict filled with per-block outputs ([B,2,C,h,w]) and 'pose_emb'."""
    B, _, H, W = batch["x"].shape
    assert (H, W) == (cfg.H, cfg.W)
    logsnr_emb, pose_embs, pe = conditioning(sd, cfg, batch, cond_mask)
    if taps is not None:
        taps["pose_emb"] = pe
        taps["logsnr_emb"] = logsnr_emb
    h = torch.stack([batch["x"], batch["z"]], 1).float().reshape(B * 2, 3, H, W)
    h = F.conv2d(h, sd["conv.weight"], sd["conv.bias"], padding=1).reshape(B, 2, cfg.ch, H, W)
    hs = [h]
    nres = cfg.num_resolutions
    for lvl in range(nres):
        emb = logsnr_emb[..., None, None] + pose_embs[lvl]
        for i in range(cfg.num_res_blocks):
            h = xunet_block(sd, f"xunetblocks.{lvl}.{i}", h, emb, lvl in cfg.attn_resolutions, cfg, training, taps)
            hs.append(h)
        if lvl != nres - 1:
            h = resnet_block(sd, f"xunetblocks.{lvl}.{cfg.num_res_blocks}", h, emb, "down", cfg.dropout, training, taps)
            hs.append(h)
    emb = logsnr_emb[..., None, None] + pose_embs[-1]
    h = xunet_block(sd, "middle", h, emb, nres in cfg.attn_resolutions, cfg, training, taps)
    for lvl in reversed(range(nres)):
        emb = logsnr_emb[..., None, None] + pose_embs[lvl]
        for i in range(cfg.num_res_blocks + 1):
            h = torch.cat([h, hs.pop()], 2)
            h = xunet_block(sd, f"upsample.{lvl}.{i}", h, emb, lvl in cfg.attn_resolutions, cfg, training, taps)
        if lvl != 0:
            h = resnet_block(sd, f"upsample.{lvl}.{cfg.num_res_blocks + 1}", h, emb, "up", cfg.dropout, training, taps)
    assert not hs
    h = F.silu(group_norm(sd, "lastgn", h))
    y = F.conv2d(h.reshape(B * 2, cfg.ch, H, W), sd["lastconv.weight"], sd["lastconv.bias"], padding=1)
    return y.reshape(B, 2, 3, H, W)[:, 1]


def strip_module_prefix(sd):
    """Checkpoints saved from DataParallel carry `module.` (sampling.py:52-56)."""
    return {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}


def rerandomize_zero_init(sd, seed=1234):
    """H1/Q2: conv2 / attn `linear` / lastconv weights are zero-initialised (xunet.py:131,191,474)
    which makes parity degenerate; overwrite them with U(+-1/sqrt(fan_in)) in place."""
    g = torch.Generator().manual_seed(seed)
    for k, v in sd.items():
        if k.endswith("conv2.weight") or k.endswith("linear.weight") or k == "lastconv.weight":
            fan_in = v[0].numel()
            bound = 1.0 / math.sqrt(fan_in)
            v.copy_((torch.rand(v.shape, generator=g) * 2 - 1) * bound)
    return sd


def param_shapes(cfg):
    """State-dict schema of `XUNet(**cfg)` (names + shapes), restated from the constructors at
    xunet.py:64-66,77-80,111-129,161,187-190,273-299,385,393-474.  make_golden.py asserts it equals
    the reference module's own state_dict()."""
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

    def gn(pfx, c):
        S[pfx + ".gn.weight"] = (c,)
        S[pfx + ".gn.bias"] = (c,)

    def resnet(pfx, cin, f):
        gn(pfx + ".groupnorm0", cin)
        gn(pfx + ".groupnorm1", f)
        S[pfx + ".conv1.weight"] = (f, cin, 3, 3)
        S[pfx + ".conv1.bias"] = (f,)
        S[pfx + ".film.dense.weight"] = (2 * f, 1024)      # FiLM emb_ch default, xunet.py:77,119
        S[pfx + ".film.dense.bias"] = (2 * f,)
        S[pfx + ".conv2.weight"] = (f, f, 3, 3)
        S[pfx + ".conv2.bias"] = (f,)
        if cin != f:
            S[pfx + ".dense.weight"] = (f, cin, 1, 1)
            S[pfx + ".dense.bias"] = (f,)

    def attn(pfx, c):
        gn(pfx + ".groupnorm", c)
        a = pfx + ".attn_layer.attn"
        S[a + ".in_proj_weight"] = (3 * c, c)
        S[a + ".in_proj_bias"] = (3 * c,)
        S[a + ".out_proj.weight"] = (c, c)
        S[a + ".out_proj.bias"] = (c,)
        S[pfx + ".linear.weight"] = (c, c, 1, 1)
        S[pfx + ".linear.bias"] = (c,)

    def xblock(pfx, cin, f, use_attn):
        resnet(pfx + ".resnetblock", cin, f)
        if use_attn:
            attn(pfx + ".attnblock_self", f)
            attn(pfx + ".attnblock_cross", f)

    nres, nb = cfg.num_resolutions, cfg.num_res_blocks
    for lvl in range(nres):
        for i in range(nb):
            xblock(f"xunetblocks.{lvl}.{i}", cfg.dim_in[lvl] if i == 0 else cfg.dim_out[lvl], cfg.dim_out[lvl],
                   lvl in cfg.attn_resolutions)
        if lvl != nres - 1:
            resnet(f"xunetblocks.{lvl}.{nb}", cfg.dim_out[lvl], cfg.dim_out[lvl])
    xblock("middle", cfg.dim_out[-1], cfg.dim_out[-1], nres in cfg.attn_resolutions)
    for lvl in reversed(range(nres)):
        for i in range(nb + 1):
            if i == 0:
                ph = cfg.dim_out[lvl + 1] if lvl + 1 < nres else cfg.dim_out[lvl]
                pe = cfg.dim_out[lvl]
            elif i == nb:
                ph, pe = cfg.dim_out[lvl], cfg.dim_in[lvl]
            else:
                ph, pe = cfg.dim_out[lvl], cfg.dim_out[lvl]
            xblock(f"upsample.{lvl}.{i}", ph + pe, cfg.dim_out[lvl], lvl in cfg.attn_resolutions)
        if lvl != 0:
            resnet(f"upsample.{lvl}.{nb + 1}", cfg.dim_out[lvl], cfg.dim_out[lvl])
    gn("lastgn", cfg.ch)
    S["lastconv.weight"] = (3, cfg.ch, 3, 3)
    S["lastconv.bias"] = (3,)
    return S


def init_state_dict(cfg, seed=0, dtype=torch.float32):
    """Deterministic NON-degenerate weights (no zero-init families, H1): used by fixtures, parity
    tests, smoke and bench so that no weights file has to travel.  Not the reference's init."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in param_shapes(cfg).items():
        if k.endswith(".gn.weight"):
            v = 1.0 + 0.1 * torch.randn(shp, generator=g)
        elif k.endswith("bias"):
            v = 0.05 * torch.randn(shp, generator=g)
        elif k.endswith("pos_emb") or k.endswith("first_emb") or k.endswith("other_emb"):
            v = torch.randn(shp, generator=g) / math.sqrt(144)
        else:
            fan_in = int(np.prod(shp[1:]))
            v = torch.randn(shp, generator=g) / math.sqrt(fan_in)
        sd[k] = v.to(dtype)
    return sd
