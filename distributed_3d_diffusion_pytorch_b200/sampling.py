"""Drop-in for the sampler functions of the reference's sampling.py:59-155 (stochastic-conditioning
ancestral sampler with classifier-free guidance), device-resident:

* the conditional and unconditional UNet calls of one denoise step (sampling.py:97-99) are ONE forward
  of batch 2b (rows are independent: GroupNorm is per image, attention is per pair);
* the guidance mix / x0 clip / posterior mean / noise add (sampling.py:101-127) is ONE kernel;
* nothing crosses to the host inside a step (the reference moves 4 tensors per step and does the update
  on the CPU).

Objects are independent, views of one object are sequential (`record` grows), so multi-GPU sampling
shards OBJECTS across ranks with no collective (`shard_objects`)."""
import random

import numpy as np
import torch

from .diff3d import _ops, logsnr_schedule_cosine


def _dev(model):
    m = model.module if hasattr(model, "module") else model
    return m.flat_parameters().device


def _t(a, dev, dtype):
    if not torch.is_tensor(a):
        a = torch.as_tensor(np.asarray(a))
    return a.to(device=dev, dtype=dtype)


def _guidance(w, b, dev):
    """guidance weights as fp32 [b]: the update kernel reads w[i] for image i, so a python scalar (the reference's `w=2.0`
    default, lightning/diff3d.py:200) is broadcast and any other length is refused here -- the C ABI cannot see the length"""
    w = _t(w, dev, torch.float32)
    if w.dim() == 0:
        w = w.expand(b)
    assert w.dim() == 1 and w.numel() == b, f"guidance weights: expected {b} values (one per image), got shape {tuple(w.shape)}"
    return w.contiguous()


def xt2batch(x, logsnr, z, R, T, K):
    """sampling.py:65-76 (K is one [3,3] matrix repeated over the batch)."""
    b = x.shape[0]
    return {
        'x': x.cuda(),
        'z': z.cuda(),
        'logsnr': torch.stack([logsnr_schedule_cosine(torch.zeros_like(logsnr)), logsnr], dim=1).cuda(),
        'R': R.cuda(),
        't': T.cuda(),
        'K': K[None].repeat(b, 1, 1).cuda(),
    }


@torch.no_grad()
def predict_noise_pair(model, x, z, R, T, K, logsnr, x_uncond=None, generator=None):
    """eps_cond, eps_uncond of sampling.py:97-99 from one batch-2b forward."""
    dev = _dev(model)
    b = z.shape[0]
    x, z = _t(x, dev, torch.float32), _t(z, dev, torch.float32)
    if x_uncond is None:
        x_uncond = torch.randn(x.shape, device=dev, dtype=torch.float32, generator=generator)
    lam = torch.as_tensor(float(logsnr)).repeat(2 * b)
    K2 = _t(K, dev, torch.float64)
    K2 = K2[None].repeat(2 * b, 1, 1) if K2.dim() == 2 else K2.repeat(2, 1, 1)
    batch = {
        'x': torch.cat([x, _t(x_uncond, dev, torch.float32)], 0),
        'z': torch.cat([z, z], 0),
        'logsnr': torch.stack([logsnr_schedule_cosine(torch.zeros_like(lam)), lam], dim=1).to(dev),
        'R': _t(R, dev, torch.float64).repeat(2, 1, 1, 1),
        't': _t(T, dev, torch.float64).repeat(2, 1, 1),
        'K': K2,
    }
    cond_mask = torch.cat([torch.ones(b, dtype=torch.bool), torch.zeros(b, dtype=torch.bool)]).to(dev)
    eps = model(batch, cond_mask=cond_mask)
    return eps[:b], eps[b:]


@torch.no_grad()
def p_mean_variance(model, x, z, R, T, K, logsnr, logsnr_next, w):
    """sampling.py:79-112: returns (model_mean, posterior_variance) like the reference (device tensors)."""
    dev = _dev(model)
    eps_c, eps_u = predict_noise_pair(model, x, z, R, T, K, logsnr)
    z = _t(z, dev, torch.float32)
    mean = torch.empty_like(z)
    # the same kernel as p_sample with a zero noise tensor gives the mean alone
    zero = torch.zeros_like(z) if float(logsnr_next) != 0.0 else None
    _ops(dev).sampler_update(z.contiguous(), eps_c.contiguous(), eps_u.contiguous(), _guidance(w, z.shape[0], dev), zero, mean, float(logsnr),
                             float(logsnr_next))
    lam, lam_n = torch.as_tensor(float(logsnr)), torch.as_tensor(float(logsnr_next))
    var = (-lam_n).sigmoid() * (-torch.special.expm1(lam - lam_n))
    return mean, var.to(dev)


@torch.no_grad()
def p_sample(model, z, x, R, T, K, logsnr, logsnr_next, w, noise=None, generator=None):
    """sampling.py:115-127: one denoise step.  `noise` (optional) is the randn_like of :127."""
    dev = _dev(model)
    eps_c, eps_u = predict_noise_pair(model, x, z, R, T, K, logsnr, generator=generator)
    z = _t(z, dev, torch.float32).contiguous()
    if noise is None and float(logsnr_next) != 0.0:
        noise = torch.randn(z.shape, device=dev, dtype=torch.float32, generator=generator)
    out = torch.empty_like(z)
    _ops(dev).sampler_update(z, eps_c.contiguous(), eps_u.contiguous(), _guidance(w, z.shape[0], dev),
                             _t(noise, dev, torch.float32).contiguous() if noise is not None else None, out, float(logsnr), float(logsnr_next))
    return out


@torch.no_grad()
def sample(model, record, target_R, target_T, K, w, timesteps=256, generator=None):
    """sampling.py:130-155: 256-step ancestral loop; each step conditions on a random entry of `record`
    (python `random.choice`, as the reference).  Returns np.ndarray [b,3,H,W]."""
    dev = _dev(model)
    first = _t(record[0][0], dev, torch.float32)
    b = first.shape[0]
    w = _guidance(w, b, dev)
    img = torch.randn(first.shape, device=dev, dtype=torch.float32, generator=generator)
    logsnrs = logsnr_schedule_cosine(torch.linspace(1., 0., timesteps + 1)[:-1])
    logsnr_nexts = logsnr_schedule_cosine(torch.linspace(1., 0., timesteps + 1)[1:])
    target_R, target_T = _t(target_R, dev, torch.float64), _t(target_T, dev, torch.float64)
    cache = {}
    for logsnr, logsnr_next in zip(logsnrs, logsnr_nexts):
        # random.choice(record) of sampling.py:138, in index form so device copies of the entries can be cached
        i = random.randrange(len(record))
        if i not in cache:
            ci, cR, cT = record[i]
            cache[i] = (_t(ci, dev, torch.float32), _t(cR, dev, torch.float64), _t(cT, dev, torch.float64))
        cimg, cR, cT = cache[i]
        R = torch.stack([cR, target_R], 0)[None].repeat(b, 1, 1, 1)
        T = torch.stack([cT, target_T], 0)[None].repeat(b, 1, 1)
        img = p_sample(model, z=img, x=cimg, R=R, T=T, K=K, logsnr=logsnr, logsnr_next=logsnr_next, w=w, generator=generator)
    return img.cpu().numpy()


@torch.no_grad()
def sample_fixed_condition(model, img, R, T, K, w, timesteps=256):
    """train.py:63-78 flavour: conditioning frame fixed to img[:,0]; returns the list of per-step arrays."""
    dev = _dev(model)
    x = _t(img, dev, torch.float32)[:, 0]
    z = torch.randn_like(x)
    imgs = []
    logsnrs = logsnr_schedule_cosine(torch.linspace(1., 0., timesteps + 1)[:-1])
    logsnr_nexts = logsnr_schedule_cosine(torch.linspace(1., 0., timesteps + 1)[1:])
    for logsnr, logsnr_next in zip(logsnrs, logsnr_nexts):
        z = p_sample(model, z=z, x=x, R=R, T=T, K=K, logsnr=logsnr, logsnr_next=logsnr_next, w=w)
        imgs.append(z.cpu().numpy())
    return imgs


def shard_objects(num_objects, rank, world_size):
    """object i -> rank i % world_size (no data-path collective)."""
    return [i for i in range(num_objects) if i % world_size == rank]


@torch.no_grad()
def sample_object(model, views, K, w, timesteps=256, max_views=None, generator=None):
    """Driver loop of sampling.py:158-176 for one object: `views` = [(img [3,H,W], R [3,3], T [3]), ...];
    view 0 is the given image, every following view is synthesised autoregressively.  Returns the record."""
    w_t = torch.as_tensor(np.asarray(w))
    b = w_t.shape[0]
    img0 = np.asarray(views[0][0])
    record = [[np.repeat(img0[None], b, axis=0), np.asarray(views[0][1]), np.asarray(views[0][2])]]
    for vi, (_, R, T) in enumerate(views[1:]):
        if max_views is not None and vi >= max_views:
            break
        img = sample(model, record=record, target_R=torch.as_tensor(np.asarray(R)), target_T=torch.as_tensor(np.asarray(T)), K=K, w=w_t,
                     timesteps=timesteps, generator=generator)
        record.append([img, np.asarray(R), np.asarray(T)])
    return record


class GraphSampler:
    """One denoise step (sampling.py:115-127) on static device buffers, captured once in a CUDA graph and replayed for
    every step / view: conditional + unconditional UNet forward as one batch-2b pass, then the fused update kernel.
    Per step the host only picks the conditioning entry (python `random`, as the reference) and copies two scalars.

    cache_conditioning: the conditioning path up to `+ logsnr_emb` (pose embedding + the four strided convs,
    xunet.py:311-349) depends only on the cameras, so it is evaluated once per (record entry, frame role), once per target
    view and once for the unconditional half, kept on the device in fp32 and reused by all denoising steps; the batch holds
    only FOUR distinct conditionings (cond frame 0, target frame 1, and the two unconditional ones), so FiLM's Linear is
    evaluated for 4 images instead of 2b.  Results are bit-identical to the uncached step.  Record images are kept on the
    device too."""

    def __init__(self, model, b, use_graph=True, cache_conditioning=True):
        m = model.module if hasattr(model, "module") else model
        self.model, self.b, self.dev = m, b, _dev(m)
        self.ops = _ops(self.dev)
        H, W = m.H, m.W
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.z = torch.zeros(b, 3, H, W, **f32)
        self.xz = torch.zeros(2 * b, 3, H, W, **f32)          # z duplicated for the cond / uncond halves
        self.x = torch.zeros(2 * b, 3, H, W, **f32)           # [conditioning image | randn]
        self.noise = torch.zeros(b, 3, H, W, **f32)
        self.R = torch.zeros(2 * b, 2, 3, 3, dtype=torch.float64, device=self.dev)
        self.T = torch.zeros(2 * b, 2, 3, dtype=torch.float64, device=self.dev)
        self.K = torch.zeros(2 * b, 3, 3, dtype=torch.float64, device=self.dev)
        self.w = torch.zeros(b, **f32)
        self.sched = torch.zeros(2, **f32)
        self.logsnr2 = torch.zeros(2 * b, 2, **f32)
        self.logsnr2[:, 0] = float(logsnr_schedule_cosine(torch.zeros(1))[0])
        self.mask = torch.cat([torch.ones(b, dtype=torch.uint8), torch.zeros(b, dtype=torch.uint8)]).to(self.dev)
        self.out = torch.zeros(b, 3, H, W, **f32)
        # a private engine: its buffers (and, while training goes on elsewhere, the saved activations of the model's own
        # engines) are never shared with a training step of the same batch size
        m._engine(2 * b)                                       # makes sure the model's kernel backend exists (raises on CPU)
        from .engine import Engine
        self.engine = Engine(m.cfg, m._ops, 2 * b)
        self.use_graph, self.graph, self.steps_done = use_graph, None, 0
        self._version = None
        self.cache_conditioning = cache_conditioning
        if cache_conditioning:
            self.engine_c = Engine(m.cfg, m._ops, 2)           # the 2-pair pass that yields the four distinct conditionings
            E = m.cfg.emb_ch
            self.cond_c = [torch.zeros(4, H >> l, W >> l, E, **f32) for l in range(m.cfg.num_resolutions)]
            # image n = 2*pair + frame; pairs < b are conditional, the rest unconditional
            self.image = torch.tensor([2 * int(n >= 2 * b) + (n & 1) for n in range(4 * b)], dtype=torch.int32, device=self.dev)
            self.rep = torch.tensor([0, 1, 2 * b, 2 * b + 1], dtype=torch.int32, device=self.dev)
            self._R2 = torch.zeros(2, 2, 3, 3, dtype=torch.float64, device=self.dev)
            self._T2 = torch.zeros(2, 2, 3, dtype=torch.float64, device=self.dev)
            self._K2 = torch.zeros(2, 3, 3, dtype=torch.float64, device=self.dev)
            self._mask2 = torch.tensor([1, 0], dtype=torch.uint8, device=self.dev)
        self._entries, self._target_key, self._cur_keys = {}, None, None

    def refresh_weights(self, force=False):
        """Repack the bf16 GEMM operands from the model's current fp32 parameters (the captured step only READS them) and drop
        the cached conditioning, if the parameters changed since the last call (optimizer steps and load_state_dict bump
        `model._param_version`; after editing parameters in place by hand, pass force=True).  `sample()` calls this first."""
        v = getattr(self.model, "_param_version", 0)
        if not force and self._version == v:
            return
        P = dict(zip(self.model._names, self.model._plist))
        self.engine._assign_cat_sources()
        self.engine._pack_weights(P, False, True)
        if self.cache_conditioning:
            self.engine_c._assign_cat_sources()
            self.engine_c._pack_weights(P, False, True)
        self._entries, self._target_key, self._cur_keys = {}, None, None
        self._version = v

    def _compute(self):
        b = self.b
        self.x[b:].normal_()                                     # batch['x'] = randn_like(x)        (sampling.py:98)
        self.noise.normal_()                                     # randn_like(x) of p_sample          (sampling.py:127)
        self.xz[:b].copy_(self.z)
        self.xz[b:].copy_(self.z)
        self.logsnr2[:, 1].copy_(self.sched[0].expand(2 * b))
        P = dict(zip(self.model._names, self.model._plist))
        cond = dict(c=self.cond_c, image=self.image, rep=self.rep) if self.cache_conditioning else None
        eps = self.engine.forward(P, self.x, self.xz, self.logsnr2, self.R, self.T, self.K, self.mask, training=self.model.training,
                                  need_bwd=False, seed=0, repack=False, cond=cond)     # operands packed by refresh_weights()
        self.ops.sampler_update_dev(self.z, eps[:b], eps[b:], self.w, self.noise, self.out, self.sched)
        self.z.copy_(self.out)

    def _cond_pass(self):
        """conditioning convs of the pairs (cond cam, target cam) x (conditional, unconditional): 4 images = the 4 slots"""
        P = dict(zip(self.model._names, self.model._plist))
        return self.engine_c.cond_features(P, self._R2, self._T2, self._K2, self._mask2, repack=False)

    def set_views(self, cond_img, cond_R, cond_T, target_R, target_T, K, cond_key=None, target_key=None):
        """conditioning view (image + camera), target camera and intrinsics of the next step(s).  cond_key / target_key
        (hashable, optional) identify the conditioning entry and the target view, so that the entry's device image and the
        cached conditioning are reused; without a key the respective part is re-evaluated on every call."""
        self.refresh_weights()
        b = self.b
        if cond_key is not None and target_key is not None and (cond_key, target_key) == self._cur_keys:
            return                                   # same views as the previous step: every buffer is already in place
        self._cur_keys = (cond_key, target_key)
        ent = self._entries.get(cond_key) if cond_key is not None else None
        if ent is None:
            ent = {"img": _t(cond_img, self.dev, torch.float32).clone(), "R": _t(cond_R, self.dev, torch.float64).clone(),
                   "T": _t(cond_T, self.dev, torch.float64).clone(), "c": None, "ref": cond_img}     # ref: keeps id(cond_img) unique
            if cond_key is not None:
                self._entries[cond_key] = ent
        tR, tT, Kd = _t(target_R, self.dev, torch.float64), _t(target_T, self.dev, torch.float64), _t(K, self.dev, torch.float64)
        Kd = Kd if Kd.dim() == 2 else Kd[0]
        self.x[:b].copy_(ent["img"])
        self.R[:, 0].copy_(ent["R"].expand(2 * b, 3, 3))
        self.R[:, 1].copy_(tR.expand(2 * b, 3, 3))
        self.T[:, 0].copy_(ent["T"].expand(2 * b, 3))
        self.T[:, 1].copy_(tT.expand(2 * b, 3))
        self.K.copy_(Kd.expand(2 * b, 3, 3))
        if not self.cache_conditioning:
            return
        tkey = target_key
        new_target = tkey is None or tkey != self._target_key
        if new_target or ent["c"] is None:
            self._R2[:, 0].copy_(ent["R"].expand(2, 3, 3))
            self._R2[:, 1].copy_(tR.expand(2, 3, 3))
            self._T2[:, 0].copy_(ent["T"].expand(2, 3))
            self._T2[:, 1].copy_(tT.expand(2, 3))
            self._K2.copy_(Kd.expand(2, 3, 3))
            feats = self._cond_pass()
            if new_target:                       # slots 1..3: target frame and the two unconditional images
                for dst, src in zip(self.cond_c, feats):
                    dst[1:].copy_(src[1:])
                self._target_key = tkey
            ent["c"] = [src[0].clone() for src in feats]
        for dst, src in zip(self.cond_c, ent["c"]):      # slot 0: this entry as the conditioning frame
            dst[0].copy_(src)

    def step(self, logsnr, logsnr_next):
        self.sched[0] = float(logsnr)
        self.sched[1] = float(logsnr_next)
        self.refresh_weights()
        if self.use_graph and self.graph is None and self.steps_done >= 1:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._compute()
            self.graph = g
        if self.graph is not None:
            self.graph.replay()
        else:
            self._compute()
        self.steps_done += 1

    @torch.no_grad()
    def sample(self, record, target_R, target_T, K, w, timesteps=256):
        """sampling.py:130-155 on the captured step.  Entries of `record` are recognised by the identity of their image array
        (the record of an object only grows), so each is uploaded and its conditioning evaluated once per object."""
        self.refresh_weights()
        self.w.copy_(_guidance(w, self.b, self.dev))
        self.z.normal_()
        ts = torch.linspace(1., 0., timesteps + 1)
        logsnrs, logsnr_nexts = logsnr_schedule_cosine(ts[:-1]), logsnr_schedule_cosine(ts[1:])
        tkey = object()                                              # one target view per call
        live = {id(e[0]) for e in record}
        for k in [k for k in self._entries if k not in live]:       # entries of an object that is no longer sampled
            del self._entries[k]
        for l0, l1 in zip(logsnrs, logsnr_nexts):
            ci, cR, cT = record[random.randrange(len(record))]       # random.choice(record), sampling.py:138
            self.set_views(ci, cR, cT, target_R, target_T, K, cond_key=id(ci), target_key=tkey)
            self.step(l0, l1)
        return self.z.cpu().numpy()
