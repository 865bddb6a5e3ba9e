"""Drop-in for the reference's diffusion glue: the free functions of train.py:30-177 and the
`Diff3D` training module of lightning/diff3d.py (same names, argument meaning and error behaviour).
Elementwise math (q_sample, conditioning-frame replacement, loss) runs in the b3d kernels; the host
side keeps the reference's RNG call order so that a seed reproduces the same inputs."""
import numpy as np
import torch

from . import _lib
from .ops import CudaOps
from .xunet import XUNet

_LOSS_TYPES = {"l2": 0, "l1": 1, "huber": 2}
_ops_cache = {}


def _ops(device):
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.B3DError("b3d diffusion kernels need CUDA tensors (no CPU fallback)")
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _ops_cache:
        _ops_cache[key] = CudaOps(torch.device("cuda", key))
    return _ops_cache[key]


def logsnr_schedule_cosine(t, *, logsnr_min=-20., logsnr_max=20.):
    """train.py:30-34 (host-side scalar/table math, identical expression)."""
    b = np.arctan(np.exp(-.5 * logsnr_max))
    a = np.arctan(np.exp(-.5 * logsnr_min)) - b
    return -2. * torch.log(torch.tan(a * t + b))


def xt2batch(x, logsnr, z, R, T, K):
    """train.py:36-46."""
    return {
        'x': x.cuda(),
        'z': z.cuda(),
        'logsnr': torch.stack([logsnr_schedule_cosine(torch.zeros_like(logsnr)), logsnr], dim=1).cuda(),
        'R': R.cuda(),
        't': T.cuda(),
        'K': K.cuda(),
    }


def q_sample(z, logsnr, noise):
    """Forward process q(x_t|x_0), train.py:50-60, as one kernel."""
    ops = _ops(z.device)
    z = z.contiguous().float()
    out = torch.empty_like(z)
    ops.q_sample(z, noise.contiguous().float(), logsnr.to(z.device, torch.float32).contiguous(), out)
    return out


class _LossFn(torch.autograd.Function):
    """mean-reduced l2 / l1 / smooth-l1 and its gradient in one pass (train.py:102-112)"""

    @staticmethod
    def forward(ctx, noise, pred, loss_type):
        ops = _ops(pred.device)
        noise, pred = noise.contiguous().float(), pred.contiguous().float()
        dpred = torch.empty_like(pred)
        loss = torch.empty(1, dtype=torch.float32, device=pred.device)
        scratch = _scratch(pred.device)
        ops.loss(noise, pred, dpred, loss_type, scratch, loss)
        ctx.save_for_backward(dpred)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dpred,) = ctx.saved_tensors
        return None, dpred * g, None


_scratch_cache = {}


def _scratch(device):
    key = str(device)
    if key not in _scratch_cache:
        _scratch_cache[key] = _ops(device).loss_scratch()
    return _scratch_cache[key]


def diffusion_loss(noise, predicted_noise, loss_type="l2"):
    if loss_type not in _LOSS_TYPES:
        raise NotImplementedError()
    return _LossFn.apply(noise, predicted_noise, _LOSS_TYPES[loss_type])


def p_losses(denoise_model, img, R, T, K, logsnr, noise=None, loss_type="l2", cond_prob=0.1):
    """train.py:80-114: same draws in the same order (randn_like for noise, CPU rand for the mask,
    randn_like for the replacement frame)."""
    if loss_type not in _LOSS_TYPES:
        raise NotImplementedError()
    B = img.shape[0]
    x = img[:, 0]
    z = img[:, 1]
    if noise is None:
        noise = torch.randn_like(x)
    cond_mask = (torch.rand((B,)) > cond_prob).cuda()
    x_rand = torch.randn_like(x)
    ops = _ops(img.device)
    z_noisy, x_condition = torch.empty_like(z, dtype=torch.float32), torch.empty_like(x, dtype=torch.float32)
    ops.q_sample(z.contiguous().float(), noise.contiguous().float(), logsnr.to(img.device, torch.float32).contiguous(), z_noisy,
                 x.contiguous().float(), x_rand.contiguous().float(), cond_mask, x_condition)
    batch = xt2batch(x=x_condition, logsnr=logsnr, z=z_noisy, R=R, T=T, K=K)
    predicted_noise = denoise_model(batch, cond_mask=cond_mask.cuda())
    return diffusion_loss(noise, predicted_noise, loss_type)


def warmup(optimizer, step, last_step, last_lr):
    """train.py:169-177."""
    if step < last_step:
        optimizer.param_groups[0]['lr'] = step / last_step * last_lr
    else:
        optimizer.param_groups[0]['lr'] = last_lr


class FusedAdam:
    """Adam(lr, betas=(0.9, 0.99)) (train.py:235, lightning/diff3d.py:106) as ONE kernel over the model's flat
    fp32 parameter arena; exposes `param_groups[0]['lr']` so `warmup` works unchanged.

    ema_decay (optional): the exponential moving average of the parameters that the reference only describes
    (lightning/diff3d.py:11-19), kept in a second flat arena by the same kernel (`ema_state_dict()` gives it under the
    model's parameter names).

    Checkpoints: `state_dict()` emits, and `load_state_dict()` accepts, the layout of `torch.optim.Adam.state_dict()` that
    the reference stores under ckpt['optim'] (train.py:247-251,289,298): per-parameter `exp_avg` / `exp_avg_sq` / `step` in
    parameter-registration order.  The compact flat layout {step, m, v, lr} of earlier versions still loads."""

    def __init__(self, model: XUNet, lr=1e-4, betas=(0.9, 0.99), eps=1e-8, ema_decay=None):
        self.model = model
        self.param_groups = [{"lr": lr, "betas": betas, "eps": eps, "params": list(model.parameters())}]
        self.step_count = 0
        self.m = self.v = None
        self.ema_decay, self.ema = ema_decay, None

    def zero_grad(self, set_to_none=True):
        pass                                    # the engine overwrites the flat gradient arena every backward

    def _arenas(self):
        flat = self.model.flat_parameters()
        if self.m is None or self.m.shape != flat.shape or self.m.device != flat.device:
            self.m, self.v = torch.zeros_like(flat), torch.zeros_like(flat)
        if self.ema_decay is not None and (self.ema is None or self.ema.shape != flat.shape or self.ema.device != flat.device):
            self.ema = flat.clone()
        return flat

    def step(self, grad_scale=1.0, lo=None, hi=None, first=True, grad=None):
        """One Adam step.  (lo, hi): update only that element range of the arena -- a data-parallel trainer steps each gradient
        range as soon as its all-reduce has finished, beside the rest of the backward pass; `first` marks the first range of a
        step (the step count advances once per step).  grad: the gradient of that range in a buffer of its own (fp32 or bf16:
        the all-reduced communication buffer, read in place) instead of the model's flat gradient arena."""
        flat, g = self._arenas(), self.model.flat_gradients()
        if first:
            self.step_count += 1
            self.model._param_version = getattr(self.model, "_param_version", 0) + 1
        pg = self.param_groups[0]
        sl = slice(lo, hi)
        gs = g[sl] if grad is None else grad
        assert gs.numel() == flat[sl].numel()
        _ops(flat.device).adam(flat[sl], gs, self.m[sl], self.v[sl], pg["lr"], pg["betas"][0], pg["betas"][1], pg["eps"], self.step_count,
                               grad_scale, ema=self.ema[sl] if self.ema is not None else None, ema_decay=self.ema_decay or 0.0)

    def ema_state_dict(self):
        """EMA weights under the model's state_dict names (load them into an XUNet to sample with them)"""
        assert self.ema is not None, "FusedAdam was built without ema_decay"
        m = self.model
        return {n: self.ema[o:o + p.numel()].view(p.shape).clone() for n, p, o in zip(m._names, m._plist, m._offs)}

    def state_dict(self, torch_format=True):
        m = self.model
        flat = self._arenas()
        if not torch_format:
            return {"step": self.step_count, "m": self.m, "v": self.v, "lr": self.param_groups[0]["lr"], "ema": self.ema}
        pg = self.param_groups[0]
        state = {}
        if self.step_count > 0:
            for i, (p, o) in enumerate(zip(m._plist, m._offs)):
                state[i] = {"step": torch.tensor(float(self.step_count)), "exp_avg": self.m[o:o + p.numel()].view(p.shape).clone(),
                            "exp_avg_sq": self.v[o:o + p.numel()].view(p.shape).clone()}
        group = {"lr": pg["lr"], "betas": tuple(pg["betas"]), "eps": pg["eps"], "weight_decay": 0, "amsgrad": False, "maximize": False,
                 "foreach": None, "capturable": False, "differentiable": False, "fused": None, "params": list(range(len(m._plist)))}
        sd = {"state": state, "param_groups": [group]}
        if self.ema is not None:
            sd["b3d_ema"] = self.ema.clone()          # extra key: ignored by torch.optim.Adam.load_state_dict
        return sd

    def load_state_dict(self, sd):
        m = self.model
        flat = self._arenas()
        if isinstance(sd, dict) and "m" in sd and "v" in sd:                      # compact flat layout
            self.step_count = int(sd["step"])
            self.m.copy_(sd["m"])
            self.v.copy_(sd["v"])
            self.param_groups[0]["lr"] = sd["lr"]
            if sd.get("ema") is not None and self.ema is not None:
                self.ema.copy_(sd["ema"])
            return
        if not (isinstance(sd, dict) and "state" in sd and "param_groups" in sd):
            raise ValueError("FusedAdam.load_state_dict: expected a torch.optim.Adam state_dict ({'state', 'param_groups'}) or the flat "
                             f"{{'step','m','v','lr'}} layout, got keys {sorted(sd.keys()) if isinstance(sd, dict) else type(sd)}")
        groups = sd["param_groups"]
        ids = [i for g_ in groups for i in g_["params"]]
        if len(ids) != len(m._plist):
            raise ValueError(f"optimizer state covers {len(ids)} parameters, the model has {len(m._plist)}")
        steps = set()
        self.m.zero_()
        self.v.zero_()
        for pid, p, o, name in zip(ids, m._plist, m._offs, m._names):
            st = sd["state"].get(pid)
            if st is None:
                continue                                   # parameter that never received a gradient: zero moments
            if tuple(st["exp_avg"].shape) != tuple(p.shape):
                raise ValueError(f"optimizer state of {name}: shape {tuple(st['exp_avg'].shape)} != parameter shape {tuple(p.shape)}")
            self.m[o:o + p.numel()].view(p.shape).copy_(st["exp_avg"])
            self.v[o:o + p.numel()].view(p.shape).copy_(st["exp_avg_sq"])
            steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError(f"per-parameter Adam step counts differ ({sorted(steps)}): one fused step count cannot represent them")
        self.step_count = steps.pop() if steps else 0
        g0 = groups[0]
        self.param_groups[0]["lr"] = g0["lr"]
        self.param_groups[0]["betas"] = tuple(g0.get("betas", self.param_groups[0]["betas"]))
        self.param_groups[0]["eps"] = g0.get("eps", self.param_groups[0]["eps"])
        if g0.get("weight_decay", 0) != 0 or g0.get("amsgrad", False):
            raise ValueError("FusedAdam implements the reference's plain Adam (no weight decay, no amsgrad)")
        if sd.get("b3d_ema") is not None and self.ema is not None:
            self.ema.copy_(sd["b3d_ema"])


def save_checkpoint(path, model, optimizer, step, epoch=None, module_prefix=False):
    """The reference's checkpoint file (train.py:289,298): {'optim': Adam state_dict, 'model': state_dict, 'step'[, 'epoch']}.
    module_prefix=True writes the `module.`-prefixed keys of a checkpoint saved from DataParallel/DDP (what the public
    pretrained checkpoint and sampling.py:52-56 use)."""
    msd = {("module." + k if module_prefix else k): v.detach().cpu().clone() for k, v in model.state_dict().items()}
    osd = optimizer.state_dict()
    if "state" in osd:
        osd = {**osd, "state": {i: {k: v.cpu() if torch.is_tensor(v) else v for k, v in st.items()} for i, st in osd["state"].items()}}
        if osd.get("b3d_ema") is not None:
            osd["b3d_ema"] = osd["b3d_ema"].cpu()
    ckpt = {"optim": osd, "model": msd, "step": step}
    if epoch is not None:
        ckpt["epoch"] = epoch
    torch.save(ckpt, path)
    return ckpt


def load_checkpoint(path, model, optimizer=None, map_location="cpu"):
    """train.py:246-251: restores model (with or without the `module.` prefix), optimizer state and returns the step"""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    model.load_state_dict(ckpt["model"])
    if optimizer is not None and ckpt.get("optim") is not None:
        optimizer.load_state_dict(ckpt["optim"])
    return ckpt.get("step", 0)


try:                                            # Lightning is optional (not in this image)
    import pytorch_lightning as _pl
    _Base = _pl.LightningModule
except Exception:                               # noqa: BLE001
    _Base = torch.nn.Module


class Diff3D(_Base):
    """lightning/diff3d.py:9-127.  Works without Lightning: `training_step` then drives its own optimizer
    (created by `configure_optimizers`)."""

    def __init__(self, pretrained_model=None, n_samples=10000000, image_size=64, batch_size=128, lr=1e-4, use_scheduler=False):
        super().__init__()
        self.n_samples = n_samples
        self.batch_size = batch_size
        self.image_size = image_size
        self.lr = lr
        self.use_scheduler = use_scheduler
        self.step = 0
        self.pretrained_optim = None
        self.optimizer = None
        self.xunet_denoiser = XUNet(H=self.image_size, W=self.image_size, ch=128)
        if pretrained_model is not None:
            print('[-] Loading pre-trained model: ', pretrained_model)
            ckpt = torch.load(pretrained_model, map_location="cpu", weights_only=False)
            self.xunet_denoiser.load_state_dict(ckpt['model'])
            self.pretrained_optim = ckpt['optim']

    # Lightning provides these; without it they are plain attributes
    if _Base is torch.nn.Module:
        @property
        def device(self):
            return self.xunet_denoiser.flat_parameters().device

        def optimizers(self):
            if self.optimizer is None:
                self.optimizer = self.configure_optimizers()[0]
            return self.optimizer

        def lr_schedulers(self):
            return self._scheduler

    def forward(self, in_x, noise=None, cond_prob=0.1):
        """lightning/diff3d.py:52-74: draws logsnr, noises z, replaces the conditioning frame, predicts eps."""
        img, R, T, K = in_x
        dev = self.device
        img = img.to(dev)
        B = img.shape[0]
        x = img[:, 0]
        z = img[:, 1]
        logsnr = self.logsnr_schedule_cosine(torch.rand((B,)))
        if noise is None:
            noise = torch.randn_like(x)
        cond_mask = torch.Tensor(torch.rand((B,)) > cond_prob).to(dev)
        x_rand = torch.randn_like(x).to(dev)
        ops = _ops(dev)
        z_noisy, x_condition = torch.empty_like(z, dtype=torch.float32), torch.empty_like(x, dtype=torch.float32)
        ops.q_sample(z.contiguous().float(), noise.to(dev).contiguous().float(), logsnr.to(dev, torch.float32), z_noisy,
                     x.contiguous().float(), x_rand.contiguous().float(), cond_mask, x_condition)
        batch = self.xt2batch(x=x_condition, logsnr=logsnr.to(dev), z=z_noisy, R=R.to(dev), T=T.to(dev), K=K.to(dev))
        return self.xunet_denoiser(batch, cond_mask=cond_mask).to(dev)

    def training_step(self, batch, batch_idx, loss_type="l2"):
        """lightning/diff3d.py:77-101 (returns the loss; under Lightning the trainer does backward + step)."""
        if loss_type not in _LOSS_TYPES:
            raise NotImplementedError()
        if not self.use_scheduler:
            self.warmup()
        self.optimizers().zero_grad()
        noise = torch.randn_like(batch[0][:, 0]).to(self.device)
        predicted_noise = self.forward(in_x=batch, noise=noise)
        loss = diffusion_loss(noise, predicted_noise, loss_type)
        self.step += 1
        if self.use_scheduler:
            self.lr_schedulers().step()
        return loss

    def configure_optimizers(self):
        """lightning/diff3d.py:104-116; the optimizer is the fused flat-arena Adam."""
        self.xunet_denoiser.fused_grads = True
        optimizer = FusedAdam(self.xunet_denoiser, lr=self.lr, betas=(0.9, 0.99))
        if self.pretrained_optim is not None:          # torch Adam layout (the reference's ckpt['optim']) or the flat one; else raises
            optimizer.load_state_dict(self.pretrained_optim)
        if self.use_scheduler:
            self._scheduler = _CosineLR(optimizer, T_max=300)
            return [optimizer], [self._scheduler]
        return [optimizer]

    def warmup(self):
        """lightning/diff3d.py:118-127."""
        last_step = self.n_samples / self.batch_size,
        if self.step < last_step[0]:
            self.optimizers().param_groups[0]['lr'] = (self.step / last_step[0]) * self.lr
        else:
            self.optimizers().param_groups[0]['lr'] = self.lr

    # same helper names as the reference's methods
    def logsnr_schedule_cosine(self, t, *, logsnr_min=-20., logsnr_max=20.):
        return logsnr_schedule_cosine(t, logsnr_min=logsnr_min, logsnr_max=logsnr_max)

    def xt2batch(self, x, logsnr, z, R, T, K):
        """lightning/diff3d.py:137-147 (no device moves)."""
        return {'x': x, 'z': z,
                'logsnr': torch.stack([self.logsnr_schedule_cosine(torch.zeros_like(logsnr)), logsnr], dim=1),
                'R': R, 't': T, 'K': K}

    def q_sample(self, z, logsnr, noise):
        return q_sample(z, logsnr, noise)

    @torch.no_grad()
    def sample(self, model, img, R, T, K, w, timesteps=256):
        from . import sampling
        return sampling.sample_fixed_condition(model, img, R, T, K, w, timesteps)

    @torch.no_grad()
    def p_sample(self, model, x, z, R, T, K, logsnr, logsnr_next, w):
        from . import sampling
        return sampling.p_sample(model, z=z, x=x, R=R, T=T, K=K, logsnr=logsnr, logsnr_next=logsnr_next, w=w)

    @torch.no_grad()
    def p_mean_variance(self, model, x, z, R, T, K, logsnr, logsnr_next, w=2.0):
        from . import sampling
        return sampling.p_mean_variance(model, x, z, R, T, K, logsnr, logsnr_next, w)


class _CosineLR:
    """torch.optim.lr_scheduler.CosineAnnealingLR(T_max) for FusedAdam (lightning/diff3d.py:111-113)"""

    def __init__(self, opt, T_max, eta_min=0.0):
        self.opt, self.T_max, self.eta_min, self.t = opt, T_max, eta_min, 0
        self.base = opt.param_groups[0]["lr"]

    def step(self):
        self.t += 1
        self.opt.param_groups[0]["lr"] = self.eta_min + (self.base - self.eta_min) * (1 + np.cos(np.pi * self.t / self.T_max)) / 2
