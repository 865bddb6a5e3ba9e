"""TEST INFRASTRUCTURE ONLY -- fp32 restatement of the reference's diffusion glue and sampler
update (never imported by the product package).  Pinned against the reference's own functions by
tests/golden/make_golden.py.  Paths are relative to /root/reference."""
import numpy as np
import torch
import torch.nn.functional as F


def logsnr_schedule_cosine(t, logsnr_min=-20.0, logsnr_max=20.0):
    """train.py:30-34 / sampling.py:59-63 (a, b are float64 numpy scalars; tensor math fp32)."""
    b = np.arctan(np.exp(-0.5 * logsnr_max))
    a = np.arctan(np.exp(-0.5 * logsnr_min)) - b
    return -2.0 * torch.log(torch.tan(a * t + b))


def q_sample(z, logsnr, noise):
    """train.py:50-60."""
    alpha = logsnr.sigmoid().sqrt()[:, None, None, None]
    sigma = (-logsnr).sigmoid().sqrt()[:, None, None, None]
    return alpha * z + sigma * noise


def make_batch(x, logsnr, z, R, T, K):
    """xt2batch, train.py:36-46 (device moves dropped)."""
    return {"x": x, "z": z,
            "logsnr": torch.stack([logsnr_schedule_cosine(torch.zeros_like(logsnr)), logsnr], 1),
            "R": R, "t": T, "K": K}


def loss_fn(noise, pred, loss_type="l2"):
    """train.py:102-112."""
    if loss_type == "l1":
        return F.l1_loss(noise, pred)
    if loss_type == "l2":
        return F.mse_loss(noise, pred)
    if loss_type == "huber":
        return F.smooth_l1_loss(noise, pred)
    raise NotImplementedError()


def sampler_update(z, eps_c, eps_u, w, logsnr, logsnr_next, noise):
    """p_mean_variance + p_sample arithmetic, sampling.py:79-127, given the two UNet outputs.
    Returns (x_next, mean, var).  noise is the randn_like drawn at sampling.py:127."""
    w = w.float()[:, None, None, None]
    c = -torch.special.expm1(logsnr - logsnr_next)
    sq_a, sq_an = logsnr.sigmoid(), logsnr_next.sigmoid()
    sq_s, sq_sn = (-logsnr).sigmoid(), (-logsnr_next).sigmoid()
    alpha, sigma, alpha_next = sq_a.sqrt(), sq_s.sqrt(), sq_an.sqrt()
    eps = (1 + w) * eps_c - w * eps_u
    z_start = ((z - sigma * eps) / alpha).clamp(-1.0, 1.0)
    mean = alpha_next * (z * (1 - c) / alpha + c * z_start)
    var = sq_sn * c
    if logsnr_next == 0:          # sampling.py:124 (fires only at t=0.5, Q11)
        return mean, mean, var
    return mean + var.sqrt() * noise, mean, var
