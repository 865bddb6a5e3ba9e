"""Data-parallel training step of the X-UNet (reference train.py:254-292 / lightning/diff3d.py:77-101),
one process per GPU:

    q_sample + conditioning-frame replacement -> XUNet forward -> loss -> hand-written backward
    -> gradient all-reduce over NCCL (N>1; the reference's DDP never armed it) -> fused Adam

The compute part (everything up to the gradients) works on static device buffers and is captured once
in four CUDA graphs (forward + backward phase 0 | phase 1 | phase 2a | 2b); draws (logsnr, noise, cond mask, dropout) are made
on the device inside the graph (dropout through a device-side step counter), so replays are fresh steps, not repeats.
The UPDATE of each gradient range -- all-reduce (N>1), Adam, and the repack of the bf16 GEMM operands built from those
parameters -- runs on a second stream as soon as the phase that finalises the range has finished, beside the later
phases; the next forward packs nothing."""
import math
import os

import numpy as np
import torch
import torch.distributed as dist

from . import parallel
from .diff3d import FusedAdam
from .ops import CudaOps
from .xunet import XUNet


def phase_ranges(names, offs, numel, prefixes):
    """Element ranges [lo, hi) of the flat gradient arena that each backward phase finalises (engine.PHASE_PREFIXES).
    The arena is laid out in parameter-registration order: conditioningprocessor, conv (phase 2) | xunetblocks, middle
    (phase 1) | upsample, lastgn, lastconv (phase 0).  Refuses any layout in which the phases are not three contiguous runs
    covering [0, numel) exactly once: a gap would leave gradients un-reduced, silently."""
    def phase_of(n):
        hits = [ph for ph, pre in enumerate(prefixes) if n.startswith(pre)]
        assert len(hits) == 1, f"parameter {n} belongs to {len(hits)} backward phases"
        return hits[0]
    seq = [phase_of(n) for n in names]
    runs = [(ph, o) for i, (ph, o) in enumerate(zip(seq, offs)) if i == 0 or seq[i - 1] != ph]
    assert [ph for ph, _ in runs] == [2, 1, 0], f"flat arena is not ordered phase 2 | phase 1 | phase 0: {[ph for ph, _ in runs]}"
    assert runs[0][1] == 0
    return {2: (0, runs[1][1]), 1: (runs[1][1], runs[2][1]), 0: (runs[2][1], numel)}


class Trainer:
    def __init__(self, model_kw, B, device, world=1, seed=0, use_graph=True, lr=1e-4, cond_prob=0.1, init_sd=None, ema_decay=None,
                 grad_comm_dtype=torch.bfloat16):
        self.B, self.dev, self.world = B, torch.device(device), world
        # every rank draws its own noise / logsnr / cond-mask / dropout streams: the rank enters the seed here, so
        # callers pass the same `seed` on all ranks (the initial weights are rank 0's, broadcast below)
        self.seed = int(seed) + parallel.rank()
        torch.manual_seed(self.seed)
        self.model = XUNet(**model_kw).to(self.dev)
        if init_sd is not None:
            self.model.load_state_dict(init_sd)
        parallel.broadcast_parameters(self.model.flat_parameters())   # DDP's rank-0 broadcast (train.py:230)
        self.model.train()
        self.model.fused_grads = True
        self.opt = FusedAdam(self.model, lr=lr, betas=(0.9, 0.99), ema_decay=ema_decay)
        self.ops = CudaOps(self.dev)
        # launch-latency hiding pays for small per-GPU batches only (measured: +1.9% at 16 pairs, -0.6% at 128 of 64x64)
        from . import _lib
        pdl = 2 * B * model_kw["H"] * model_kw["W"] <= 64 * 64 * 64
        if os.environ.get("B3D_PDL") is not None:         # measurement switch
            pdl = os.environ["B3D_PDL"] == "1"
        _lib.set_pdl(pdl)
        self.cond_prob = cond_prob
        # payload of the gradient all-reduce: bf16 halves the bytes on NVLink (273 instead of 547 MB at CFG-64); the fp32 arena is
        # cast on the communication stream, reduced, and cast back before that range's Adam step.  torch.float32 = exact sum.
        self.grad_comm_dtype, self._gcomm = grad_comm_dtype, None
        H, W = model_kw["H"], model_kw["W"]
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.img = torch.zeros(B, 2, 3, H, W, **f32)
        self.xi, self.zi = torch.zeros(B, 3, H, W, **f32), torch.zeros(B, 3, H, W, **f32)
        self.R = torch.zeros(B, 2, 3, 3, dtype=torch.float64, device=self.dev)
        self.T = torch.zeros(B, 2, 3, dtype=torch.float64, device=self.dev)
        self.K = torch.zeros(B, 3, 3, dtype=torch.float64, device=self.dev)
        self.u = torch.zeros(B, **f32)
        self.logsnr = torch.zeros(B, **f32)
        self.logsnr2 = torch.zeros(B, 2, **f32)
        self.noise = torch.zeros(B, 3, H, W, **f32)
        self.xrand = torch.zeros(B, 3, H, W, **f32)
        self.mask = torch.zeros(B, dtype=torch.uint8, device=self.dev)
        self.z_noisy = torch.zeros(B, 3, H, W, **f32)
        self.x_cond = torch.zeros(B, 3, H, W, **f32)
        self.deps = torch.zeros(B, 3, H, W, **f32)
        self.loss = torch.zeros(1, **f32)
        self.scratch = self.ops.loss_scratch()
        self.step_ctr = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.engine = self.model._engine(B)
        self.engine.seed_ptr = self.step_ctr
        # logsnr_schedule_cosine constants (train.py:30-34)
        b = np.arctan(np.exp(-0.5 * 20.0))
        self._a, self._b = float(np.arctan(np.exp(0.5 * 20.0)) - b), float(b)
        self._lam0 = float(-2.0 * np.log(np.tan(self._b)))        # schedule(0) = +20 for the clean conditioning frame
        self.use_graph, self.graph, self.graphed = use_graph, None, False
        self.launches_per_step = 0
        self._packed, self._pv = False, 0
        self._count_update = False
        self.step_idx = 0

    # ------------------------------------------------------------------ one step, device resident
    def _draws_forward_loss(self):
        """draws + forward + loss on the static buffers (graph-capturable: no allocation, no sync)"""
        self._draws()
        self._forward_loss()

    def _draws(self):
        self.step_ctr.add_(1)
        self.u.uniform_()                                            # logsnr = schedule(rand(B))      (train.py:272)
        torch.log(torch.tan(self.u * self._a + self._b), out=self.logsnr)
        self.logsnr.mul_(-2.0)
        self.noise.normal_()                                         # randn_like(x)                   (train.py:90)
        self.u.uniform_()                                            # cond_mask = rand(B) > cond_prob (train.py:95)
        self.mask.copy_(self.u > self.cond_prob)
        self.xrand.normal_()                                         # randn_like(x) for uncond rows   (train.py:96)

    def _forward_loss(self):
        self.xi.copy_(self.img[:, 0])
        self.zi.copy_(self.img[:, 1])
        self.ops.q_sample(self.zi, self.noise, self.logsnr, self.z_noisy, self.xi, self.xrand, self.mask, self.x_cond)
        self.logsnr2[:, 0].fill_(self._lam0)                         # xt2batch                        (train.py:36-46)
        self.logsnr2[:, 1].copy_(self.logsnr)
        P = dict(zip(self.model._names, self.model._plist))
        # the bf16 operands were repacked behind the previous step's Adam (see _update): only the very first forward packs
        eps = self.engine.forward(P, self.x_cond, self.z_noisy, self.logsnr2, self.R, self.T, self.K, self.mask, training=True, need_bwd=True,
                                  seed=self.seed % 1000003, repack=not self._packed)
        self.ops.loss(self.noise, eps, self.deps, 0, self.scratch, self.loss)

    def _backward(self, phase=None):
        self.engine.backward(self.model._grad_views(), self.deps, grad_flat=self.model._grad_flat, phase=phase)

    def _compute(self):
        self._draws_forward_loss()
        self._backward()

    def _phase_ranges(self):
        m = self.model
        return phase_ranges(m._names, m._offs, m._grad_flat.numel(), self.engine.PHASE_PREFIXES)

    def _finish(self):
        parallel.allreduce_mean_(self.model._grad_flat)
        self.opt.step()
        self.repack()
        self.step_idx += 1

    def repack(self):
        """bf16 GEMM operands <- fp32 parameters, all groups, on the current stream (after every un-graphed step; call it after
        editing parameters by hand)"""
        for g in range(4):
            self.engine.pack_group(g)
        self._packed = True
        self._pv = getattr(self.model, "_param_version", 0)

    def _update(self, ph, lo, hi):
        """(communication stream) all-reduce of gradient range [lo, hi), its Adam step, and the repack of the operands built
        from those parameters -- all beside the later backward phases, which read packed operands and fp32 GroupNorm
        parameters of THEIR OWN blocks only.  The level-wide FiLM operands are the exception (the first part of phase 2, the
        FiLM data gradient, reads all of them): step_resident repacks them when that part has finished."""
        g, gb, scale = self.model._grad_flat, None, 1.0
        if self.world > 1:
            if self.grad_comm_dtype == torch.float32:
                parallel.allreduce_mean_(g[lo:hi])                     # NCCL over NVLink, concurrent with the next phase
            else:       # the Adam kernel reads the reduced bf16 buffer in place (the fp32 arena keeps this rank's own gradients)
                if self._gcomm is None:
                    self._gcomm = torch.empty(max(h_ - l_ for l_, h_ in self.ranges.values()), dtype=self.grad_comm_dtype, device=self.dev)
                gb = self._gcomm[:hi - lo]
                gb.copy_(g[lo:hi])
                parallel.allreduce_sum_(gb)                            # the 1/world is Adam's grad_scale
                scale = 1.0 / self.world
        self.opt.step(grad_scale=scale, lo=lo, hi=hi, first=ph == 0, grad=gb)
        self.engine.pack_group(ph)

    def _set_batch(self, batch):
        img, R, T, K = batch
        self.img.copy_(img, non_blocking=True)
        self.R.copy_(R, non_blocking=True)
        self.T.copy_(T, non_blocking=True)
        self.K.copy_(K, non_blocking=True)

    def _capture(self, fn):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g

    def step_resident(self, batch):
        """batch already lives in HBM (device tensors)"""
        from . import _lib
        self._set_batch(batch)
        if self._packed and getattr(self.model, "_param_version", 0) != self._pv:
            self.repack()          # the parameters changed outside this trainer's own update (load_state_dict, a foreign optimizer)
        if self.use_graph and self.graph is None and self.step_idx >= 1:
            torch.cuda.synchronize()
            n0 = _lib.launch_count()
            # four graphs: the update of each finished gradient range (all-reduce, Adam, repack) overlaps the next phase
            self.model._grad_views()
            self.graph = [self._capture(lambda: (self._draws_forward_loss(), self._backward(0))), self._capture(lambda: self._backward(1)),
                          self._capture(lambda: self._backward("2a")), self._capture(lambda: self._backward("2b"))]
            self.ranges = self._phase_ranges()
            self.comm = torch.cuda.Stream(self.dev)
            self.launches_per_step = _lib.launch_count() - n0     # + the un-captured update launches, counted on the first replay
            self._count_update = True
            self.graphed = True
        if self.graph is None:
            self._compute()
            self._finish()
            return
        main = torch.cuda.current_stream(self.dev)
        n_up = _lib.launch_count()
        for ph in (0, 1, 2):
            self.graph[ph].replay()
            self.comm.wait_stream(main)                                # gradients of this phase are final
            if ph == 2:       # "2a" held the last reads of the level-wide FiLM operands: repack them beside "2b"
                with torch.cuda.stream(self.comm):
                    self.engine.pack_group(3)
                self.graph[3].replay()
                self.comm.wait_stream(main)
            lo, hi = self.ranges[ph]
            with torch.cuda.stream(self.comm):
                self._update(ph, lo, hi)
        main.wait_stream(self.comm)                                    # the next forward reads the repacked operands
        self._pv = getattr(self.model, "_param_version", 0)            # (FusedAdam.step bumped it: this update's operands are current)
        if self._count_update:                                         # Adam per range, repack per group, composed attention operands
            self.launches_per_step += _lib.launch_count() - n_up
            self._count_update = False
        self.step_idx += 1

    def step_e2e(self, host_batch):
        """the same step through the copy path: pinned host inputs -> device, loss read back"""
        self.step_resident(host_batch)
        return float(self.loss.item())

    # ------------------------------------------------------------------ profiled step (per-kernel CUDA events)
    def profile_step(self, batch, dump=None):
        self._set_batch(batch)
        ops = self.engine.ops
        ops.prof = []
        self.ops.prof = ops.prof
        try:
            self._compute()
            torch.cuda.synchronize()
        finally:
            rec, ops.prof, self.ops.prof = ops.prof, None, None
        out, total = {}, 0.0
        if dump:
            import json
            json.dump([dict(name=n_, ms=s_.elapsed_time(e_), flop=f_, info=i_) for n_, s_, e_, f_, i_ in rec], open(dump, "w"))
        for name, s, e, flop, _ in rec:
            grp = "b3d_gemm" if name == "b3d_gemm" else ("groupnorm" if name.startswith("b3d_gn") else "elementwise")
            d = out.setdefault(grp, {"ms": 0.0, "n": 0, "flop": 0.0})
            ms = s.elapsed_time(e)
            d["ms"] += ms
            d["n"] += 1
            d["flop"] += flop
            total += ms
        out["total_ms"] = total
        self._finish()
        return out
