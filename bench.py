"""Headline benchmark: X-UNet (CFG-64: H=W=64, ch=128) training step, global batch 128, bf16 tensor-core
path, synthetic SRN-shaped data, random-init weights.  One process per GPU (torchrun for N>1).

    python bench.py --gpus N --steps K --warmup W            # this implementation
    python bench.py --impl reference --steps K --warmup W     # the reference path on the host cores (CPU)

A step = q_sample + conditioning-frame replacement, XUNet forward, loss, hand-written backward, gradient
allreduce (N>1), fused Adam.  Prints ONE JSON line (see README / DESIGN.md for the keys)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "xunet_train_samples_per_sec_64"
UNIT = "samples/s"
MODEL_KW = dict(H=64, W=64, ch=128)
TRAIN_FLOP_PER_SAMPLE = 707.7e9          # 3 x 235.9 GFLOP forward


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region"""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_step_time(B, steps, warmup, threads=None):
    """the reference algorithm (oracle port, fp32, torch CPU) on the host cores: one training step = fwd + loss + bwd"""
    from oracle import diffusion_ref as dr, xunet_ref as xr
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    # all host cores: torchrun exports OMP_NUM_THREADS=1, which would cripple this arm at N > 1
    torch.set_num_threads(threads or os.cpu_count() or 1)
    cfg = xr.Cfg(**MODEL_KW)
    sd = {k: v.requires_grad_(True) for k, v in xr.init_state_dict(cfg, seed=0).items()}
    img, R, T, K = synthetic.make_views(B, MODEL_KW["H"], seed=1)
    g = torch.Generator().manual_seed(2)
    times = []
    for it in range(warmup + steps):
        t0 = time.time()
        logsnr = dr.logsnr_schedule_cosine(torch.rand(B, generator=g))
        noise = torch.randn(img[:, 0].shape, generator=g)
        cm = torch.rand(B, generator=g) > 0.1
        x = torch.where(cm[:, None, None, None], img[:, 0], torch.randn(img[:, 0].shape, generator=g))
        batch = dr.make_batch(x, logsnr, dr.q_sample(img[:, 1], logsnr, noise), R, T, K)
        for v in sd.values():
            v.grad = None
        loss = dr.loss_fn(noise, xr.xunet_forward(sd, cfg, batch, cm, training=True))
        loss.backward()
        if it >= warmup:
            times.append(time.time() - t0)
    return sum(times) / len(times), torch.get_num_threads()


def gpu_eager_baseline(dev, batch=64, steps=5, warmup=2):
    """The bar that matters beside the CPU number: the reference's eager PyTorch path
    on the SAME B200 -- the pinned port of its module (oracle/xunet_ref.xunet_forward, plain torch ops on cuDNN / cuBLAS)
    + loss + autograd backward, (i) fp32 with TF32 off, (ii) under torch.autocast(bfloat16).  /root/reference itself does
    not travel to the GPU box, hence kind "port".  Batch: the largest power of two <= `batch` that fits."""
    from oracle import diffusion_ref as dr, xunet_ref as xr
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    cfg = xr.Cfg(**MODEL_KW)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    out = {"kind": "port", "unit": UNIT, "steps": steps, "warmup": warmup,
           "what": "oracle port of xunet.py + mse + autograd backward, eager torch on cuDNN/cuBLAS, no optimizer step"}
    sd = {k: v.to(dev).requires_grad_(True) for k, v in xr.init_state_dict(cfg, seed=0).items()}
    for key, autocast in (("fp32", False), ("bf16_autocast", True)):
        B = batch
        while B >= 1:
            try:
                img, R, T, K = synthetic.make_views(B, MODEL_KW["H"], seed=1, device=dev)
                g = torch.Generator(device=dev).manual_seed(2)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                for it in range(warmup + steps):
                    if it == warmup:
                        torch.cuda.synchronize()
                        e0.record()
                    logsnr = dr.logsnr_schedule_cosine(torch.rand(B, generator=g, device=dev))
                    noise = torch.randn(img[:, 0].shape, generator=g, device=dev)
                    cm = torch.rand(B, generator=g, device=dev) > 0.1
                    x = torch.where(cm[:, None, None, None], img[:, 0], torch.randn(img[:, 0].shape, generator=g, device=dev))
                    b_ = dr.make_batch(x, logsnr, dr.q_sample(img[:, 1], logsnr, noise), R, T, K)
                    for v in sd.values():
                        v.grad = None
                    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                        eps = xr.xunet_forward(sd, cfg, b_, cm, training=True)
                    loss = dr.loss_fn(noise, eps.float())
                    loss.backward()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / steps
                out[key] = {"value": B / (ms / 1e3), "ms_per_step": ms, "batch": B}
                break
            except torch.cuda.OutOfMemoryError:
                B //= 2
            finally:
                for v in sd.values():
                    v.grad = None
                torch.cuda.empty_cache()
        else:
            out[key] = None
    return out


def bench_cfg128(dev, args, batch=32, steps=3, warmup=3):
    """BASELINE configs[2] on one GPU: XUNet() defaults (H=W=128, ch=256), bf16, the same training step, at the largest
    per-GPU batch of the 8-GPU run that matters here (32 = global 128 over 4 GPUs fits one B200 with room to spare)."""
    from distributed_3d_diffusion_pytorch_b200 import synthetic
    from distributed_3d_diffusion_pytorch_b200.trainer import Trainer
    kw = dict(H=128, W=128, ch=256)
    tr = Trainer(kw, batch, dev, world=1, seed=4321, use_graph=not args.no_graph)
    res = [tuple(t.to(dev) for t in synthetic.make_views(batch, 128, seed=50 + i)) for i in range(2)]
    for i in range(warmup):
        tr.step_resident(res[i % 2])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        tr.step_resident(res[i % 2])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    prof = tr.profile_step(res[0])
    achieved = prof["b3d_gemm"]["flop"] / (prof["b3d_gemm"]["ms"] / 1e3) / 1e12
    pk = peaks()
    out = {"metric": "xunet_train_samples_per_sec_128", "value": batch / (ms / 1e3), "unit": UNIT, "ms_per_step": ms, "batch": batch, "steps": steps,
           "warmup": warmup, "whole_step_tflops": 8921e9 * batch / (ms / 1e3) / 1e12, "gemm_tflops": achieved, "gemm_frac_of_sustained_peak": achieved / pk["tf_sustained"],
           "breakdown_ms": {k: round(v["ms"], 3) for k, v in prof.items() if isinstance(v, dict)}, "loss": float(tr.loss.item())}
    del tr, res
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = args.ref_batch
    t, cores = cpu_reference_step_time(B, args.steps, args.warmup)
    v = B / t
    sample = f"CFG-64 fp32 fwd+bwd, batch {B} per step, {args.steps} timed steps after {args.warmup} warm-up (oracle port of the reference path; /root/reference is Python and does not travel to the GPU box)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg64_train_step", "global_batch": B, "device": "cpu"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b3d")
    ap.add_argument("--global-batch", type=int, default=128)
    ap.add_argument("--ref-batch", type=int, default=4, help="BASELINE configs[0]: batch 4 on the CPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-sampler", action="store_true")
    ap.add_argument("--fp32-allreduce", action="store_true", help="all-reduce the gradients in fp32 (default: bf16 payload)")
    ap.add_argument("--no-extras", action="store_true", help="skip the CFG-128 extra line and the eager-GPU baseline (N = 1 only)")
    ap.add_argument("--model", default="cfg64", choices=["cfg64", "cfg128"], help="cfg128 = XUNet() defaults (H=W=128, ch=256), BASELINE configs[2]")
    ap.add_argument("--profile-json", default=None, help="dump the per-launch CUDA-event table of the profiled step")
    args = ap.parse_args()
    global MODEL_KW, TRAIN_FLOP_PER_SAMPLE, METRIC
    if args.model == "cfg128":
        MODEL_KW, TRAIN_FLOP_PER_SAMPLE, METRIC = dict(H=128, W=128, ch=256), 8921e9, "xunet_train_samples_per_sec_128"
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch.distributed as dist
    from distributed_3d_diffusion_pytorch_b200 import _lib, synthetic
    from distributed_3d_diffusion_pytorch_b200.trainer import Trainer

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, (world, args.gpus)
    B = args.global_batch // world
    tr = Trainer(MODEL_KW, B, dev, world=world, seed=1234, use_graph=not args.no_graph,
                 grad_comm_dtype=torch.float32 if args.fp32_allreduce else torch.bfloat16)      # (the trainer adds the rank to the seed)

    # synthetic SRN-shaped data: a pool of host batches (pinned) and device-resident copies of the same
    pool = [synthetic.make_views(B, MODEL_KW["H"], seed=100 * rank + i) for i in range(4)]
    pinned = [tuple(t.pin_memory() for t in b) for b in pool]
    resident = [tuple(t.to(dev) for t in b) for b in pool]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing (value)
    for i in range(args.warmup):
        tr.step_resident(resident[i % len(resident)])
    barrier()
    n0 = _lib.launch_count()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        tr.step_resident(resident[i % len(resident)])
    e1.record()
    barrier()
    clk = clocks.stop() if rank == 0 else None
    launches = (_lib.launch_count() - n0) if not tr.graphed else tr.launches_per_step * args.steps
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    value = args.global_batch * args.steps / (ms / 1e3)

    # ---------------- end to end through the public API with host buffers (e2e)
    h2d = sum(t.numel() * t.element_size() for t in pinned[0])
    for i in range(2):
        tr.step_e2e(pinned[i % len(pinned)])
    barrier()
    e0.record()
    for i in range(args.steps):
        loss_val = tr.step_e2e(pinned[i % len(pinned)])
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = args.global_batch * args.steps / (t.item() / 1e3)

    # ---------------- every rank must hold the same parameters after the timed steps (outside the timed region): the proof
    # that the three all-reduced gradient ranges cover the whole arena and that the ranks never drifted apart
    checksum_equal = None
    if world > 1:
        cs = tr.model.flat_parameters().double().sum().reshape(1)
        allcs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        checksum_equal = bool(all(torch.equal(c, allcs[0]) for c in allcs))
        assert checksum_equal, f"rank parameters diverged: {[c.item() for c in allcs]}"

    # ---------------- roofline of the dominant kernel (tcgen05 implicit GEMM), measured live with CUDA events
    roof = None
    pk = peaks()
    if not args.no_profile:
        prof = tr.profile_step(resident[0], dump=args.profile_json)
        gemm_ms = prof["b3d_gemm"]["ms"]
        gemm_flop = prof["b3d_gemm"]["flop"]
        achieved = gemm_flop / (gemm_ms / 1e3) / 1e12
        traffic = None
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("gemm_dram_bytes_per_launch")
        roof = {"bound": "tensor", "achieved": achieved, "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sustained"],
                "traffic": traffic, "traffic_source": "profiles/roofline_traffic.json (one ncu --set full capture of one conv launch, offline; NOT measured by this run)",
                "kernel": "b3d_gemm_kernel", "launches_per_step": prof["b3d_gemm"]["n"],
                "kernel_share_of_step": gemm_ms / prof["total_ms"], "peak_source": pk["src"] + " (sustained cuBLAS bf16)",
                "whole_step_tflops": TRAIN_FLOP_PER_SAMPLE * B / (ms / args.steps / 1e3) / 1e12,
                "breakdown_ms": {k: round(v["ms"], 3) for k, v in prof.items() if isinstance(v, dict)}}

    # ---------------- second half of the metric: stochastic-conditioning sampler, denoise steps/s (one object per rank)
    sampler, gs = None, None
    if not args.no_sampler:
        from distributed_3d_diffusion_pytorch_b200.sampling import GraphSampler
        import random
        gs = GraphSampler(tr.model, 8)
        nv = 4                                             # the given view + 3 synthesised ones
        img, R, T, K = synthetic.make_views(nv, MODEL_KW["H"], seed=7 + rank)
        w = torch.arange(8, dtype=torch.float32)
        record = [[img[0, 0][None].repeat(8, 1, 1, 1).numpy(), R[0, 0].numpy(), T[0, 0].numpy()]]
        random.seed(11 + rank)
        gs.sample(record, R[1, 0], T[1, 0], K[0], w, timesteps=6)      # warm-up: eager step, graph capture, replays
        barrier()
        nst = 256
        e0.record()
        for v in range(1, nv):                             # sampling.py:158-176: the record grows by one view per pass
            out = gs.sample(record, R[v, 0], T[v, 0], K[0], w, timesteps=nst)
            record.append([out, R[v, 0].numpy(), T[v, 0].numpy()])
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_steps = nst * (nv - 1)
        sampler = {"metric": "sampler_denoise_steps_per_sec_64", "value": world * total_steps / (t.item() / 1e3), "unit": "steps/s",
                   "ms_per_step": t.item() / total_steps, "finite": bool((out == out).all()),
                   "config": f"one object per GPU, {nv - 1} target views x {nst} steps, stochastic conditioning on a growing record (random.choice per step), "
                             "8 guidance weights x (cond, uncond) = one batch-16 forward + fused update per step, CUDA graph, conditioning convs cached per "
                             "camera, FiLM evaluated for the 4 distinct conditionings, record on the device, dropout on (as sampling.py); includes the "
                             "per-view read-back of the image"}

    # ---------------- N = 1 extras, after the trainer's memory is released: BASELINE configs[2] (CFG-128 train step) and the
    # reference's eager GPU path on this same B200
    eager, cfg128, graphed = None, None, tr.graphed
    if rank == 0 and world == 1 and args.model == "cfg64" and not args.no_extras:
        del gs, tr, resident
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        cfg128 = bench_cfg128(dev, args)
        gc.collect()
        torch.cuda.empty_cache()
        eager = gpu_eager_baseline(dev)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tcpu, cores = cpu_reference_step_time(args.ref_batch, 2, 1)
        cpu = {"value": args.ref_batch / tcpu, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"CFG-64 fp32 fwd+bwd, batch {args.ref_batch}, 2 timed steps after 1 warm-up (oracle port of the reference path)"}

    if rank == 0:
        print("peak device memory: %.1f GB" % (torch.cuda.max_memory_allocated(dev) / 2 ** 30), file=sys.stderr)
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": value / 29.9,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": args.model + "_train_step", "model": "XUNet(H=%d,W=%d,ch=%d)" % (MODEL_KW["H"], MODEL_KW["W"], MODEL_KW["ch"]), "global_batch": args.global_batch, "per_gpu_batch": B,
                       "parallelism": f"dp{world}", "grad_allreduce": ("fp32 avg" if args.fp32_allreduce else "bf16 sum (1/N in Adam's gradient scale)") if world > 1 else None, "dropout": 0.1, "optimizer": "fused Adam", "cuda_graph": graphed,
                       "l2": "no flush needed: per-step working set (activations) is GBs, far above the 126 MB L2",
                       "baseline_note": "vs_baseline = value / 29.9 samples/s (README.md:39, 8x RTX 3090, derived)"},
            "clocks": clk, "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "last_loss": loss_val},
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "sampler": sampler,
            "param_checksum_equal": checksum_equal, "gpu_eager_baseline": eager, "cfg128": cfg128,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
