"""Runs every kernel case on the GPU against the fp32 torch statements (tests/emu_ops.py) and prints
one line per case; never stops at the first failure (GPU round trips are expensive).
    python tests/gpu_diag.py [filter] > gpurun_out/diag.log"""
import json
import os
import sys
import time
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import gemm_cases  # noqa: E402
import op_cases  # noqa: E402
from emu_ops import EmuOps  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.ops import CudaOps  # noqa: E402


def cmp(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    if not torch.isfinite(a).all():
        return float("inf"), float("inf")
    den = b.norm().item() + 1e-12
    return (a - b).norm().item() / den, (a - b).abs().max().item()


def tol_for(t):
    return 8e-3 if t.dtype == torch.bfloat16 else 2e-4


def run_gemm_case(case, cuda, emu):
    d_c, outs_c = case.build(cuda)
    d_e, outs_e = case.build(emu)
    cuda.gemm(d_c)
    torch.cuda.synchronize()
    emu.gemm(d_e)
    res = [cmp(a, b) for a, b in zip(outs_c, outs_e)]
    ok = all(r[0] <= tol_for(a) for r, a in zip(res, outs_c))
    return ok, res


def run_op_case(fn, cuda, emu):
    oc = fn(cuda)
    torch.cuda.synchronize()
    oe = fn(emu)
    res = [cmp(a, b) for a, b in zip(oc, oe)]
    ok = all(r[0] <= tol_for(a) for r, a in zip(res, oc))
    return ok, res


def main():
    flt = sys.argv[1] if len(sys.argv) > 1 else ""
    cuda, emu = CudaOps("cuda:0"), EmuOps("cpu", stats_dtype=torch.bfloat16)
    summary = {}
    nfail = 0
    items = [(c.name, lambda c=c: run_gemm_case(c, cuda, emu)) for c in gemm_cases.all_cases()]
    items += [(k, lambda f=f: run_op_case(f, cuda, emu)) for k, f in op_cases.all_cases().items()]
    for name, thunk in items:
        if flt and flt not in name:
            continue
        t0 = time.time()
        try:
            ok, res = thunk()
            msg = " ".join(f"[{r[0]:.2e},{r[1]:.2e}]" for r in res)
        except Exception as e:  # noqa: BLE001
            ok, msg = False, "EXC " + repr(e)[:300]
            if os.environ.get('B3D_DIAG_TRACE'):
                traceback.print_exc()
            if "CUDA error" in repr(e) or "illegal" in repr(e) or "launch failure" in repr(e):
                print(f"FAIL {name}: {msg}", flush=True)
                print("sticky CUDA error: stopping", flush=True)
                summary[name] = msg
                nfail += 1
                break
        nfail += (not ok)
        summary[name] = ("ok " if ok else "FAIL ") + msg
        print(f"{'ok  ' if ok else 'FAIL'} {name:42s} {msg}  ({time.time() - t0:.1f}s)", flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(summary, open("gpurun_out/diag.json", "w"), indent=1)
    print(f"{nfail} failures of {len(summary)}")
    return 1 if nfail else 0


if __name__ == "__main__":
    sys.exit(main())
