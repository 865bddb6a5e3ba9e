"""-m gpu: every CUDA kernel against the plain-torch fp32 statement of the same op (tests/emu_ops.py),
called through the C ABI (ops.CudaOps -> libb3d.so)."""
import pytest
import torch

import gemm_cases
import op_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def backends():
    from distributed_3d_diffusion_pytorch_b200.ops import CudaOps
    from emu_ops import EmuOps
    return CudaOps("cuda:0"), EmuOps("cpu", stats_dtype=torch.bfloat16)


def _cmp(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert torch.isfinite(a).all()
    return (a - b).norm().item() / (b.norm().item() + 1e-12)


@pytest.mark.parametrize("case", gemm_cases.all_cases(), ids=lambda c: c.name)
def test_gemm(case, backends):
    cuda, emu = backends
    d_c, outs_c = case.build(cuda)
    d_e, outs_e = case.build(emu)
    cuda.gemm(d_c)
    torch.cuda.synchronize()
    emu.gemm(d_e)
    for a, b in zip(outs_c, outs_e):
        assert _cmp(a, b) <= (8e-3 if a.dtype == torch.bfloat16 else 2e-4)


@pytest.mark.parametrize("name", list(op_cases.all_cases().keys()))
def test_op(name, backends):
    cuda, emu = backends
    fn = op_cases.all_cases()[name]
    oc = fn(cuda)
    torch.cuda.synchronize()
    oe = fn(emu)
    for i, (a, b) in enumerate(zip(oc, oe)):
        tol = 8e-3 if a.dtype == torch.bfloat16 else 2e-4
        if name.startswith("pose_fp32"):
            tol = 1e-5           # north_star: ray / posenc outputs within 1e-5 relative
        assert _cmp(a, b) <= tol, (name, i)
