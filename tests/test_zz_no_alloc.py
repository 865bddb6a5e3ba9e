"""-m gpu, collected last on purpose (file name): the allocation-free loss entry point `b3d_loss_static`.

Some hosts of this library cannot give a kernel scratch memory: a trainer that split the whole HBM into its
own arenas up front, or a service that bans cudaMalloc/cudaFree after start-up because they stall every
stream.  For them the diffusion loss also exists with its reduction scratch inside the library image.  This
test calls that entry point straight through the C ABI (ctypes on libb3d.so, no wrapper) and checks loss and
gradient against the ordinary `b3d_loss` on two consecutive calls (the second one reuses the shared scratch
with different data), then checks that an oversized request is refused with a readable message."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-4          # fp32 outputs, same bound as the other kernel tests (summation order differs between the two grids)


def _lib():
    from distributed_3d_diffusion_pytorch_b200 import _lib
    lib = _lib.load()
    lib.b3d_loss_static_max_elems.restype = ctypes.c_int64
    return lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("loss_type", [0, 1, 2])
def test_static_scratch_loss_matches_caller_scratch_loss_twice(loss_type):
    lib = _lib()
    dev = torch.device("cuda:0")
    scratch = torch.zeros(int(lib.b3d_loss_scratch_bytes()) // 4, dtype=torch.float32, device=dev)
    g = torch.Generator().manual_seed(5 + loss_type)
    for call, shape in enumerate([(16, 3, 64, 64), (3, 3, 24, 24)]):         # two calls in a row, different sizes and data
        noise = torch.randn(shape, generator=g).to(dev)
        pred = (torch.randn(shape, generator=g) * 1.5).to(dev)
        n = noise.numel()
        ref_loss, ref_d = torch.zeros(1, device=dev), torch.empty_like(pred)
        rc = lib.b3d_loss(_p(noise), _p(pred), _p(ref_d), ctypes.c_int64(n), loss_type, ctypes.c_float(1.0), _p(scratch), _p(ref_loss), _stream())
        assert rc == 0, lib.b3d_last_error().decode()
        out_loss, out_d = torch.full((1,), float("nan"), device=dev), torch.full_like(pred, float("nan"))
        rc = lib.b3d_loss_static(_p(noise), _p(pred), _p(out_d), ctypes.c_int64(n), loss_type, ctypes.c_float(1.0), _p(out_loss), _stream())
        assert rc == 0, lib.b3d_last_error().decode()
        torch.cuda.synchronize()
        assert abs(out_loss.item() - ref_loss.item()) <= TOL * abs(ref_loss.item()), (call, out_loss.item(), ref_loss.item())
        assert torch.equal(out_d, ref_d), call
        # and against torch, so both entry points are anchored to the definition (train.py:102-112)
        fn = [torch.nn.functional.mse_loss, torch.nn.functional.l1_loss, torch.nn.functional.smooth_l1_loss][loss_type]
        assert abs(out_loss.item() - fn(noise, pred).item()) <= TOL * abs(ref_loss.item())


def test_static_scratch_loss_refuses_oversize():
    lib = _lib()
    dev = torch.device("cuda:0")
    mx = int(lib.b3d_loss_static_max_elems())
    assert mx == 1024 * 4096
    a = torch.zeros(8, device=dev)
    out = torch.zeros(1, device=dev)
    rc = lib.b3d_loss_static(_p(a), _p(a), None, ctypes.c_int64(mx + 1), 0, ctypes.c_float(1.0), _p(out), _stream())
    assert rc != 0
    msg = lib.b3d_last_error().decode()
    assert "maximum" in msg and str(mx) in msg, msg
    # the largest accepted size still works
    big = torch.randn(mx, device=dev)
    rc = lib.b3d_loss_static(_p(big), _p(big * 0.5), None, ctypes.c_int64(mx), 0, ctypes.c_float(1.0), _p(out), _stream())
    assert rc == 0, lib.b3d_last_error().decode()
    torch.cuda.synchronize()
    assert abs(out.item() - (0.25 * big * big).mean().item()) <= TOL * out.item()
