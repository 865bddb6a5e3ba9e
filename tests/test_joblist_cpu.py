"""Host logic of the one-launch weight repack: ops.JobList groups the recorded jobs by source tensor into b3d_pack_group
entries (include/b3d.h) and builds the block -> group table the kernel reads.  No GPU needed."""
import ctypes

import torch

from distributed_3d_diffusion_pytorch_b200.ops import JobList, _PackDst, _PackGroup, _PackJob


def _rec(jl, w, mode, out, ci0=0, cn=None, Cpad=None, k_off=0):
    Co, Ci = w.shape[0], w.shape[1]
    taps = w.numel() // (Co * Ci)
    cn = Ci if cn is None else cn
    jl.add(w, out, Co, Ci, taps, mode, ci0, cn, Cpad, k_off, out.stride(0), 0)


def test_struct_layouts_match_the_header():
    # b3d_pack_dst: ptr + 4 x int32 + 2 x int64; b3d_pack_group: ptr + 4 x int32 + 2 x int64 + 8 destinations
    assert ctypes.sizeof(_PackDst) == 8 + 16 + 16
    assert ctypes.sizeof(_PackGroup) == 8 + 16 + 16 + 8 * ctypes.sizeof(_PackDst)
    assert ctypes.sizeof(_PackJob) == 16 + 32 + 32


def test_jobs_of_one_source_become_one_group_read_once():
    jl = JobList("pack")
    w3 = torch.zeros(40, 70, 3, 3)                       # 2 x 3 tiles of 32 x 32
    w1 = torch.zeros(64, 64)
    f, t, f2 = torch.zeros(40, 9 * 128), torch.zeros(70, 9 * 64), torch.zeros(40, 9 * 128 + 128)
    _rec(jl, w3, 0, f, Cpad=128)
    _rec(jl, w1, 0, torch.zeros(64, 64), Cpad=64)
    _rec(jl, w3, 1, t, Cpad=64)
    jl.add(torch.zeros(3000), torch.zeros(3000), 0, 0, 1, 3, 0, 0, 0, 0, 0, 3000)      # fp32 copy: its own group
    _rec(jl, w3, 0, f2, ci0=6, cn=64, Cpad=128, k_off=128)
    groups, block_group = jl.groups()
    assert len(groups) == 3
    g0, g1, g2 = groups
    assert (g0.src, g0.Co, g0.Ci, g0.taps, g0.n_dst, g0.block0) == (w3.data_ptr(), 40, 70, 9, 3, 0)
    assert [g0.d[k].mode for k in range(3)] == [0, 1, 0] and [g0.d[k].dst for k in range(3)] == [f.data_ptr(), t.data_ptr(), f2.data_ptr()]
    assert (g0.d[2].ci0, g0.d[2].cn, g0.d[2].Cpad, g0.d[2].k_off, g0.d[2].ld) == (6, 64, 128, 128, f2.stride(0))
    assert (g1.n_dst, g1.block0, g1.taps) == (1, 6, 1) and (g2.d[0].mode, g2.total, g2.block0) == (3, 3000, 6 + 4)
    assert block_group == [0] * 6 + [1] * 4 + [2] * 3            # ceil(40/32)*ceil(70/32), 2*2, ceil(3000/1024)


def test_more_than_eight_destinations_split_into_two_groups():
    jl = JobList("pack")
    w = torch.zeros(32, 32)
    for _ in range(11):
        _rec(jl, w, 0, torch.zeros(32, 64), Cpad=64)
    groups, block_group = jl.groups()
    assert [g.n_dst for g in groups] == [8, 3] and [g.block0 for g in groups] == [0, 1] and block_group == [0, 1]
