"""Tensor-level bindings of the b3d C ABI (include/b3d.h): one method = one entry point.

torch is plumbing here (device memory, streams); all arithmetic happens in libb3d.so.  Tensors are
passed as (data_ptr, strides) -- any view whose innermost stride is 1 is accepted, so channel
slices / head splits need no copies.  There is no fallback path: without the CUDA library every
method raises (see _lib.load)."""
import contextlib
import ctypes
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from . import _lib

c_void = ctypes.c_void_p
c_i32 = ctypes.c_int32
c_i64 = ctypes.c_int64
c_f32 = ctypes.c_float


class _View4(ctypes.Structure):
    _fields_ = [("ptr", c_void), ("dim", c_i32 * 4), ("stride", c_i64 * 3)]


class _GemmDesc(ctypes.Structure):
    _fields_ = [
        ("a", _View4 * 3), ("a_taps", c_i32 * 3), ("n_src", c_i32),
        ("b", _View4),
        ("a_major", c_i32), ("b_major", c_i32),
        ("conv_stride", c_i32), ("pad", c_i32), ("b_taps", c_i32),
        ("W", c_i32), ("H", c_i32), ("N", c_i32),
        ("m_rows", c_i32), ("n_cols", c_i32),
        ("batched", c_i32), ("b_batch_xor", c_i32), ("out_batch_xor", c_i32), ("split_k", c_i32),
        ("out", c_void), ("out_fp32", c_i32), ("out_stride", c_i64 * 3), ("out_tap_stride", c_i64),
        ("out2", c_void), ("bias", c_void), ("rowvec", c_void), ("rowvec_ld", c_i64), ("resid", c_void),
        ("alpha", c_f32), ("out_scale", c_f32), ("act", c_i32), ("out_atomic", c_i32), ("bn_tile", c_i32),
        ("gn_sums", c_void), ("gn_cpg", c_i32), ("gn_rows_per_image", c_i32),
    ]


@dataclass
class GemmDesc:
    """Python mirror of b3d_gemm_desc.  Operand/output tensors are 4-D views [n, y, x, c] with
    stride(c) == 1 (see include/b3d.h for the meaning of the majors)."""
    a: List[torch.Tensor]
    b: torch.Tensor
    out: torch.Tensor                       # [n, y, rows(x), cols]
    n_cols: int
    a_taps: List[int] = field(default_factory=lambda: [1])
    a_major: int = 0
    b_major: int = 0
    conv_stride: int = 1
    pad: int = 0
    b_taps: int = 1
    W: int = 0
    H: int = 1
    N: int = 1
    m_rows: int = 0
    batched: bool = False
    b_batch_xor: int = 0
    out_batch_xor: int = 0
    split_k: int = 1
    out_tap_stride: int = 0
    out2: Optional[torch.Tensor] = None
    bias: Optional[torch.Tensor] = None
    rowvec: Optional[torch.Tensor] = None   # fp32 [N, >=n_cols]
    resid: Optional[torch.Tensor] = None    # same shape/strides as out
    alpha: float = 1.0
    out_scale: float = 1.0
    act: int = 0
    out_atomic: bool = False
    bn_tile: int = 0
    # GroupNorm (sum, sumsq) of the bf16 output taken in the epilogue: fp32 [images, 32, 2], accumulated (see include/b3d.h)
    gn_sums: Optional[torch.Tensor] = None
    gn_cpg: int = 0
    gn_rows_per_image: int = 0


def _ptr(t):
    return c_void(t.data_ptr()) if t is not None else c_void(0)


def _view4(t):
    assert t.dim() == 4 and (t.stride(3) == 1 or t.shape[3] == 1), (t.shape, t.stride())
    v = _View4()
    v.ptr = t.data_ptr()
    n, y, x, c = t.shape
    v.dim[0], v.dim[1], v.dim[2], v.dim[3] = c, x, y, n
    v.stride[0], v.stride[1], v.stride[2] = t.stride(2), t.stride(1), t.stride(0)
    return v


def _u8(mask):
    return mask if mask.dtype == torch.uint8 else mask.to(torch.uint8)


class _PackJob(ctypes.Structure):
    _fields_ = [("src", c_void), ("dst", c_void), ("Co", c_i32), ("Ci", c_i32), ("taps", c_i32), ("mode", c_i32), ("ci0", c_i32),
                ("cn", c_i32), ("Cpad", c_i32), ("accumulate", c_i32), ("k_off", c_i64), ("ld", c_i64), ("total", c_i64), ("block0", c_i64)]


class _PackDst(ctypes.Structure):
    _fields_ = [("dst", c_void), ("mode", c_i32), ("ci0", c_i32), ("cn", c_i32), ("Cpad", c_i32), ("k_off", c_i64), ("ld", c_i64)]


class _PackGroup(ctypes.Structure):
    _fields_ = [("src", c_void), ("Co", c_i32), ("Ci", c_i32), ("taps", c_i32), ("n_dst", c_i32), ("total", c_i64), ("block0", c_i64),
                ("d", _PackDst * 8)]


class JobList:
    """Recorded pack / unpack jobs (see b3d_pack_job) that run as ONE kernel launch.  Built once per engine: parameter,
    gradient and staging buffers have fixed addresses, so the device-side table is reused every step.
    Pack lists: the jobs that read the same source tensor become one b3d_pack_group (the source is read once for all its
    destinations); pad columns are not written by the kernel, so the recorder zeroes every destination once (`zero_pads`)."""

    def __init__(self, kind):
        self.kind, self.jobs, self.keep, self.table, self.blocks = kind, [], [], None, 0
        self.block_table, self.n_entries, self.zeroed = None, 0, set()

    def groups(self):
        """(ctypes array of _PackGroup, block -> group list): jobs grouped by source, first-appearance order, <= 8 each"""
        by_src, order = {}, []
        for j in self.jobs:
            key = (j.src, j.Co, j.Ci, j.taps) if j.mode != 3 else ("copy", len(order))
            lst = by_src.get(key)
            if lst is None or len(lst[-1]) == 8:
                if lst is None:
                    by_src[key] = lst = []
                lst.append([])
                order.append(lst[-1])
            lst[-1].append(j)
        out, block_group, blocks = [], [], 0
        for gi, js in enumerate(order):
            g, j0 = _PackGroup(), js[0]
            g.src, g.Co, g.Ci, g.taps, g.n_dst, g.total, g.block0 = j0.src, j0.Co, j0.Ci, j0.taps, len(js), j0.total, blocks
            for k, j in enumerate(js):
                g.d[k].dst, g.d[k].mode, g.d[k].ci0, g.d[k].cn, g.d[k].Cpad, g.d[k].k_off, g.d[k].ld = j.dst, j.mode, j.ci0, j.cn, j.Cpad, j.k_off, j.ld
            nb = (j0.total + 1023) // 1024 if j0.mode == 3 else ((j0.Co + 31) // 32) * ((j0.Ci + 31) // 32)
            block_group += [gi] * nb
            blocks += nb
            out.append(g)
        return (_PackGroup * len(out))(*out), block_group

    def add(self, src, dst, Co, Ci, taps, mode, ci0, cn, Cpad, k_off, ld, total, accumulate=0):
        j = _PackJob()
        j.src, j.dst = src.data_ptr(), dst.data_ptr()
        j.Co, j.Ci, j.taps, j.mode, j.ci0, j.cn, j.Cpad, j.accumulate = Co, Ci, taps, mode, ci0, cn, Cpad or 0, int(accumulate)
        j.k_off, j.ld, j.total, j.block0 = k_off, ld, total, self.blocks
        if self.kind == "pack" and mode != 3:      # one 32 x 32 (x taps) source tile per block (pack_multi_kernel)
            t32 = lambda v: (v + 31) // 32
            self.blocks += {0: t32(Co) * t32(j.Cpad), 1: t32(j.Cpad) * t32(Ci), 2: t32(Co) * t32(Ci)}[mode]
        elif self.kind == "unpack":                 # one 32 x 32 (x taps) destination tile per block (unpack_multi_kernel)
            self.blocks += ((Co + 31) // 32) * ((cn + 31) // 32)
        else:
            self.blocks += (total + 1023) // 1024
        self.jobs.append(j)
        self.keep.append((src, dst))
        self.table = None

    def signature(self):
        return tuple((j.src, j.dst) for j in self.jobs)


class CudaOps:
    """The product backend: every method launches kernels of libb3d.so on the current stream."""

    act_dtype = torch.bfloat16
    name = "cuda"

    def __init__(self, device):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.B3DError("b3d kernels need a CUDA device (sm_100a); there is no CPU path")

    # ---- plumbing
    def stream(self):
        return c_void(torch.cuda.current_stream(self.device).cuda_stream)

    def empty(self, shape, dtype=None):
        return torch.empty(shape, dtype=dtype or self.act_dtype, device=self.device)

    def zeros(self, shape, dtype=None):
        return torch.zeros(shape, dtype=dtype or self.act_dtype, device=self.device)

    # ---- second stream for launches whose results nothing downstream reads until the end of a phase (weight / bias
    # gradients): they overlap the data-gradient chain, filling SMs that small or tail-end launches leave idle.
    # Works eagerly and under stream capture (the fork/join become graph edges).
    use_side = True
    n_side = 1            # side streams used round-robin, one `with side():` region each (regions are independent of each other)
    _sides, _side_dirty, _side_rr = None, None, 0

    @contextlib.contextmanager
    def side(self):
        """launches inside run on a side stream, after everything issued so far on the current stream"""
        if self.prof is not None or not self.use_side:
            yield
            return
        if self._sides is None or len(self._sides) < self.n_side:
            self._sides = (self._sides or []) + [torch.cuda.Stream(self.device) for _ in range(self.n_side - len(self._sides or []))]
            self._side_dirty = (self._side_dirty or []) + [False] * (self.n_side - len(self._side_dirty or []))
        i = self._side_rr % self.n_side
        self._side_rr += 1
        st = self._sides[i]
        st.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(st):
            yield
        self._side_dirty[i] = True

    def join_side(self):
        """the current stream waits for everything issued on the side streams"""
        if self._side_dirty:
            for i, dirty in enumerate(self._side_dirty):
                if dirty:
                    torch.cuda.current_stream(self.device).wait_stream(self._sides[i])
                    self._side_dirty[i] = False

    prof = None          # when a list: (name, start_event, end_event, flop, info) per launch (bench.py's profiled step)
    _flop, _info = 0.0, ""

    def _call(self, name, *args):
        if self.prof is None:
            _lib.check(getattr(self.lib, name)(*args), name)
            return
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        _lib.check(getattr(self.lib, name)(*args), name)
        e.record()
        self.prof.append((name, s, e, self._flop, self._info))
        self._flop, self._info = 0.0, ""

    @staticmethod
    def gemm_flop(d):
        """algorithmic FLOPs of one b3d_gemm launch (true extents, no padding)"""
        if d.a_major == 0 and d.b_major == 0:
            k = sum(t.shape[3] * tp for t, tp in zip(d.a, d.a_taps))
            return 2.0 * d.N * d.H * d.W * d.n_cols * k
        if d.a_major == 0:
            return 2.0 * d.N * d.H * d.W * d.n_cols * d.a[0].shape[3]
        if d.batched:
            return 2.0 * d.m_rows * d.n_cols * d.W * d.H * d.N
        return 2.0 * d.m_rows * d.n_cols * d.N * d.H * d.W * d.b_taps

    # ---- tensor-core path
    def gemm_stats_supported(self, d: GemmDesc):
        """can b3d_gemm take the GroupNorm statistics of d.out in its epilogue (d.gn_cpg / d.gn_rows_per_image set)?"""
        return bool(self.lib.b3d_gemm_stats_supported(ctypes.byref(self._cdesc(d))))

    def _cdesc(self, d: GemmDesc):
        cd = _GemmDesc()
        assert 1 <= len(d.a) <= 3
        for i, t in enumerate(d.a):
            assert t.dtype == torch.bfloat16
            cd.a[i] = _view4(t)
            cd.a_taps[i] = d.a_taps[i] if i < len(d.a_taps) else 1
        cd.n_src = len(d.a)
        assert d.b.dtype == torch.bfloat16
        cd.b = _view4(d.b)
        cd.a_major, cd.b_major = d.a_major, d.b_major
        cd.conv_stride, cd.pad, cd.b_taps = d.conv_stride, d.pad, d.b_taps
        cd.W, cd.H, cd.N = d.W, d.H, d.N
        cd.m_rows, cd.n_cols = d.m_rows, d.n_cols
        cd.batched, cd.b_batch_xor, cd.out_batch_xor, cd.split_k = int(d.batched), d.b_batch_xor, d.out_batch_xor, d.split_k
        o = d.out
        assert o.dim() == 4 and o.stride(3) == 1 and o.dtype in (torch.bfloat16, torch.float32)
        cd.out = o.data_ptr()
        cd.out_fp32 = int(o.dtype == torch.float32)
        cd.out_stride[0], cd.out_stride[1], cd.out_stride[2] = o.stride(2), o.stride(1), o.stride(0)
        cd.out_tap_stride = d.out_tap_stride
        if d.out2 is not None:
            assert d.out2.stride() == o.stride() and d.out2.dtype == torch.bfloat16
            cd.out2 = d.out2.data_ptr()
        if d.bias is not None:
            assert d.bias.dtype == torch.float32 and d.bias.numel() >= d.n_cols
            cd.bias = d.bias.data_ptr()
        if d.rowvec is not None:
            assert d.rowvec.dtype == torch.float32 and d.rowvec.dim() == 2 and d.rowvec.stride(1) == 1
            cd.rowvec = d.rowvec.data_ptr()
            cd.rowvec_ld = d.rowvec.stride(0)
        if d.resid is not None:
            assert d.resid.stride() == o.stride() and d.resid.dtype == torch.bfloat16
            cd.resid = d.resid.data_ptr()
        cd.alpha, cd.out_scale, cd.act = d.alpha, d.out_scale, d.act
        cd.out_atomic, cd.bn_tile = int(d.out_atomic), d.bn_tile
        cd.gn_cpg, cd.gn_rows_per_image = d.gn_cpg, d.gn_rows_per_image
        if d.gn_sums is not None:
            assert d.gn_sums.dtype == torch.float32 and d.gn_sums.is_contiguous()
            cd.gn_sums = d.gn_sums.data_ptr()
        return cd

    def gemm(self, d: GemmDesc):
        cd = self._cdesc(d)
        o = d.out
        if self.prof is not None:
            self._flop = self.gemm_flop(d)
            self._info = (f"maj{d.a_major}{d.b_major} WHN={d.W}x{d.H}x{d.N} rows={d.m_rows} cols={d.n_cols} "
                          f"C={[t.shape[3] for t in d.a]} taps={d.a_taps if d.a_major == 0 else d.b_taps} s={d.conv_stride} "
                          f"split={d.split_k} bn={d.bn_tile} batched={int(d.batched)} fp32={int(o.dtype == torch.float32)}")
        self._call("b3d_gemm", ctypes.byref(cd), self.stream())

    def attn_fwd(self, qkv, heads, cross, scale, out, probs=None, lse=None):
        """fused softmax(QK^T*scale)V; qkv [N,L,3C] bf16 -> out [N,L,C]; probs (optional) [N,heads,L,Lp] view with Lp % 8 == 0;
        lse (optional) fp32 [N,heads,L]: the row statistics a P-recomputing backward needs"""
        N, L, C3 = qkv.shape
        d = C3 // 3 // heads
        assert qkv.is_contiguous() and out.is_contiguous() and qkv.dtype == torch.bfloat16
        assert lse is None or (lse.dtype == torch.float32 and lse.is_contiguous() and lse.shape == (N, heads, L))
        ld = probs.stride(2) if probs is not None else 0
        self._call("b3d_attn_fwd", _ptr(qkv), c_i32(N), c_i32(L), c_i32(heads), c_i32(d), c_i32(int(cross)), c_f32(scale), _ptr(out),
                   _ptr(probs), c_i64(ld), _ptr(lse), self.stream())

    def attn_bwd_supported(self, L, d):
        return bool(self.lib.b3d_attn_bwd_supported(c_i32(L), c_i32(d)))

    def attn_bwd(self, qkv, probs, o, dout, dqkv, heads, cross, scale, lse=None):
        """fused attention backward: qkv [N,L,3C], o / dout [N,L,C] -> dqkv [N,L,3C]; from probs [N,heads,L,Lp] view (saved by
        attn_fwd) or, probs=None, from lse [N,heads,L] (P is recomputed)"""
        N, L, C3 = qkv.shape
        d = C3 // 3 // heads
        assert qkv.is_contiguous() and o.is_contiguous() and dout.is_contiguous() and dqkv.is_contiguous()
        assert (probs is None) != (lse is None)
        self._call("b3d_attn_bwd", _ptr(qkv), _ptr(probs), c_i64(probs.stride(2) if probs is not None else 0), _ptr(lse), _ptr(o), _ptr(dout),
                   _ptr(dqkv), c_i32(N), c_i32(L), c_i32(heads), c_i32(d), c_i32(int(cross)), c_f32(scale), self.stream())

    # ---- conditioning
    def pose_embedding(self, K, R, t, mask, pos_emb, first_emb, other_emb, out):
        B = K.shape[0]
        N, H, W, ld = out.shape
        assert N == 2 * B and out.is_contiguous() and K.dtype == torch.float64
        self._call("b3d_pose_embedding", _ptr(K.contiguous()), _ptr(R.contiguous()), _ptr(t.contiguous()), _ptr(_u8(mask)),
                   _ptr(pos_emb), _ptr(first_emb), _ptr(other_emb), _ptr(out), c_i32(int(out.dtype == torch.float32)),
                   c_i64(ld), c_i32(B), c_i32(H), c_i32(W), self.stream())

    def pose_embedding_bwd(self, dP, dpos, dfirst, dother):
        _, H, W, ld = dP.shape
        assert dP.dtype == torch.float32 and dP.is_contiguous()
        self._call("b3d_pose_embedding_bwd", _ptr(dP), c_i64(ld), c_i32(H), c_i32(W), _ptr(dpos), _ptr(dfirst), _ptr(dother),
                   self.stream())

    def ddpm_embedding(self, logsnr, out):
        n, e = out.shape
        assert logsnr.dtype == torch.float32 and logsnr.numel() == n and out.is_contiguous()
        self._call("b3d_ddpm_embedding", _ptr(logsnr), _ptr(out), c_i32(n), c_i32(e), self.stream())

    def pack_input(self, x, z, out):
        B, _, H, W = x.shape
        self._call("b3d_pack_input", _ptr(x.contiguous()), _ptr(z.contiguous()), _ptr(out), c_i32(B), c_i32(H * W), self.stream())

    def unpack_input_grad(self, g, dx, dz):
        B, _, H, W = dx.shape
        self._call("b3d_unpack_input_grad", _ptr(g), _ptr(dx), _ptr(dz), c_i32(B), c_i32(H * W), self.stream())

    def unpack_output(self, y, eps):
        B, _, H, W = eps.shape
        self._call("b3d_unpack_output", _ptr(y), c_i64(y.shape[-1]), _ptr(eps), c_i32(B), c_i32(H * W), self.stream())

    def pack_output_grad(self, deps, dy):
        B, _, H, W = deps.shape
        self._call("b3d_pack_output_grad", _ptr(deps.contiguous()), _ptr(dy), c_i64(dy.shape[-1]), c_i32(B), c_i32(H * W), self.stream())

    # ---- parameter layouts
    def pack_weight(self, w, mode, out, ci0=0, cn=None, Cpad=None, k_off=0):
        Co, Ci = w.shape[0], w.shape[1]
        taps = w.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        assert w.dtype == torch.float32 and w.is_contiguous() and out.dim() == 2 and out.stride(1) == 1
        self._call("b3d_pack_weight", _ptr(w), c_i32(Co), c_i32(Ci), c_i32(taps), c_i32(mode), c_i32(ci0), c_i32(cn), c_i32(Cpad or 0),
                   c_i64(k_off), _ptr(out), c_i64(out.stride(0)), self.stream())

    def unpack_wgrad(self, g, k_off, Cpad, dw, ci0=0, cn=None, accumulate=False):
        Co, Ci = dw.shape[0], dw.shape[1]
        taps = dw.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        assert g.dtype == torch.float32 and dw.is_contiguous()
        self._call("b3d_unpack_wgrad", _ptr(g), c_i64(g.stride(0)), c_i64(k_off), c_i32(Cpad), c_i32(Co), c_i32(Ci), c_i32(taps), c_i32(ci0),
                   c_i32(cn), _ptr(dw), c_i32(int(accumulate)), self.stream())

    # recorded variants: same arguments, but the work is appended to a JobList and runs later in one launch
    def record_pack_weight(self, jl, w, mode, out, ci0=0, cn=None, Cpad=None, k_off=0):
        Co, Ci = w.shape[0], w.shape[1]
        taps = w.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        assert w.dtype == torch.float32 and w.is_contiguous() and out.dim() == 2 and out.stride(1) == 1 and taps <= 9
        total = Co * taps * Cpad if mode == 0 else (Ci * taps * Cpad if mode == 1 else Co * Ci * taps)
        if jl.kind == "pack" and id(out) not in jl.zeroed:      # pad columns / rows are never written by the grouped kernel
            base = out._base if out._base is not None else out
            if id(base) not in jl.zeroed:
                base.zero_()
                jl.zeroed.add(id(base))
            jl.zeroed.add(id(out))
        jl.add(w, out, Co, Ci, taps, mode, ci0, cn, Cpad, k_off, out.stride(0), total)

    def record_copy(self, jl, src, dst):
        assert src.dtype == torch.float32 and dst.dtype == torch.float32 and src.is_contiguous() and dst.is_contiguous()
        jl.add(src, dst, 0, 0, 1, 3, 0, 0, 0, 0, 0, src.numel())

    def record_unpack_wgrad(self, jl, g, k_off, Cpad, dw, ci0=0, cn=None, accumulate=False):
        Co, Ci = dw.shape[0], dw.shape[1]
        taps = dw.numel() // (Co * Ci)
        cn = Ci if cn is None else cn
        assert g.dtype == torch.float32 and dw.is_contiguous()
        jl.add(g, dw, Co, Ci, taps, 0, ci0, cn, Cpad, k_off, g.stride(0), Co * cn * taps, accumulate)

    def run_jobs(self, jl):
        if not jl.jobs:
            return
        if jl.table is None:
            if jl.kind == "pack":
                arr, block_of = jl.groups()
            else:
                arr, block_of = (_PackJob * len(jl.jobs))(*jl.jobs), []
                for i, j in enumerate(jl.jobs):
                    nxt = jl.jobs[i + 1].block0 if i + 1 < len(jl.jobs) else jl.blocks
                    block_of += [i] * (nxt - j.block0)
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            jl.table, jl.n_entries = host.to(self.device), len(arr)
            jl.block_table = torch.tensor(block_of, dtype=torch.int32).to(self.device)
        name = "b3d_pack_weights_multi" if jl.kind == "pack" else "b3d_unpack_wgrads_multi"
        self._call(name, _ptr(jl.table), c_i32(jl.n_entries), _ptr(jl.block_table), c_i64(jl.block_table.numel()), self.stream())

    def scale_colsum(self, a, out, scale, colsum):
        C = a.shape[-1]
        assert a.is_contiguous() and out.is_contiguous() and colsum.dtype == torch.float32
        self._call("b3d_scale_colsum", _ptr(a), _ptr(out), c_f32(scale), c_i64(a.numel() // C), c_i32(C), _ptr(colsum), self.stream())

    def matvec(self, W, v, out, transpose=False, add=None):
        """out = W v (+ add), or W^T v (+ add): fp32, W [R, C] contiguous"""
        R, C = W.shape
        assert W.dtype == torch.float32 and W.is_contiguous() and v.dtype == torch.float32 and out.dtype == torch.float32
        assert v.numel() == (R if transpose else C) and out.numel() == (C if transpose else R) and v.is_contiguous() and out.is_contiguous()
        self._call("b3d_matvec", _ptr(W), c_i32(R), c_i32(C), _ptr(v), _ptr(add) if add is not None else None, _ptr(out), c_i32(int(transpose)),
                   self.stream())

    def rank1_add(self, A, u, v):
        """A += outer(u, v): fp32, A [R, C] contiguous"""
        R, C = A.shape
        assert A.dtype == torch.float32 and A.is_contiguous() and u.numel() == R and v.numel() == C and u.dtype == v.dtype == torch.float32
        self._call("b3d_rank1_add", _ptr(A), c_i32(R), c_i32(C), _ptr(u), _ptr(v), self.stream())

    def cast(self, src, out):
        assert src.dtype == torch.float32 and src.is_contiguous() and out.is_contiguous()
        self._call("b3d_cast_bf16", _ptr(src), _ptr(out), c_i64(src.numel()), self.stream())

    # ---- resampling / elementwise / reductions
    def avgpool2(self, x, out, scale=0.25):
        N, H, W, C = x.shape
        self._call("b3d_avgpool2", _ptr(x), _ptr(out), c_i32(N), c_i32(H), c_i32(W), c_i32(C), c_f32(scale), self.stream())

    def upsample2(self, x, out, scale=1.0):
        N, Ho, Wo, C = out.shape
        self._call("b3d_upsample2", _ptr(x), _ptr(out), c_i32(N), c_i32(Ho), c_i32(Wo), c_i32(C), c_f32(scale), self.stream())

    def silu_bwd(self, dy, x, dx):
        self._call("b3d_silu_bwd", _ptr(dy), _ptr(x), _ptr(dx), c_i64(dy.numel()), self.stream())

    def silu_bwd_colsum(self, dy, x, dx, rows_per_group, group_out, total_out):
        """dx = dy*silu'(x) on [rows, C] views; group_out [groups, C] += per-group column sums, total_out [C] += all rows.
        x = dx = None: the two sums of dy alone"""
        C = dy.shape[-1]
        rows = dy.numel() // C
        assert dy.is_contiguous() and (x is None) == (dx is None) and (x is None or (x.is_contiguous() and dx.is_contiguous()))
        self._call("b3d_silu_bwd_colsum", _ptr(dy), _ptr(x), _ptr(dx), c_i32(C), c_i64(rows_per_group), c_i32(rows // rows_per_group),
                   _ptr(group_out), c_i64(group_out.stride(0) if group_out is not None else 0), _ptr(total_out), self.stream())

    def add_scale(self, a, b, out, scale):
        self._call("b3d_add_scale", _ptr(a), _ptr(b), _ptr(out), c_f32(scale), c_i64(a.numel()), self.stream())

    def colsum(self, x2d, rows_per_group, out, scale=1.0):
        rows, C = x2d.shape
        groups = rows // rows_per_group
        assert x2d.stride(1) == 1 and out.dtype == torch.float32 and out.stride(-1) == 1
        out_ld = out.stride(0) if out.dim() == 2 else 0
        self._call("b3d_colsum", _ptr(x2d), c_i64(x2d.stride(0)), c_i32(C), c_i64(rows_per_group), c_i32(groups), _ptr(out), c_i64(out_ld),
                   c_f32(scale), self.stream())

    def batch_frame_sum(self, x, out):
        B = x.shape[0] // 2
        self._call("b3d_batch_frame_sum", _ptr(x), _ptr(out), c_i32(B), c_i64(out.numel() // 2), self.stream())

    def conv_tap_scatter(self, G, C, h, w, s, dP, accumulate):
        _, H, W, ldp = dP.shape
        self._call("b3d_conv_tap_scatter", _ptr(G), c_i64(G.stride(0)), c_i32(C), c_i32(h), c_i32(w), c_i32(s), c_i32(H), c_i32(W), _ptr(dP),
                   c_i64(ldp), c_i32(int(accumulate)), self.stream())

    @staticmethod
    def _rows_ld(t):
        """[..., rows, L] view whose leading dims are packed over a padded row stride"""
        L, ld = t.shape[-1], t.stride(-2)
        assert t.stride(-1) == 1 and all(t.stride(i) == t.stride(i + 1) * t.shape[i + 1] for i in range(t.dim() - 2))
        return t.numel() // L, L, ld

    def softmax(self, S, P):
        rows, L, ld_s = self._rows_ld(S)
        rows_p, _, ld_p = self._rows_ld(P)
        assert rows == rows_p and P.dtype == torch.bfloat16
        self._call("b3d_softmax", _ptr(S), c_i32(int(S.dtype == torch.float32)), c_i64(ld_s), _ptr(P), c_i64(ld_p), c_i64(rows), c_i32(L),
                   self.stream())

    def softmax_bwd(self, P, dP, dS):
        rows, L, ld_p = self._rows_ld(P)
        _, _, ld_dp = self._rows_ld(dP)
        _, _, ld_ds = self._rows_ld(dS)
        self._call("b3d_softmax_bwd", _ptr(P), c_i64(ld_p), _ptr(dP), c_i32(int(dP.dtype == torch.float32)), c_i64(ld_dp), _ptr(dS),
                   c_i64(ld_ds), c_i64(rows), c_i32(L), self.stream())

    # ---- GroupNorm family.  x0/x1: [N, HW, C*] views with stride(-1) == 1 (x1 may be None)
    @staticmethod
    def _src(x):
        if x is None:
            return c_void(0), c_i32(0), c_i64(0)
        assert x.dim() == 3 and x.stride(2) == 1 and x.stride(0) == x.shape[1] * x.stride(1)
        return _ptr(x), c_i32(x.shape[2]), c_i64(x.stride(1))

    def gn_stats(self, x0, x1, sums):
        N, HW, _ = x0.shape
        self._call("b3d_gn_stats", *self._src(x0), *self._src(x1), c_i32(N), c_i32(HW), _ptr(sums), self.stream())

    def gn_apply(self, x0, x1, sums, gamma, beta, mode, out, ss=None, F=0, drop_p=0.0, seed=0, seed_ptr=None, ss_image=None):
        """seed_ptr: optional int64 device scalar added to `seed` (a step counter that lives on the device, so a
        captured CUDA graph draws a fresh dropout mask on every replay)"""
        N, HW, _ = x0.shape
        self._call("b3d_gn_apply", *self._src(x0), *self._src(x1), c_i32(N), c_i32(HW), _ptr(sums), _ptr(gamma), _ptr(beta), c_i32(mode),
                   _ptr(ss), c_i64(ss.stride(0) if ss is not None else 0), c_i32(F), c_f32(drop_p), ctypes.c_uint64(seed), _ptr(seed_ptr),
                   _ptr(out), c_i64(out.stride(1)), _ptr(ss_image), self.stream())

    def gn_stats_apply(self, x0, x1, sums, gamma, beta, mode, out, ss=None, F=0, drop_p=0.0, seed=0, seed_ptr=None, ss_image=None):
        """gn_stats + gn_apply of the same input as one call (image-chunked for L2 reuse when a chunk budget is set)"""
        N, HW, _ = x0.shape
        self._call("b3d_gn_stats_apply", *self._src(x0), *self._src(x1), c_i32(N), c_i32(HW), _ptr(sums), _ptr(gamma), _ptr(beta), c_i32(mode),
                   _ptr(ss), c_i64(ss.stride(0) if ss is not None else 0), c_i32(F), c_f32(drop_p), ctypes.c_uint64(seed), _ptr(seed_ptr),
                   _ptr(out), c_i64(out.stride(1)), _ptr(ss_image), self.stream())

    def gn_bwd(self, x0, x1, sums, gamma, beta, mode, dy, dgamma, dbeta, ws, dx0, dx1=None, ss=None, F=0, dss=None, drop_p=0.0, seed=0,
               add0=None, add1a=None, add1b=None, seed_ptr=None, dss_sum=None, dx_sum=None, add0_scale=1.0):
        N, HW, _ = x0.shape
        ld = lambda t, i=1: c_i64(t.stride(i) if t is not None else 0)
        self._call("b3d_gn_bwd", *self._src(x0), *self._src(x1), c_i32(N), c_i32(HW), _ptr(sums), _ptr(gamma), _ptr(beta), c_i32(mode),
                   _ptr(dy), ld(dy), _ptr(ss), ld(ss, 0), c_i32(F), _ptr(dss), ld(dss, 0), c_f32(drop_p), ctypes.c_uint64(seed),
                   _ptr(seed_ptr), _ptr(dgamma), _ptr(dbeta), _ptr(dss_sum), _ptr(dx_sum), _ptr(ws), _ptr(dx0), ld(dx0), _ptr(dx1), ld(dx1),
                   _ptr(add0), ld(add0),
                   _ptr(add1a), ld(add1a), _ptr(add1b), ld(add1b), c_f32(add0_scale), self.stream())

    # ---- diffusion glue
    def q_sample(self, z, noise, logsnr, z_out, x=None, xrand=None, mask=None, x_out=None):
        B = z.shape[0]
        per = z.numel() // B
        assert all(t is None or t.is_contiguous() for t in (z, noise, z_out, x, xrand, x_out))
        self._call("b3d_q_sample", _ptr(z), _ptr(noise), _ptr(logsnr), _ptr(x), _ptr(xrand), _ptr(_u8(mask) if mask is not None else None),
                   _ptr(z_out), _ptr(x_out), c_i32(B), c_i64(per), self.stream())

    def loss_scratch(self):
        return torch.zeros(int(self.lib.b3d_loss_scratch_bytes()) // 4, dtype=torch.float32, device=self.device)

    def loss(self, noise, pred, dpred, loss_type, scratch, loss_out, grad_scale=1.0):
        self._call("b3d_loss", _ptr(noise), _ptr(pred), _ptr(dpred), c_i64(noise.numel()), c_i32(loss_type), c_f32(grad_scale), _ptr(scratch),
                   _ptr(loss_out), self.stream())

    def sampler_update(self, z, eps_c, eps_u, w, noise, out, logsnr, logsnr_next):
        B = z.shape[0]
        assert w.dtype == torch.float32 and w.is_contiguous() and w.numel() == B, "w: one fp32 guidance weight per image"
        self._call("b3d_sampler_update", _ptr(z), _ptr(eps_c), _ptr(eps_u), _ptr(w), _ptr(noise), _ptr(out), c_f32(float(logsnr)),
                   c_f32(float(logsnr_next)), c_i32(B), c_i64(z.numel() // B), self.stream())

    def sampler_update_dev(self, z, eps_c, eps_u, w, noise, out, sched):
        B = z.shape[0]
        assert sched.dtype == torch.float32 and sched.numel() >= 2
        assert w.dtype == torch.float32 and w.is_contiguous() and w.numel() == B, "w: one fp32 guidance weight per image"
        self._call("b3d_sampler_update_dev", _ptr(z), _ptr(eps_c), _ptr(eps_u), _ptr(w), _ptr(noise), _ptr(out), _ptr(sched), c_i32(B),
                   c_i64(z.numel() // B), self.stream())

    def cond_silu(self, c, lemb, lemb_row, out):
        """out[j] = silu(c[j] + lemb[lemb_row[j]]): c fp32 [D, h, w, E] (cached conditioning conv + bias), lemb fp32 [N, E], out bf16"""
        D, E = c.shape[0], c.shape[-1]
        assert c.dtype == torch.float32 and c.is_contiguous() and out.is_contiguous() and lemb.is_contiguous() and lemb_row.dtype == torch.int32
        self._call("b3d_cond_silu", _ptr(c), _ptr(lemb), _ptr(lemb_row), _ptr(out), c_i32(D), c_i64(c.numel() // (D * E)), c_i32(E), self.stream())

    def adam(self, p, g, m, v, lr, beta1, beta2, eps, step, grad_scale=1.0, ema=None, ema_decay=0.0):
        """ema (optional, fp32 like p): ema = ema_decay * ema + (1 - ema_decay) * p_new, in the same pass; g: fp32 or bf16"""
        if g.dtype == torch.bfloat16:
            assert g.numel() == p.numel() and g.is_contiguous()
            self._call("b3d_adam_bf16grad", _ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(ema) if ema is not None else None, c_i64(p.numel()),
                       c_f32(lr), c_f32(beta1), c_f32(beta2), c_f32(eps), c_i32(step), c_f32(grad_scale), c_f32(ema_decay), self.stream())
        elif ema is None:
            self._call("b3d_adam", _ptr(p), _ptr(g), _ptr(m), _ptr(v), c_i64(p.numel()), c_f32(lr), c_f32(beta1), c_f32(beta2), c_f32(eps),
                       c_i32(step), c_f32(grad_scale), self.stream())
        else:
            assert ema.dtype == torch.float32 and ema.numel() == p.numel()
            self._call("b3d_adam_ema", _ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(ema), c_i64(p.numel()), c_f32(lr), c_f32(beta1), c_f32(beta2),
                       c_f32(eps), c_i32(step), c_f32(grad_scale), c_f32(ema_decay), self.stream())
