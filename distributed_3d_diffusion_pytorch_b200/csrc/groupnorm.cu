// GroupNorm(32 groups, eps 1e-5, per image) on bf16 NHWC tensors, with the consumers the model needs
// fused into the apply pass: SiLU (ResnetBlock GN0, lastgn), FiLM (+dropout) (ResnetBlock GN1), plain
// (AttnBlock).  Inputs may be the implicit channel-concat of two tensors (up-path skip connections).
// Replaces nn.GroupNorm / F.silu / FiLM apply / Dropout: xunet.py:61-71,87,140,143-145,198,535.
// HBM-bound: every pass is 16-byte vectorised over channels and reads each tensor once.
#include "../../include/b3d.h"
#include <cooperative_groups.h>

#include "b3d_common.cuh"

void b3d_count_launch(int n);

namespace {

constexpr int G = 32;
constexpr int kT = 256;

struct Src2 {
  const bf16* p0; const bf16* p1;
  int C0, C1; long long ld0, ld1;
};

// Thread layout shared by all kernels: a block covers `rows_per_block` pixels of ONE image and all C channels;
// thread t owns the 8-channel vector v = t % (C/8) for every row it visits (so gamma/beta/mean/rstd of its
// channels live in registers) and visits rows rsub, rsub + rows_par, ... with U independent 16-byte loads per
// tensor in flight (HBM latency is hidden by memory-level parallelism, not by occupancy).
// Per-channel affine of the normalisation kept in registers: z = gamma*xhat + beta = x*a + b.
struct Aff {
  float a[8], b[8];
};

__device__ __forceinline__ void block_stats(const float* stats, int n, int cpg, int HW, float* s_mean, float* s_rstd) {
  if (threadIdx.x < G) {
    const float cnt = (float)cpg * (float)HW;
    const float sum = stats[((long long)n * G + threadIdx.x) * 2], sq = stats[((long long)n * G + threadIdx.x) * 2 + 1];
    const float mean = sum / cnt;
    const float var = fmaxf(sq / cnt - mean * mean, 0.f);
    s_mean[threadIdx.x] = mean; s_rstd[threadIdx.x] = rsqrtf(var + 1e-5f);
  }
}
__device__ __forceinline__ void load_aff(Aff& ch, const float* gamma, const float* beta, int c0, int cpg, const float* s_mean,
                                         const float* s_rstd) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    ch.a[j] = s_rstd[g] * gamma[c0 + j];
    ch.b[j] = beta[c0 + j] - s_mean[g] * ch.a[j];
  }
}
struct RowPtr {   // per-thread base pointer of its channel vector inside the (possibly two-source) input
  const bf16* p; long long ld;
};
__device__ __forceinline__ RowPtr src_ptr(const Src2& s, int c0) {
  if (c0 < s.C0) return RowPtr{s.p0 + c0, s.ld0};
  return RowPtr{s.p1 + (c0 - s.C0), s.ld1};
}
__device__ __forceinline__ bf16x8 ld8(const bf16* p) { return *reinterpret_cast<const bf16x8*>(p); }
__device__ __forceinline__ bf16x8 zero8() { bf16x8 z; z.u = make_uint4(0u, 0u, 0u, 0u); return z; }

// Block-level sum of per-thread 8-channel partials over the rows_par threads that own the same channel vector.
// Plain stores + a strided read-back: shared-memory fp32 atomicAdd compiles to a CAS spin loop (ATOMS.CAST.SPIN) and,
// with rows_par-way contention per address, cost as much as the whole streaming loop.
// red: smem [256*8]; result: smem [C].  All threads of the block must call it.
__device__ __forceinline__ void block_sum8(const float* vals, bool active, int v, int rsub, int rows_par, int C, float* red, float* result) {
  if (active) {
#pragma unroll
    for (int j = 0; j < 8; ++j) red[rsub * C + v * 8 + j] = vals[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rows_par; ++r) s += red[r * C + c];
    result[c] = s;
  }
  __syncthreads();
}

// ---- stats: raw (sum, sumsq) per (image, group) accumulated with fp32 atomics into ws[N,32,2] (pre-zeroed)
// (s_part != nullptr: the block's partial (sum, sumsq) of group g go to s_part[2g], s_part[2g+1] in shared memory instead -- the
// cluster-fused kernels below combine them across the blocks of an image)
__device__ __forceinline__ void stats_body(const Src2& s, int HW, int rows_per_block, int n, int bx, float* sm, float* __restrict__ ws,
                                           float* s_part) {
  // sm: red [256*8] | sum [C] | sumsq [C]
  constexpr int U = 4;
  const int C = s.C0 + s.C1, cpg = C / G, vecs = C / 8;
  float* red = sm; float* r_a = sm + 2048; float* r_b = r_a + C;
  const int r0 = bx * rows_per_block, r1 = min(HW, r0 + rows_per_block);
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs;
  const bool active = rsub < rows_par;
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = 0.f; b[i] = 0.f; }
  if (active) {
    const RowPtr rp = src_ptr(s, v * 8);
    const bf16* base = rp.p + (long long)n * HW * rp.ld;
    for (int r = r0 + rsub; r < r1; r += rows_par * U) {
      bf16x8 x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { const int rr = r + u * rows_par; x[u] = rr < r1 ? ld8(base + (long long)rr * rp.ld) : zero8(); }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float f[8]; b3d_unpack8(x[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) { a[i] += f[i]; b[i] = fmaf(f[i], f[i], b[i]); }
      }
    }
  }
  block_sum8(a, active, v, rsub, rows_par, C, red, r_a);
  block_sum8(b, active, v, rsub, rows_par, C, red, r_b);
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    float sa = 0.f, sb = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) { sa += r_a[c]; sb += r_b[c]; }
    if (s_part != nullptr) { s_part[2 * g] = sa; s_part[2 * g + 1] = sb; }
    else {
      atomicAdd(&ws[((long long)n * G + g) * 2], sa);
      atomicAdd(&ws[((long long)n * G + g) * 2 + 1], sb);
    }
  }
}
__global__ void __launch_bounds__(kT) gn_stats_kernel(Src2 s, int HW, int rows_per_block, float* __restrict__ ws, int n0) {
  b3d_pdl_sync();
  extern __shared__ float sm[];
  stats_body(s, HW, rows_per_block, blockIdx.y + n0, blockIdx.x, sm, ws, nullptr);
}

// ---- dropout keep-mask of one 8-channel vector: b3d_drop_pair (b3d_common.cuh) hashes (seed, pair index) to 32 bits,
// 16 per element; an element is dropped when its field is below thr16 = round(p * 65536).
struct Drop {
  unsigned long long seedmix;   // (seed + *seed_ptr * B3D_SEED_STEP_STRIDE) * golden ratio
  unsigned thr_hi;              // thr16 << 16
  float keep;                   // 1 / (1 - p)
};
__device__ __forceinline__ Drop make_drop(float p, unsigned long long seed, const unsigned long long* seed_ptr) {
  Drop d;
  // the step counter enters with a stride far above any host seed (those number the blocks of a step and the ranks), so
  // no (block, step) pair shares its mask with another one
  d.seedmix = (seed + (seed_ptr ? *seed_ptr * B3D_SEED_STEP_STRIDE : 0ULL)) * 0x9E3779B97F4A7C15ULL;
  d.thr_hi = b3d_drop_thr16(p) << 16;
  d.keep = 1.f / (1.f - p);
  return d;
}
__device__ __forceinline__ void drop_mask8(const Drop& d, unsigned long long elem0, float* m) {
  const unsigned long long base = d.seedmix + (elem0 >> 1);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const unsigned h = b3d_drop_pair(base + k);
    m[2 * k] = ((h << 16) < d.thr_hi) ? 0.f : d.keep;
    m[2 * k + 1] = (h < d.thr_hi) ? 0.f : d.keep;
  }
}

struct ApplyP {
  Src2 s;
  const float* stats;  // raw sums [N,32,2]
  const float* gamma; const float* beta;
  bf16* out; long long ld_out;
  int HW, mode;  // 0 plain, 1 silu, 2 film
  int n0;        // first image of this launch (image-chunked launches, see gn_chunk_images)
  const bf16* ss; long long ld_ss; int F;  // film: scale = ss[row, c], shift = ss[row, F + c]
  const int* ss_image;                     // optional [N]: image n reads the ss rows of image ss_image[n] (images that share their conditioning)
  float drop_p; unsigned long long seed; const unsigned long long* seed_ptr;  // effective seed = seed + *seed_ptr * B3D_SEED_STEP_STRIDE (device step counter)
};

constexpr int kApplyU0 = 4, kApplyU2 = 2;   // rows in flight per thread (mode 0/1: one 16 B load per row; film: three)
constexpr int kBwdU1 = 4, kBwdU2 = 2;       // reduce pass (mode 0/1: two loads per row; film: three)
constexpr int kBwdApplyU = 2;
#ifndef B3D_GN_BWD2_BLOCKS
#define B3D_GN_BWD2_BLOCKS 2
#endif
#ifndef B3D_GN_BWD1_BLOCKS
#define B3D_GN_BWD1_BLOCKS 3
#endif
constexpr int kBwdBlocks1 = B3D_GN_BWD1_BLOCKS;
constexpr int kBwdBlocks2 = B3D_GN_BWD2_BLOCKS;   // resident blocks per SM of the film-mode backward kernels

template <int MODE>
__device__ __forceinline__ void apply_body(const ApplyP& p, int rows_per_block, int n, int bx, const float* s_mean, const float* s_rstd) {
  constexpr int U = MODE == 2 ? kApplyU2 : kApplyU0;
  const int C = p.s.C0 + p.s.C1, cpg = C / G, vecs = C / 8;
  const int r0 = bx * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs, c0 = v * 8;
  if (rsub >= rows_par) return;
  Aff ch; load_aff(ch, p.gamma, p.beta, c0, cpg, s_mean, s_rstd);
  const RowPtr rp = src_ptr(p.s, c0);
  const long long img = (long long)n * p.HW;
  const long long ss_off = (MODE == 2 && p.ss_image != nullptr) ? ((long long)p.ss_image[n] - n) * p.HW : 0;   // row offset into ss
  const bool drop = MODE == 2 && p.drop_p > 0.f;
  Drop dr{};
  if (drop) dr = make_drop(p.drop_p, p.seed, p.seed_ptr);
  for (int r = r0 + rsub; r < r1; r += rows_par * U) {
    bf16x8 x[U], sc[U], sh[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = img + r + u * rows_par;
      const bool ok = r + u * rows_par < r1;
      x[u] = ok ? ld8(rp.p + row * rp.ld) : zero8();
      if (MODE == 2) {
        sc[u] = ok ? ld8(p.ss + (row + ss_off) * p.ld_ss + c0) : zero8();
        sh[u] = ok ? ld8(p.ss + (row + ss_off) * p.ld_ss + p.F + c0) : zero8();
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (r + u * rows_par >= r1) break;
      const long long row = img + r + u * rows_par;
      float f[8], a[8], b[8], m[8];
      b3d_unpack8(x[u], f);
      if (MODE == 2) {
        b3d_unpack8(sc[u], a); b3d_unpack8(sh[u], b);
        if (drop) drop_mask8(dr, (unsigned long long)row * C + c0, m);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float y = fmaf(f[j], ch.a[j], ch.b[j]);
        if (MODE == 1) y = b3d_silu(y);
        else if (MODE == 2) {
          y = fmaf(y, a[j], y) + b[j];
          if (drop) y *= m[j];
        }
        f[j] = y;
      }
      *reinterpret_cast<bf16x8*>(p.out + row * p.ld_out + c0) = b3d_pack8(f);
    }
  }
}
template <int MODE>
__global__ void __launch_bounds__(kT, MODE == 2 ? 3 : 4) gn_apply_kernel(ApplyP p, int rows_per_block) {
  b3d_pdl_sync();
  __shared__ float s_mean[G], s_rstd[G];
  const int C = p.s.C0 + p.s.C1, n = blockIdx.y + p.n0;
  block_stats(p.stats, n, C / G, p.HW, s_mean, s_rstd);
  __syncthreads();
  apply_body<MODE>(p, rows_per_block, n, blockIdx.x, s_mean, s_rstd);
}


struct BwdP {
  Src2 s;
  const float* stats; const float* gamma; const float* beta;
  const bf16* dy; long long ld_dy;
  int HW, mode;
  int n0;                           // first image of this launch (image-chunked launches, see gn_chunk_images)
  const bf16* ss; long long ld_ss; int F;
  bf16* dss; long long ld_dss;      // film: d(scale) at [row, c], d(shift) at [row, F + c]
  float drop_p; unsigned long long seed; const unsigned long long* seed_ptr;
  float* dgamma; float* dbeta;      // fp32 [C], accumulated
  float* dss_sum;                   // optional fp32 [2F]: column sums of dss (the FiLM dense bias gradient)
  float* dx_sum;                    // optional fp32 [C]: column sums of dx before the addends (the producer's bias gradient)
  float* ws;                        // [N,32,2] (S1, S2), pre-zeroed
  // pass 2 outputs
  bf16* dx0; bf16* dx1; long long ld_dx0, ld_dx1;
  const bf16* add0; long long ld_add0;   // optional, full C wide, added to dx times add0_scale
  float add0_scale;
  const bf16* add1a; const bf16* add1b; long long ld_add1a, ld_add1b;  // optional second addend, split like the sources
};

// dz = d(loss)/d(gamma*xhat+beta) for one 8-channel vector, from the loaded x, dy (and FiLM scale / dropout mask m)
template <int MODE>
__device__ __forceinline__ void bwd_dz(const Aff& ch, const float* f, const float* dy, const float* sc, const float* m, bool drop,
                                       float* dz, float* dsc, float* dsh) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float z = fmaf(f[j], ch.a[j], ch.b[j]);
    float d = dy[j];
    if (MODE == 1) d *= b3d_silu_grad(z);
    else if (MODE == 2) {
      if (drop) d *= m[j];
      dsc[j] = d * z; dsh[j] = d;
      d = fmaf(d, sc[j], d);
    }
    dz[j] = d;
  }
}

// pass 1: d_scale/d_shift (film), dgamma/dbeta, per-(image,group) S1 = sum dz*gamma, S2 = sum dz*gamma*xhat.
// The loop accumulates sum dz*x and sum dz per channel; xhat = (x - mean)*rstd is applied to the block's partial sums
// (one image per block, so mean/rstd are block constants per group).
// (s_part != nullptr: the block's partial S1, S2 of group g go to s_part[2g], s_part[2g+1] in shared memory instead of the
// global workspace -- see the cluster-fused kernels)
template <int MODE>
__device__ __forceinline__ void bwd_reduce_body(const BwdP& p, int rows_per_block, int n, int bx, float* sm, const float* s_mean,
                                                const float* s_rstd, float* s_part) {
  // sm: red [256*8] | dgamma [C] | dbeta [C] | dscale-sum [C] | dshift-sum [C]
  constexpr int U = MODE == 2 ? kBwdU2 : kBwdU1;
  const int C = p.s.C0 + p.s.C1, cpg = C / G, vecs = C / 8;
  float* red = sm; float* r_g = sm + 2048; float* r_b = r_g + C; float* r_sc = r_b + C; float* r_sh = r_sc + C;
  const bool want_ss = MODE == 2 && p.dss_sum != nullptr;
  const int r0 = bx * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs, c0 = v * 8;
  const bool active = rsub < rows_par;
  float ag[8], ab[8], asc[8], ash[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { ag[j] = 0.f; ab[j] = 0.f; asc[j] = 0.f; ash[j] = 0.f; }
  if (active) {
    Aff ch; load_aff(ch, p.gamma, p.beta, c0, cpg, s_mean, s_rstd);
    const RowPtr rp = src_ptr(p.s, c0);
    const long long img = (long long)n * p.HW;
    const bool drop = MODE == 2 && p.drop_p > 0.f;
    Drop dr{};
    if (drop) dr = make_drop(p.drop_p, p.seed, p.seed_ptr);
    for (int r = r0 + rsub; r < r1; r += rows_par * U) {
      bf16x8 x[U], dy[U], sc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = img + r + u * rows_par;
        const bool ok = r + u * rows_par < r1;
        x[u] = ok ? ld8(rp.p + row * rp.ld) : zero8();
        dy[u] = ok ? ld8(p.dy + row * p.ld_dy + c0) : zero8();
        if (MODE == 2) sc[u] = ok ? ld8(p.ss + row * p.ld_ss + c0) : zero8();
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (r + u * rows_par >= r1) break;
        const long long row = img + r + u * rows_par;
        float f[8], d[8], s8[8], m[8], dz[8], dsc[8], dsh[8];
        b3d_unpack8(x[u], f); b3d_unpack8(dy[u], d);
        if (MODE == 2) {
          b3d_unpack8(sc[u], s8);
          if (drop) drop_mask8(dr, (unsigned long long)row * C + c0, m);
        }
        bwd_dz<MODE>(ch, f, d, s8, m, drop, dz, dsc, dsh);
        if (MODE == 2) {
          *reinterpret_cast<bf16x8*>(p.dss + row * p.ld_dss + c0) = b3d_pack8(dsc);
          *reinterpret_cast<bf16x8*>(p.dss + row * p.ld_dss + p.F + c0) = b3d_pack8(dsh);
#pragma unroll
          for (int j = 0; j < 8; ++j) { asc[j] += dsc[j]; ash[j] += dsh[j]; }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { ag[j] = fmaf(dz[j], f[j], ag[j]); ab[j] += dz[j]; }
      }
    }
  }
  block_sum8(ag, active, v, rsub, rows_par, C, red, r_g);
  block_sum8(ab, active, v, rsub, rows_par, C, red, r_b);
  if (want_ss) {
    block_sum8(asc, active, v, rsub, rows_par, C, red, r_sc);
    block_sum8(ash, active, v, rsub, rows_par, C, red, r_sh);
  }
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const float dg = s_rstd[g] * (r_g[c] - s_mean[g] * r_b[c]);   // sum dz*xhat
    r_g[c] = dg;
    atomicAdd(&p.dgamma[c], dg);
    atomicAdd(&p.dbeta[c], r_b[c]);
    if (want_ss) { atomicAdd(&p.dss_sum[c], r_sc[c]); atomicAdd(&p.dss_sum[p.F + c], r_sh[c]); }
  }
  __syncthreads();
  if (threadIdx.x < G) {
    float s1 = 0.f, s2 = 0.f;
    for (int c = threadIdx.x * cpg; c < (threadIdx.x + 1) * cpg; ++c) { const float gam = p.gamma[c]; s1 += r_b[c] * gam; s2 += r_g[c] * gam; }
    if (s_part != nullptr) { s_part[2 * threadIdx.x] = s1; s_part[2 * threadIdx.x + 1] = s2; }
    else {
      atomicAdd(&p.ws[((long long)n * G + threadIdx.x) * 2], s1);
      atomicAdd(&p.ws[((long long)n * G + threadIdx.x) * 2 + 1], s2);
    }
  }
}
template <int MODE>
__global__ void __launch_bounds__(kT, MODE == 2 ? kBwdBlocks2 : kBwdBlocks1) gn_bwd_reduce_kernel(BwdP p, int rows_per_block) {
  b3d_pdl_sync();
  extern __shared__ float sm[];
  __shared__ float s_mean[G], s_rstd[G];
  const int C = p.s.C0 + p.s.C1, n = blockIdx.y + p.n0;
  block_stats(p.stats, n, C / G, p.HW, s_mean, s_rstd);
  __syncthreads();
  bwd_reduce_body<MODE>(p, rows_per_block, n, blockIdx.x, sm, s_mean, s_rstd, nullptr);
}

// pass 2: dx = rstd * (dz*gamma - (S1 + xhat*S2)/cnt) [+ add0] [+ add1]  =  dz*a + x*c1 + c0 per channel
// group constants of pass 2 from the group totals S1, S2 (thread g < 32 handles group g; it wrote s_mean / s_rstd[g] itself)
__device__ __forceinline__ void bwd_group_consts(float S1, float S2, int cpg, int HW, const float* s_mean, const float* s_rstd, float* s_c1,
                                                 float* s_c0) {
  const float cnt = (float)cpg * (float)HW;
  const float s1 = S1 / cnt, s2 = S2 / cnt;
  const float rstd = s_rstd[threadIdx.x], c1 = -rstd * rstd * s2;
  s_c1[threadIdx.x] = c1;
  s_c0[threadIdx.x] = -rstd * s1 - c1 * s_mean[threadIdx.x];
}
template <int MODE>
__device__ __forceinline__ void bwd_apply_body(const BwdP& p, int rows_per_block, int n, int bx, float* sm, const float* s_mean,
                                               const float* s_rstd, const float* s_c1, const float* s_c0) {
  // sm: red [256*8] | column sums of dx [C]   (used when p.dx_sum is set)
  constexpr int U = kBwdApplyU;
  const int C = p.s.C0 + p.s.C1, cpg = C / G, vecs = C / 8;
  const int r0 = bx * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs, c0 = v * 8;
  const bool active = rsub < rows_par;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (active) {
    Aff ch; load_aff(ch, p.gamma, p.beta, c0, cpg, s_mean, s_rstd);
    float k1[8], k0[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { const int g = (c0 + j) / cpg; k1[j] = s_c1[g]; k0[j] = s_c0[g]; }
    const RowPtr rp = src_ptr(p.s, c0);
    const bool first = c0 < p.s.C0;
    const bf16* ad1 = first ? p.add1a : p.add1b;
    const long long ld_ad1 = first ? p.ld_add1a : p.ld_add1b;
    const int c1 = first ? c0 : c0 - p.s.C0;
    bf16* dxp = first ? p.dx0 : p.dx1;
    const long long ld_dx = first ? p.ld_dx0 : p.ld_dx1;
    const long long img = (long long)n * p.HW;
    const bool drop = MODE == 2 && p.drop_p > 0.f;
    Drop dr{};
    if (drop) dr = make_drop(p.drop_p, p.seed, p.seed_ptr);
    for (int r = r0 + rsub; r < r1; r += rows_par * U) {
      bf16x8 x[U], dy[U], sc[U], a0[U], a1[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = img + r + u * rows_par;
        const bool ok = r + u * rows_par < r1;
        x[u] = ok ? ld8(rp.p + row * rp.ld) : zero8();
        dy[u] = ok ? ld8(p.dy + row * p.ld_dy + c0) : zero8();
        if (MODE == 2) sc[u] = ok ? ld8(p.ss + row * p.ld_ss + c0) : zero8();
        a0[u] = (ok && p.add0) ? ld8(p.add0 + row * p.ld_add0 + c0) : zero8();
        a1[u] = (ok && ad1) ? ld8(ad1 + row * ld_ad1 + c1) : zero8();
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (r + u * rows_par >= r1) break;
        const long long row = img + r + u * rows_par;
        float f[8], d[8], s8[8], m[8], dz[8], dsc[8], dsh[8], t0[8], t1[8];
        b3d_unpack8(x[u], f); b3d_unpack8(dy[u], d);
        if (MODE == 2) {
          b3d_unpack8(sc[u], s8);
          if (drop) drop_mask8(dr, (unsigned long long)row * C + c0, m);
        }
        bwd_dz<MODE>(ch, f, d, s8, m, drop, dz, dsc, dsh);
        b3d_unpack8(a0[u], t0); b3d_unpack8(a1[u], t1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float g = fmaf(dz[j], ch.a[j], fmaf(f[j], k1[j], k0[j]));
          acc[j] += g;
          d[j] = fmaf(t0[j], p.add0_scale, g) + t1[j];
        }
        *reinterpret_cast<bf16x8*>(dxp + row * ld_dx + c1) = b3d_pack8(d);
      }
    }
  }
  if (p.dx_sum) {
    block_sum8(acc, active, v, rsub, rows_par, C, sm, sm + 2048);
    for (int c = threadIdx.x; c < C; c += blockDim.x) atomicAdd(&p.dx_sum[c], sm[2048 + c]);
  }
}
template <int MODE>
__global__ void __launch_bounds__(kT, MODE == 2 ? kBwdBlocks2 : kBwdBlocks1) gn_bwd_apply_kernel(BwdP p, int rows_per_block) {
  b3d_pdl_sync();
  extern __shared__ float sm[];
  __shared__ float s_mean[G], s_rstd[G], s_c1[G], s_c0[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y + p.n0;
  block_stats(p.stats, n, cpg, p.HW, s_mean, s_rstd);
  if (threadIdx.x < G)
    bwd_group_consts(p.ws[((long long)n * G + threadIdx.x) * 2], p.ws[((long long)n * G + threadIdx.x) * 2 + 1], cpg, p.HW, s_mean, s_rstd, s_c1,
                     s_c0);
  __syncthreads();
  bwd_apply_body<MODE>(p, rows_per_block, n, blockIdx.x, sm, s_mean, s_rstd, s_c1, s_c0);
}

// ---- Cluster-fused variants for tensors that stay in L2 between the passes (small per-GPU batches, the sampler): ONE launch
// per GroupNorm instead of two.  The blocks of an image form a thread-block cluster; each block runs pass 1 over its rows,
// the per-group partial sums meet through distributed shared memory (no global atomics, no workspace round trip), and after
// the cluster barrier each block runs pass 2 over the same rows, which it re-reads from L2.  At 16 pairs per GPU the model's
// 404 GroupNorm launches per step sit on the serial data-dependency chain; this halves them.
__device__ __forceinline__ void cluster_totals(const float* s_part, float* s_tot) {
  namespace cg = cooperative_groups;
  cg::cluster_group cl = cg::this_cluster();
  cl.sync();                                             // every block's partials are in its shared memory
  if (threadIdx.x < 2 * G) {
    float t = 0.f;
    const unsigned nb = cl.num_blocks();
    for (unsigned r = 0; r < nb; ++r) t += cl.map_shared_rank(s_part, r)[threadIdx.x];
    s_tot[threadIdx.x] = t;
  }
  cl.sync();                                             // nobody reads a peer's shared memory after this point
}
template <int MODE>
__global__ void __launch_bounds__(kT, MODE == 2 ? 3 : 4) gn_fwd_fused_kernel(ApplyP p, float* __restrict__ ws, int rows_per_block) {
  b3d_pdl_sync();
  extern __shared__ float sm[];
  __shared__ float s_part[2 * G], s_tot[2 * G], s_mean[G], s_rstd[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y + p.n0;
  stats_body(p.s, p.HW, rows_per_block, n, blockIdx.x, sm, nullptr, s_part);
  cluster_totals(s_part, s_tot);
  if (threadIdx.x < G) {
    const float cnt = (float)cpg * (float)p.HW;
    const float sum = s_tot[2 * threadIdx.x], sq = s_tot[2 * threadIdx.x + 1];
    const float mean = sum / cnt;
    s_mean[threadIdx.x] = mean;
    s_rstd[threadIdx.x] = rsqrtf(fmaxf(sq / cnt - mean * mean, 0.f) + 1e-5f);
    if (blockIdx.x == 0) {                               // the raw sums stay available to the backward pass
      ws[((long long)n * G + threadIdx.x) * 2] = sum;
      ws[((long long)n * G + threadIdx.x) * 2 + 1] = sq;
    }
  }
  __syncthreads();
  apply_body<MODE>(p, rows_per_block, n, blockIdx.x, s_mean, s_rstd);
}
template <int MODE>
__global__ void __launch_bounds__(kT, MODE == 2 ? kBwdBlocks2 : kBwdBlocks1) gn_bwd_fused_kernel(BwdP p, int rows_per_block) {
  b3d_pdl_sync();
  extern __shared__ float sm[];
  __shared__ float s_part[2 * G], s_tot[2 * G], s_mean[G], s_rstd[G], s_c1[G], s_c0[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y + p.n0;
  block_stats(p.stats, n, cpg, p.HW, s_mean, s_rstd);
  __syncthreads();
  bwd_reduce_body<MODE>(p, rows_per_block, n, blockIdx.x, sm, s_mean, s_rstd, s_part);
  cluster_totals(s_part, s_tot);
  if (threadIdx.x < G) {
    bwd_group_consts(s_tot[2 * threadIdx.x], s_tot[2 * threadIdx.x + 1], cpg, p.HW, s_mean, s_rstd, s_c1, s_c0);
    if (blockIdx.x == 0 && p.ws != nullptr) {            // the workspace ends up holding (S1, S2) like after the two-launch path
      p.ws[((long long)n * G + threadIdx.x) * 2] = s_tot[2 * threadIdx.x];
      p.ws[((long long)n * G + threadIdx.x) * 2 + 1] = s_tot[2 * threadIdx.x + 1];
    }
  }
  __syncthreads();
  bwd_apply_body<MODE>(p, rows_per_block, n, blockIdx.x, sm, s_mean, s_rstd, s_c1, s_c0);
}

// rows of one image per block: about `waves` full waves of `blocks_per_sm` resident blocks per SM over the whole launch
// (many short waves keep the last, partial one cheap), but never fewer rows than one unrolled sweep of the block
int pick_rows_per_block(int N, int HW, int C, int U, int blocks_per_sm, int waves = 8) {
  const int vecs = C / 8;
  const int rows_par = 256 / vecs > 0 ? 256 / vecs : 1;
  static const int forced = getenv("B3D_GN_RPB") ? atoi(getenv("B3D_GN_RPB")) : 0;      // measurement switch
  if (forced > 0) { int r = (forced + rows_par - 1) / rows_par * rows_par; return r > HW ? HW : r; }
  int per_img = (b3d_num_sms() * blocks_per_sm * waves + N - 1) / N;
  if (per_img < 1) per_img = 1;
  int rpb = (HW + per_img - 1) / per_img;
  const int min_rows = rows_par * U;
  // every block ends with O(C) fp32 atomics on per-channel sums shared by the whole launch: give it at least 64 KB of
  // each tensor to stream (rows * C * 2 B), or small images drown in atomic traffic (8-row blocks at C = 1024 did)
  // -- as long as the launch keeps two blocks per SM (tiny tensors are latency-bound: they want parallelism instead)
  const int amort_rows = (32768 + C - 1) / C;
  const long long par_rows = (long long)N * HW / (2 * b3d_num_sms());
  const int want = (int)(amort_rows < par_rows ? amort_rows : par_rows);
  if (rpb < want) rpb = want;
  if (rpb < min_rows) rpb = min_rows;
  rpb = (rpb + min_rows - 1) / min_rows * min_rows;
  if (rpb > HW) rpb = HW;
  return rpb;
}

// Image-chunked launches: the two passes of a GroupNorm (statistics then apply; reduce then apply in the backward) run over
// `chunk` images at a time, so the second pass re-reads the tensors from L2 instead of HBM.  gn_chunk_bytes is the budget
// for the tensors BOTH passes read (0 = one launch pair over all images).
long long g_gn_chunk_bytes = -1;
int gn_chunk_images(int N, int HW, int C, int bytes_per_elem) {
  if (g_gn_chunk_bytes < 0) {
    const char* e = getenv("B3D_GN_CHUNK_MB");
    g_gn_chunk_bytes = e ? atoll(e) * (1LL << 20) : 0;
  }
  if (g_gn_chunk_bytes <= 0) return N;
  const long long per_img = (long long)HW * C * bytes_per_elem;
  long long c = g_gn_chunk_bytes / (per_img > 0 ? per_img : 1);
  if (c >= N) return N;
  if (c < 2) c = 2;                 // images come in (cond, target) pairs
  if (c > 1) c &= ~1LL;
  // a chunk must still fill the GPU: at least ~4 MB per tensor in flight, else launch latencies dominate
  return (int)(c < N ? c : N);
}

// Cluster-fused single-launch GroupNorm (gn_fwd_fused_kernel / gn_bwd_fused_kernel): used when the input tensor is small enough
// to still be in L2 for the second pass.  Returns the cluster size (blocks per image, 1..8) or 0 for the two-launch path.
// The forward uses it by default (+3 % on the sampler's denoise step); the backward does not: at 16 pairs per GPU the fused
// backward measured 2 % SLOWER than the two launches (20.73 vs 21.12 ms per step) -- its two passes per block leave fewer
// independent blocks in flight than the separately sized launches.
long long g_gn_fuse_bytes = -1, g_gn_fuse_bwd_bytes = -1;
int gn_fuse_cluster(int N, int HW, int C, int U, int* rows_per_block, bool bwd = false) {
  if (g_gn_fuse_bytes < 0) {
    const char* e = getenv("B3D_GN_FUSE_MB");
    g_gn_fuse_bytes = (e ? atoll(e) : 40) * (1LL << 20);
  }
  if (g_gn_fuse_bwd_bytes < 0) {
    const char* e = getenv("B3D_GN_FUSE_BWD_MB");
    g_gn_fuse_bwd_bytes = (e ? atoll(e) : 0) * (1LL << 20);
  }
  if ((long long)N * HW * C * 2 > (bwd ? g_gn_fuse_bwd_bytes : g_gn_fuse_bytes)) return 0;
  const int vecs = C / 8;
  const int rows_par = 256 / vecs > 0 ? 256 / vecs : 1;
  const int min_rows = rows_par * U;
  int cl = 8;
  while (cl > 1 && (HW + cl - 1) / cl < min_rows) cl >>= 1;          // every block keeps at least one unrolled sweep of rows
  int rpb = (HW + cl - 1) / cl;
  rpb = (rpb + rows_par - 1) / rows_par * rows_par;
  *rows_per_block = rpb;
  return cl;
}

int check_src(const void* x0, int C0, const void* x1, int C1) {
  const int C = C0 + C1;
  B3D_CHECK_ARG(x0 != nullptr && C0 > 0, "groupnorm: null input");
  B3D_CHECK_ARG((x1 == nullptr) == (C1 == 0), "groupnorm: second source / C1 mismatch");
  B3D_CHECK_ARG(C % G == 0 && C0 % 8 == 0 && C1 % 8 == 0, "groupnorm: C=%d+%d must be a multiple of 32 (8 per source)", C0, C1);
  B3D_CHECK_ARG(C <= 2048, "groupnorm: C = %d > 2048 (one 256-thread block must cover all channel vectors)", C);
  return 0;
}

}  // namespace

extern "C" void b3d_set_gn_chunk_mb(int mb) { g_gn_chunk_bytes = (long long)mb * (1LL << 20); }
extern "C" void b3d_set_gn_fuse_mb(int mb) { g_gn_fuse_bytes = (long long)mb * (1LL << 20); }
extern "C" void b3d_set_gn_fuse_bwd_mb(int mb) { g_gn_fuse_bwd_bytes = (long long)mb * (1LL << 20); }

extern "C" int b3d_gn_stats(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums,
                            void* stream) {
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  Src2 s{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  const int C = C0 + C1, rpb = pick_rows_per_block(N, HW, C, 4, 8, 2);   // measured: 2 long waves beat 1, 4 and 8 (block epilogue cost)
  dim3 grid((HW + rpb - 1) / rpb, N);
  B3D_CUDA(b3d_launch_pdl(gn_stats_kernel, dim3(grid), dim3(kT), (2048 + 2 * C) * sizeof(float), (cudaStream_t)stream, s, HW, rpb, ws_sums, 0));
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}

static int gn_apply_impl(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums, int with_stats,
                         const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p,
                         uint64_t seed, const uint64_t* seed_ptr, void* out, int64_t ld_out, const int32_t* ss_image, void* stream) {
  const float* sums = ws_sums;
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  B3D_CHECK_ARG(mode >= 0 && mode <= 2, "gn_apply: mode");
  B3D_CHECK_ARG(mode != 2 || (ss != nullptr && F == C0 + C1), "gn_apply: film needs ss and F == C");
  B3D_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "gn_apply: drop_p");
  ApplyP p{};
  p.s = Src2{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  p.stats = sums; p.gamma = gamma; p.beta = beta; p.out = (bf16*)out; p.ld_out = ld_out; p.HW = HW; p.mode = mode;
  p.ss = (const bf16*)ss; p.ld_ss = ld_ss; p.F = F; p.drop_p = drop_p; p.seed = seed; p.seed_ptr = (const unsigned long long*)seed_ptr;
  p.ss_image = ss_image;
  // with_stats: statistics pass and apply pass of the same images back to back, chunk by chunk (the apply pass then reads
  // the chunk from L2); without: one apply launch over all images
  const int C = C0 + C1;
  if (with_stats) {
    int rpb = 0;
    const int cl = gn_fuse_cluster(N, HW, C, mode == 2 ? kApplyU2 : 4, &rpb);
    if (cl > 0) {
      p.n0 = 0;
      dim3 grid(cl, N);
      const size_t smem = (2048 + 2 * C) * sizeof(float);
      cudaStream_t st = (cudaStream_t)stream;
      if (mode == 0) B3D_CUDA(b3d_launch_cluster_pdl(gn_fwd_fused_kernel<0>, dim3(grid), dim3(kT), smem, st, cl, p, ws_sums, rpb));
      else if (mode == 1) B3D_CUDA(b3d_launch_cluster_pdl(gn_fwd_fused_kernel<1>, dim3(grid), dim3(kT), smem, st, cl, p, ws_sums, rpb));
      else B3D_CUDA(b3d_launch_cluster_pdl(gn_fwd_fused_kernel<2>, dim3(grid), dim3(kT), smem, st, cl, p, ws_sums, rpb));
      B3D_LAUNCH_CHECK();
      b3d_count_launch(1);
      return B3D_OK;
    }
  }
  const int chunk = with_stats ? gn_chunk_images(N, HW, C, 2) : N;
  int launches = 0;
  for (int n0 = 0; n0 < N; n0 += chunk) {
    const int nn = N - n0 < chunk ? N - n0 : chunk;
    p.n0 = n0;
    if (with_stats) {
      const int rpb = pick_rows_per_block(nn, HW, C, 4, 8, 2);   // measured: 2 long waves beat 1, 4 and 8 (block epilogue cost)
      dim3 grid((HW + rpb - 1) / rpb, nn);
      B3D_CUDA(b3d_launch_pdl(gn_stats_kernel, dim3(grid), dim3(kT), (2048 + 2 * C) * sizeof(float), (cudaStream_t)stream, p.s, HW, rpb, ws_sums, n0));
      B3D_LAUNCH_CHECK();
      ++launches;
    }
    const int rpb = pick_rows_per_block(nn, HW, C, mode == 2 ? kApplyU2 : kApplyU0, mode == 2 ? 3 : 4);
    dim3 grid((HW + rpb - 1) / rpb, nn);
    if (mode == 0) B3D_CUDA(b3d_launch_pdl(gn_apply_kernel<0>, dim3(grid), dim3(kT), 0, (cudaStream_t)stream, p, rpb));
    else if (mode == 1) B3D_CUDA(b3d_launch_pdl(gn_apply_kernel<1>, dim3(grid), dim3(kT), 0, (cudaStream_t)stream, p, rpb));
    else B3D_CUDA(b3d_launch_pdl(gn_apply_kernel<2>, dim3(grid), dim3(kT), 0, (cudaStream_t)stream, p, rpb));
    B3D_LAUNCH_CHECK();
    ++launches;
  }
  b3d_count_launch(launches);
  return B3D_OK;
}

extern "C" int b3d_gn_apply(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
                            const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p,
                            uint64_t seed, const uint64_t* seed_ptr, void* out, int64_t ld_out, const int32_t* ss_image, void* stream) {
  return gn_apply_impl(x0, C0, ld0, x1, C1, ld1, N, HW, const_cast<float*>(sums), 0, gamma, beta, mode, ss, ld_ss, F, drop_p, seed, seed_ptr, out,
                       ld_out, ss_image, stream);
}

extern "C" int b3d_gn_stats_apply(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums,
                                  const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p,
                                  uint64_t seed, const uint64_t* seed_ptr, void* out, int64_t ld_out, const int32_t* ss_image, void* stream) {
  return gn_apply_impl(x0, C0, ld0, x1, C1, ld1, N, HW, ws_sums, 1, gamma, beta, mode, ss, ld_ss, F, drop_p, seed, seed_ptr, out, ld_out, ss_image,
                       stream);
}

extern "C" int b3d_gn_bwd(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
                          const float* gamma, const float* beta, int mode, const void* dy, int64_t ld_dy, const void* ss, int64_t ld_ss,
                          int F, void* dss, int64_t ld_dss, float drop_p, uint64_t seed, const uint64_t* seed_ptr, float* dgamma, float* dbeta,
                          float* dss_sum, float* dx_sum, float* ws_s12,
                          void* dx0, int64_t ld_dx0, void* dx1, int64_t ld_dx1, const void* add0, int64_t ld_add0, const void* add1a,
                          int64_t ld_add1a, const void* add1b, int64_t ld_add1b, float add0_scale, void* stream) {
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  B3D_CHECK_ARG(mode >= 0 && mode <= 2, "gn_bwd: mode");
  B3D_CHECK_ARG(mode != 2 || (ss && dss && F == C0 + C1), "gn_bwd: film needs ss, dss, F == C");
  B3D_CHECK_ARG(dx0 && (C1 == 0 || dx1), "gn_bwd: dx outputs");
  BwdP p{};
  p.s = Src2{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  p.stats = sums; p.gamma = gamma; p.beta = beta; p.dy = (const bf16*)dy; p.ld_dy = ld_dy; p.HW = HW; p.mode = mode;
  p.ss = (const bf16*)ss; p.ld_ss = ld_ss; p.F = F; p.dss = (bf16*)dss; p.ld_dss = ld_dss; p.drop_p = drop_p; p.seed = seed; p.seed_ptr = (const unsigned long long*)seed_ptr;
  p.dgamma = dgamma; p.dbeta = dbeta; p.ws = ws_s12; p.dss_sum = dss_sum; p.dx_sum = dx_sum;
  p.dx0 = (bf16*)dx0; p.dx1 = (bf16*)dx1; p.ld_dx0 = ld_dx0; p.ld_dx1 = ld_dx1;
  p.add0 = (const bf16*)add0; p.ld_add0 = ld_add0; p.add0_scale = add0_scale;
  p.add1a = (const bf16*)add1a; p.add1b = (const bf16*)add1b; p.ld_add1a = ld_add1a; p.ld_add1b = ld_add1b;
  const int C = C0 + C1;
  cudaStream_t st = (cudaStream_t)stream;
  {
    int rpb = 0;
    const int cl = gn_fuse_cluster(N, HW, C, mode == 2 ? kBwdU2 : kBwdU1, &rpb, true);
    if (cl > 0) {
      p.n0 = 0;
      dim3 grid(cl, N);
      const size_t smem = (2048 + 4 * C) * sizeof(float);
      if (mode == 0) B3D_CUDA(b3d_launch_cluster_pdl(gn_bwd_fused_kernel<0>, dim3(grid), dim3(kT), smem, st, cl, p, rpb));
      else if (mode == 1) B3D_CUDA(b3d_launch_cluster_pdl(gn_bwd_fused_kernel<1>, dim3(grid), dim3(kT), smem, st, cl, p, rpb));
      else B3D_CUDA(b3d_launch_cluster_pdl(gn_bwd_fused_kernel<2>, dim3(grid), dim3(kT), smem, st, cl, p, rpb));
      B3D_LAUNCH_CHECK();
      b3d_count_launch(1);
      return B3D_OK;
    }
  }
  // bytes both passes read per element: x, dy (+ the FiLM scale)
  const int chunk = gn_chunk_images(N, HW, C, mode == 2 ? 6 : 4);
  int launches = 0;
  for (int n0 = 0; n0 < N; n0 += chunk) {
    const int nn = N - n0 < chunk ? N - n0 : chunk;
    p.n0 = n0;
    {
      const int rpb = pick_rows_per_block(nn, HW, C, mode == 2 ? kBwdU2 : kBwdU1, mode == 2 ? kBwdBlocks2 : kBwdBlocks1);
      dim3 grid((HW + rpb - 1) / rpb, nn);
      const size_t smem = (2048 + 4 * C) * sizeof(float);
      if (mode == 0) B3D_CUDA(b3d_launch_pdl(gn_bwd_reduce_kernel<0>, dim3(grid), dim3(kT), smem, st, p, rpb));
      else if (mode == 1) B3D_CUDA(b3d_launch_pdl(gn_bwd_reduce_kernel<1>, dim3(grid), dim3(kT), smem, st, p, rpb));
      else B3D_CUDA(b3d_launch_pdl(gn_bwd_reduce_kernel<2>, dim3(grid), dim3(kT), smem, st, p, rpb));
      B3D_LAUNCH_CHECK();
    }
    {
      const int rpb = pick_rows_per_block(nn, HW, C, kBwdApplyU, mode == 2 ? kBwdBlocks2 : kBwdBlocks1);
      dim3 grid((HW + rpb - 1) / rpb, nn);
      const size_t smem = (2048 + C) * sizeof(float);
      if (mode == 0) B3D_CUDA(b3d_launch_pdl(gn_bwd_apply_kernel<0>, dim3(grid), dim3(kT), smem, st, p, rpb));
      else if (mode == 1) B3D_CUDA(b3d_launch_pdl(gn_bwd_apply_kernel<1>, dim3(grid), dim3(kT), smem, st, p, rpb));
      else B3D_CUDA(b3d_launch_pdl(gn_bwd_apply_kernel<2>, dim3(grid), dim3(kT), smem, st, p, rpb));
      B3D_LAUNCH_CHECK();
    }
    launches += 2;
  }
  b3d_count_launch(launches);
  return B3D_OK;
}
