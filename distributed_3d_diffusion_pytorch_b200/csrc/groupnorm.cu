// GroupNorm(32 groups, eps 1e-5, per image) on bf16 NHWC tensors, with the consumers the model needs
// fused into the apply pass: SiLU (ResnetBlock GN0, lastgn), FiLM (+dropout) (ResnetBlock GN1), plain
// (AttnBlock).  Inputs may be the implicit channel-concat of two tensors (up-path skip connections).
// Replaces nn.GroupNorm / F.silu / FiLM apply / Dropout: xunet.py:61-71,87,140,143-145,198,535.
// HBM-bound: every pass is 16-byte vectorised over channels and reads each tensor once.
#include "../../include/b3d.h"
#include "b3d_common.cuh"

void b3d_count_launch(int n);

namespace {

constexpr int G = 32;
constexpr int kT = 256;

struct Src2 {
  const bf16* p0; const bf16* p1;
  int C0, C1; long long ld0, ld1;
};

__device__ __forceinline__ bf16x8 load8(const Src2& s, long long row, int c) {
  if (c < s.C0) return *reinterpret_cast<const bf16x8*>(s.p0 + row * s.ld0 + c);
  return *reinterpret_cast<const bf16x8*>(s.p1 + row * s.ld1 + (c - s.C0));
}

// ---- stats: raw (sum, sumsq) per (image, group) accumulated with fp32 atomics into ws[N,32,2] (pre-zeroed)
__global__ void gn_stats_kernel(Src2 s, int HW, int rows_per_block, float* __restrict__ ws) {
  extern __shared__ float sm[];  // [C][2]
  const int C = s.C0 + s.C1, cpg = C / G, n = blockIdx.y;
  const int vecs = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_block, r1 = min(HW, r0 + rows_per_block);
  const int lanes_per_row = vecs;                 // threads cooperating on one pixel row
  const int rows_par = max(1, (int)blockDim.x / lanes_per_row);
  const int v = threadIdx.x % lanes_per_row, rsub = threadIdx.x / lanes_per_row;
  if (rsub < rows_par) {
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = 0.f; b[i] = 0.f; }
    for (int r = r0 + rsub; r < r1; r += rows_par) {
      float f[8];
      b3d_unpack8(load8(s, (long long)n * HW + r, v * 8), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) { a[i] += f[i]; b[i] += f[i] * f[i]; }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) { atomicAdd(&sm[2 * (v * 8 + i)], a[i]); atomicAdd(&sm[2 * (v * 8 + i) + 1], b[i]); }
  }
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    float sa = 0.f, sb = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) { sa += sm[2 * c]; sb += sm[2 * c + 1]; }
    atomicAdd(&ws[((long long)n * G + g) * 2], sa);
    atomicAdd(&ws[((long long)n * G + g) * 2 + 1], sb);
  }
}

struct ApplyP {
  Src2 s;
  const float* stats;  // raw sums [N,32,2]
  const float* gamma; const float* beta;
  bf16* out; long long ld_out;
  int HW, mode;  // 0 plain, 1 silu, 2 film
  const bf16* ss; long long ld_ss; int F;  // film: scale = ss[row, c], shift = ss[row, F + c]
  float drop_p; unsigned long long seed; const unsigned long long* seed_ptr;  // effective seed = seed + *seed_ptr (device step counter)
};

__global__ void gn_apply_kernel(ApplyP p, int rows_per_block) {
  __shared__ float s_mean[G], s_rstd[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y, vecs = C / 8;
  if (threadIdx.x < G) {
    const float cnt = (float)cpg * (float)p.HW;
    const float sum = p.stats[((long long)n * G + threadIdx.x) * 2], sq = p.stats[((long long)n * G + threadIdx.x) * 2 + 1];
    const float mean = sum / cnt;
    const float var = fmaxf(sq / cnt - mean * mean, 0.f);
    s_mean[threadIdx.x] = mean; s_rstd[threadIdx.x] = rsqrtf(var + 1e-5f);
  }
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const long long total = (long long)(r1 - r0) * vecs;
  const float keep = 1.f / (1.f - p.drop_p);
  for (long long i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = r0 + (int)(i / vecs), c = (int)(i % vecs) * 8;
    const long long row = (long long)n * p.HW + r;
    float f[8];
    b3d_unpack8(load8(p.s, row, c), f);
    float sc[8], sh[8];
    if (p.mode == 2) {
      b3d_unpack8(*reinterpret_cast<const bf16x8*>(p.ss + row * p.ld_ss + c), sc);
      b3d_unpack8(*reinterpret_cast<const bf16x8*>(p.ss + row * p.ld_ss + p.F + c), sh);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int g = (c + j) / cpg;
      float y = (f[j] - s_mean[g]) * s_rstd[g] * __ldg(p.gamma + c + j) + __ldg(p.beta + c + j);
      if (p.mode == 1) y = b3d_silu(y);
      else if (p.mode == 2) {
        y = y * (1.f + sc[j]) + sh[j];
        if (p.drop_p > 0.f) y = (b3d_uniform(p.seed + (p.seed_ptr ? *p.seed_ptr : 0ULL), (unsigned long long)row * C + c + j) < p.drop_p) ? 0.f : y * keep;
      }
      f[j] = y;
    }
    *reinterpret_cast<bf16x8*>(p.out + row * p.ld_out + c) = b3d_pack8(f);
  }
}

struct BwdP {
  Src2 s;
  const float* stats; const float* gamma; const float* beta;
  const bf16* dy; long long ld_dy;
  int HW, mode;
  const bf16* ss; long long ld_ss; int F;
  bf16* dss; long long ld_dss;      // film: d_scale -> dss[row, c], d_shift -> dss[row, F + c]
  float drop_p; unsigned long long seed; const unsigned long long* seed_ptr;  // effective seed = seed + *seed_ptr (device step counter)
  float* dgamma; float* dbeta;      // fp32 [C], accumulated
  float* ws;                        // [N,32,2] (S1, S2), pre-zeroed
  // pass 2 outputs
  bf16* dx0; bf16* dx1; long long ld_dx0, ld_dx1;
  const bf16* add0; long long ld_add0;   // optional, full C wide, added to dx
  const bf16* add1a; const bf16* add1b; long long ld_add1a, ld_add1b;  // optional second addend, split like the sources
};

__device__ __forceinline__ void gn_bwd_dz(const BwdP& p, long long row, int c, int C, int cpg, const float* s_mean, const float* s_rstd,
                                          float* xh, float* dz, bool write_dss) {
  float f[8], dy[8];
  b3d_unpack8(load8(p.s, row, c), f);
  b3d_unpack8(*reinterpret_cast<const bf16x8*>(p.dy + row * p.ld_dy + c), dy);
  float sc[8], dsc[8], dsh[8];
  if (p.mode == 2) b3d_unpack8(*reinterpret_cast<const bf16x8*>(p.ss + row * p.ld_ss + c), sc);
  const float keep = 1.f / (1.f - p.drop_p);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c + j) / cpg;
    xh[j] = (f[j] - s_mean[g]) * s_rstd[g];
    const float z = xh[j] * __ldg(p.gamma + c + j) + __ldg(p.beta + c + j);
    float d = dy[j];
    if (p.mode == 1) d *= b3d_silu_grad(z);
    else if (p.mode == 2) {
      if (p.drop_p > 0.f) d = (b3d_uniform(p.seed + (p.seed_ptr ? *p.seed_ptr : 0ULL), (unsigned long long)row * C + c + j) < p.drop_p) ? 0.f : d * keep;
      dsc[j] = d * z; dsh[j] = d;
      d *= (1.f + sc[j]);
    }
    dz[j] = d;
  }
  if (p.mode == 2 && write_dss) {
    *reinterpret_cast<bf16x8*>(p.dss + row * p.ld_dss + c) = b3d_pack8(dsc);
    *reinterpret_cast<bf16x8*>(p.dss + row * p.ld_dss + p.F + c) = b3d_pack8(dsh);
  }
}

__device__ __forceinline__ void load_stats(const float* stats, int n, int cpg, int HW, float* s_mean, float* s_rstd) {
  if (threadIdx.x < G) {
    const float cnt = (float)cpg * (float)HW;
    const float sum = stats[((long long)n * G + threadIdx.x) * 2], sq = stats[((long long)n * G + threadIdx.x) * 2 + 1];
    const float mean = sum / cnt;
    const float var = fmaxf(sq / cnt - mean * mean, 0.f);
    s_mean[threadIdx.x] = mean; s_rstd[threadIdx.x] = rsqrtf(var + 1e-5f);
  }
}

// pass 1: d_scale/d_shift (film), dgamma/dbeta, per-(image,group) S1 = sum dz*gamma, S2 = sum dz*gamma*xhat
__global__ void gn_bwd_reduce_kernel(BwdP p, int rows_per_block) {
  extern __shared__ float sm[];  // [C][2] dgamma,dbeta partials then [G][2]
  __shared__ float s_mean[G], s_rstd[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y, vecs = C / 8;
  float* s_gs = sm + 2 * C;
  for (int i = threadIdx.x; i < 2 * C + 2 * G; i += blockDim.x) sm[i] = 0.f;
  load_stats(p.stats, n, cpg, p.HW, s_mean, s_rstd);
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs;
  if (rsub < rows_par) {
    float ag[8], ab[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { ag[j] = 0.f; ab[j] = 0.f; }
    for (int r = r0 + rsub; r < r1; r += rows_par) {
      float xh[8], dz[8];
      gn_bwd_dz(p, (long long)n * p.HW + r, v * 8, C, cpg, s_mean, s_rstd, xh, dz, true);
#pragma unroll
      for (int j = 0; j < 8; ++j) { ag[j] += dz[j] * xh[j]; ab[j] += dz[j]; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { atomicAdd(&sm[2 * (v * 8 + j)], ag[j]); atomicAdd(&sm[2 * (v * 8 + j) + 1], ab[j]); }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float dg = sm[2 * c], db = sm[2 * c + 1], gam = p.gamma[c];
    atomicAdd(&p.dgamma[c], dg);
    atomicAdd(&p.dbeta[c], db);
    atomicAdd(&s_gs[2 * (c / cpg)], db * gam);       // S1
    atomicAdd(&s_gs[2 * (c / cpg) + 1], dg * gam);   // S2
  }
  __syncthreads();
  if (threadIdx.x < G) {
    atomicAdd(&p.ws[((long long)n * G + threadIdx.x) * 2], s_gs[2 * threadIdx.x]);
    atomicAdd(&p.ws[((long long)n * G + threadIdx.x) * 2 + 1], s_gs[2 * threadIdx.x + 1]);
  }
}

// pass 2: dx = rstd * (dz*gamma - (S1 + xhat*S2)/cnt) [+ add0] [+ add1]
__global__ void gn_bwd_apply_kernel(BwdP p, int rows_per_block) {
  __shared__ float s_mean[G], s_rstd[G], s_s1[G], s_s2[G];
  const int C = p.s.C0 + p.s.C1, cpg = C / G, n = blockIdx.y, vecs = C / 8;
  load_stats(p.stats, n, cpg, p.HW, s_mean, s_rstd);
  if (threadIdx.x < G) {
    const float cnt = (float)cpg * (float)p.HW;
    s_s1[threadIdx.x] = p.ws[((long long)n * G + threadIdx.x) * 2] / cnt;
    s_s2[threadIdx.x] = p.ws[((long long)n * G + threadIdx.x) * 2 + 1] / cnt;
  }
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_block, r1 = min(p.HW, r0 + rows_per_block);
  const long long total = (long long)(r1 - r0) * vecs;
  for (long long i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = r0 + (int)(i / vecs), c = (int)(i % vecs) * 8;
    const long long row = (long long)n * p.HW + r;
    float xh[8], dz[8], dx[8];
    gn_bwd_dz(p, row, c, C, cpg, s_mean, s_rstd, xh, dz, false);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int g = (c + j) / cpg;
      dx[j] = s_rstd[g] * (dz[j] * __ldg(p.gamma + c + j) - s_s1[g] - xh[j] * s_s2[g]);
    }
    if (p.add0) {
      float a[8]; b3d_unpack8(*reinterpret_cast<const bf16x8*>(p.add0 + row * p.ld_add0 + c), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) dx[j] += a[j];
    }
    const bool first = c < p.s.C0;
    const bf16* ad = first ? p.add1a : p.add1b;
    if (ad) {
      float a[8];
      b3d_unpack8(*reinterpret_cast<const bf16x8*>(ad + row * (first ? p.ld_add1a : p.ld_add1b) + (first ? c : c - p.s.C0)), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) dx[j] += a[j];
    }
    if (first) *reinterpret_cast<bf16x8*>(p.dx0 + row * p.ld_dx0 + c) = b3d_pack8(dx);
    else *reinterpret_cast<bf16x8*>(p.dx1 + row * p.ld_dx1 + (c - p.s.C0)) = b3d_pack8(dx);
  }
}

int pick_rows_per_block(int N, int HW, int C) {
  // aim for >= ~4 blocks per SM overall and >= 8 KB of input per block
  long long per_row = (long long)C * 2;
  int rpb = (int)((16384 + per_row - 1) / per_row);
  if (rpb < 1) rpb = 1;
  long long blocks = (long long)N * ((HW + rpb - 1) / rpb);
  while (blocks > 148LL * 16 && rpb < HW) { rpb *= 2; blocks = (long long)N * ((HW + rpb - 1) / rpb); }
  return rpb;
}

int check_src(const void* x0, int C0, const void* x1, int C1) {
  const int C = C0 + C1;
  B3D_CHECK_ARG(x0 != nullptr && C0 > 0, "groupnorm: null input");
  B3D_CHECK_ARG((x1 == nullptr) == (C1 == 0), "groupnorm: second source / C1 mismatch");
  B3D_CHECK_ARG(C % G == 0 && C0 % 8 == 0 && C1 % 8 == 0, "groupnorm: C=%d+%d must be a multiple of 32 (8 per source)", C0, C1);
  B3D_CHECK_ARG(C <= 4096, "groupnorm: C too large");
  return 0;
}

}  // namespace

extern "C" int b3d_gn_stats(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums,
                            void* stream) {
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  Src2 s{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  const int C = C0 + C1, rpb = pick_rows_per_block(N, HW, C);
  dim3 grid((HW + rpb - 1) / rpb, N);
  gn_stats_kernel<<<grid, kT, 2 * C * sizeof(float), (cudaStream_t)stream>>>(s, HW, rpb, ws_sums);
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}

extern "C" int b3d_gn_apply(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
                            const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p,
                            uint64_t seed, const uint64_t* seed_ptr, void* out, int64_t ld_out, void* stream) {
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  B3D_CHECK_ARG(mode >= 0 && mode <= 2, "gn_apply: mode");
  B3D_CHECK_ARG(mode != 2 || (ss != nullptr && F == C0 + C1), "gn_apply: film needs ss and F == C");
  B3D_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "gn_apply: drop_p");
  ApplyP p{};
  p.s = Src2{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  p.stats = sums; p.gamma = gamma; p.beta = beta; p.out = (bf16*)out; p.ld_out = ld_out; p.HW = HW; p.mode = mode;
  p.ss = (const bf16*)ss; p.ld_ss = ld_ss; p.F = F; p.drop_p = drop_p; p.seed = seed; p.seed_ptr = (const unsigned long long*)seed_ptr;
  const int rpb = pick_rows_per_block(N, HW, C0 + C1);
  dim3 grid((HW + rpb - 1) / rpb, N);
  gn_apply_kernel<<<grid, kT, 0, (cudaStream_t)stream>>>(p, rpb);
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}

extern "C" int b3d_gn_bwd(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
                          const float* gamma, const float* beta, int mode, const void* dy, int64_t ld_dy, const void* ss, int64_t ld_ss,
                          int F, void* dss, int64_t ld_dss, float drop_p, uint64_t seed, const uint64_t* seed_ptr, float* dgamma, float* dbeta, float* ws_s12,
                          void* dx0, int64_t ld_dx0, void* dx1, int64_t ld_dx1, const void* add0, int64_t ld_add0, const void* add1a,
                          int64_t ld_add1a, const void* add1b, int64_t ld_add1b, void* stream) {
  if (int rc = check_src(x0, C0, x1, C1)) return rc;
  B3D_CHECK_ARG(mode >= 0 && mode <= 2, "gn_bwd: mode");
  B3D_CHECK_ARG(mode != 2 || (ss && dss && F == C0 + C1), "gn_bwd: film needs ss, dss, F == C");
  B3D_CHECK_ARG(dx0 && (C1 == 0 || dx1), "gn_bwd: dx outputs");
  BwdP p{};
  p.s = Src2{(const bf16*)x0, (const bf16*)x1, C0, C1, ld0, ld1};
  p.stats = sums; p.gamma = gamma; p.beta = beta; p.dy = (const bf16*)dy; p.ld_dy = ld_dy; p.HW = HW; p.mode = mode;
  p.ss = (const bf16*)ss; p.ld_ss = ld_ss; p.F = F; p.dss = (bf16*)dss; p.ld_dss = ld_dss; p.drop_p = drop_p; p.seed = seed; p.seed_ptr = (const unsigned long long*)seed_ptr;
  p.dgamma = dgamma; p.dbeta = dbeta; p.ws = ws_s12;
  p.dx0 = (bf16*)dx0; p.dx1 = (bf16*)dx1; p.ld_dx0 = ld_dx0; p.ld_dx1 = ld_dx1;
  p.add0 = (const bf16*)add0; p.ld_add0 = ld_add0;
  p.add1a = (const bf16*)add1a; p.add1b = (const bf16*)add1b; p.ld_add1a = ld_add1a; p.ld_add1b = ld_add1b;
  const int C = C0 + C1, rpb = pick_rows_per_block(N, HW, C);
  dim3 grid((HW + rpb - 1) / rpb, N);
  gn_bwd_reduce_kernel<<<grid, kT, (2 * C + 2 * G) * sizeof(float), (cudaStream_t)stream>>>(p, rpb);
  B3D_LAUNCH_CHECK();
  gn_bwd_apply_kernel<<<grid, kT, 0, (cudaStream_t)stream>>>(p, rpb);
  B3D_LAUNCH_CHECK();
  b3d_count_launch(2);
  return B3D_OK;
}
