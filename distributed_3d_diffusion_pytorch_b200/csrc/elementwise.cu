// HBM-bound kernels of the hot path: camera-ray + NeRF positional encoding (K1), DDPM noise-level
// embedding, layout packers, pooling, attention softmax, bias/column reductions, diffusion glue
// (q_sample, loss, sampler update) and fused Adam.  All coalesced / 16-byte vectorised.
#include "../../include/b3d.h"
#include "b3d_common.cuh"

void b3d_count_launch(int n);

namespace {

constexpr int kT = 256;
inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
inline int grid_for(long long n, int per_block = kT, int max_blocks = 148 * 32) {
  long long g = (n + per_block - 1) / per_block;
  if (g > max_blocks) g = max_blocks;
  if (g < 1) g = 1;
  return (int)g;
}

// ---------------------------------------------------------------------------------------------
// K1: rays (fp64: inv(K), R*Kinv*p, normalise -> fp32) + posenc_nerf(pos,0,15) ++ posenc_nerf(dir,0,8)
//     + cond-mask + pos_emb + first/other_emb  ->  [2B, H, W, ld] (bf16 or fp32), channels 0..143.
// Replaces visu3d rays + xunet.py:311-336.  One block = 32 consecutive pixels of one image row.
// ---------------------------------------------------------------------------------------------
constexpr int kPoseRows = 8;   // image rows per block
template <typename OutT>
__global__ void __launch_bounds__(256) pose_embedding_kernel(const double* __restrict__ Kmat, const double* __restrict__ R,
                                                             const double* __restrict__ t, const uint8_t* __restrict__ mask,
                                                             const float* __restrict__ pos_emb, const float* __restrict__ first_emb,
                                                             const float* __restrict__ other_emb, OutT* __restrict__ out, long long ld, int H,
                                                             int W, int vec_ok) {
  // One block = 32 pixels x kPoseRows rows of one image.  Per row the 32-pixel x 144-channel tile is assembled in shared
  // memory (phase A: one value per thread-iteration, pixel index fastest, so the learned embedding pos_emb[c,y,x] is read
  // coalesced along x; the channel is warp-uniform, so the sinf path does not diverge) and leaves as 16-byte stores of 8
  // channels (phase B: consecutive threads -> consecutive 16-byte pieces of the 288-byte pixel rows).  The per-image set-up
  // (R * inv(K) in fp64, the 93 position channels) is paid once per block, not once per row.  History: scalar 2-byte stores
  // with a div/mod pair per element ran at 0.45 TB/s, one row per block at 0.71 TB/s (instruction- and set-up-bound).
  constexpr int kPitch = 148;                       // floats per pixel row of the tile (16-byte aligned rows)
  __shared__ __align__(16) float s_out[2][32 * kPitch];
  __shared__ float s_posf[93];                      // position channel + first/other embedding (constant over the image)
  __shared__ float s_fe[144];                       // first / other embedding of this frame role
  __shared__ float s_dir[2][32][3];
  __shared__ float s_mul[51], s_ph[51];             // direction channels 93..143: multiplier 2^d, phase (0 or pi/2)
  __shared__ int s_ax[51];                          //   axis; -1 - axis for the three raw components
  __shared__ double s_M[9];
  const int n = blockIdx.z, y0 = blockIdx.y * kPoseRows, x0 = blockIdx.x * 32;
  const int b = n >> 1, f = n & 1;
  const bool cond = mask[b] != 0;
  const float* fe = f == 0 ? first_emb : other_emb;
  if (threadIdx.x == 0) {
    // M = R * inv(K)   (world_from_px for directions)
    const double* k = Kmat + b * 9;
    const double a = k[0], bb = k[1], c = k[2], d = k[3], e = k[4], ff = k[5], g = k[6], h = k[7], i = k[8];
    const double det = a * (e * i - ff * h) - bb * (d * i - ff * g) + c * (d * h - e * g);
    const double id = 1.0 / det;
    double inv[9] = {(e * i - ff * h) * id, (c * h - bb * i) * id, (bb * ff - c * e) * id,
                     (ff * g - d * i) * id, (a * i - c * g) * id, (c * d - a * ff) * id,
                     (d * h - e * g) * id, (bb * g - a * h) * id, (a * e - bb * d) * id};
    const double* r = R + (long long)n * 9;
    for (int p = 0; p < 3; ++p)
      for (int q = 0; q < 3; ++q) s_M[p * 3 + q] = r[p * 3 + 0] * inv[0 * 3 + q] + r[p * 3 + 1] * inv[1 * 3 + q] + r[p * 3 + 2] * inv[2 * 3 + q];
  }
  // position channels depend on the image only: [t(3), sin(t*2^d) (d-major), sin(t*2^d + pi/2)]
  if (threadIdx.x < 93) {
    const int c = threadIdx.x;
    float v;
    if (c < 3) v = (float)t[(long long)n * 3 + c];
    else {
      const int j = (c - 3) % 45, d = j / 3, ax = j % 3;
      float xb = (float)t[(long long)n * 3 + ax] * (float)(1 << d);
      if (c - 3 >= 45) xb = xb + 1.5707963705062866f;
      v = sinf(xb);
    }
    s_posf[c] = cond ? v : 0.f;                     // (the learned embeddings are added per element below, in the reference's order)
  } else if (threadIdx.x >= 128 && threadIdx.x < 128 + 51) {
    const int cc = threadIdx.x - 128;               // direction channels: [dir(3), sin(dir*2^d) (d-major, 8 octaves), sin(. + pi/2)]
    if (cc < 3) { s_ax[cc] = -1 - cc; s_mul[cc] = 1.f; s_ph[cc] = 0.f; }
    else {
      const int j = (cc - 3) % 24, d = j / 3;
      s_ax[cc] = j % 3; s_mul[cc] = (float)(1 << d); s_ph[cc] = cc - 3 >= 24 ? 1.5707963705062866f : 0.f;
    }
  }
  if (threadIdx.x < 144) s_fe[threadIdx.x] = fe ? fe[threadIdx.x] : 0.f;
  __syncthreads();
  const int xx = threadIdx.x & 31, w8 = threadIdx.x >> 5;
  const bool in_x = x0 + xx < W;
  const int rows = min(kPoseRows, H - y0);
  // ray directions of all rows of the block: 256 threads = 8 rows x 32 pixels, fp64 like the reference's camera maths
  // (two halves of s_dir when kPoseRows rows do not fit one buffer -- here they do: s_dir is indexed [row & 1] per row, so
  // compute them row by row inside the loop by the warp that owns the row parity)
  for (int ry = 0; ry < rows; ++ry) {
    const int y = y0 + ry, pb = ry & 1;
    float* tile = s_out[pb];
    // phase 0: the learned embeddings of the tile -- 18 independent, coalesced loads per thread in flight (ncu: with one
    // load -> add -> store chain per element the warps sat on the load latency) -- while warp 0's lanes also do the
    // fp64 ray directions of the row
    {
      const float* pe_row = pos_emb ? pos_emb + (long long)y * W + x0 + xx : nullptr;
      const long long cstride = (long long)H * W;
      float pe[18];
#pragma unroll
      for (int k = 0; k < 18; ++k) pe[k] = (in_x && pe_row) ? __ldg(pe_row + (w8 + 8 * k) * cstride) : 0.f;
      if (w8 == 0) {
        const double u = (double)(x0 + xx) + 0.5, v = (double)y + 0.5;
        double dx = s_M[0] * u + s_M[1] * v + s_M[2];
        double dy = s_M[3] * u + s_M[4] * v + s_M[5];
        double dz = s_M[6] * u + s_M[7] * v + s_M[8];
        const double inv_n = 1.0 / sqrt(dx * dx + dy * dy + dz * dz);
        s_dir[pb][xx][0] = (float)(dx * inv_n);
        s_dir[pb][xx][1] = (float)(dy * inv_n);
        s_dir[pb][xx][2] = (float)(dz * inv_n);
      }
#pragma unroll
      for (int k = 0; k < 18; ++k) {          // position channels are constants of the image: they go in here
        const int c = w8 + 8 * k;
        tile[xx * kPitch + c] = (c < 93 ? s_posf[c] : 0.f) + (in_x ? pe[k] + s_fe[c] : 0.f);
      }
    }
    __syncthreads();
    // phase A: the direction channels (sinf)
    if (cond) {
#pragma unroll 1
      for (int cc = w8; cc < 51; cc += 8) {
        const int ax = s_ax[cc];
        float v;
        if (ax < 0) v = s_dir[pb][xx][-1 - ax];
        else {
          float xb = s_dir[pb][xx][ax] * s_mul[cc];   // same fp32 operation order as posenc_nerf (xunet.py:49-59): x*2^d, + pi/2, sin
          if (cc >= 27) xb = xb + s_ph[cc];
          v = sinf(xb);
        }
        tile[xx * kPitch + 93 + cc] += v;
      }
    }
    __syncthreads();
    // phase B (the next row's phase A writes the other tile buffer, so no barrier is needed after the stores)
    OutT* orow = out + (((long long)n * H + y) * W + x0) * ld;
    if (vec_ok) {
      for (int i = threadIdx.x; i < 32 * 18; i += blockDim.x) {
        const int px = i / 18, c8 = i - px * 18;
        if (x0 + px >= W) break;
        const float4 lo = *reinterpret_cast<const float4*>(&tile[px * kPitch + c8 * 8]);
        const float4 hi = *reinterpret_cast<const float4*>(&tile[px * kPitch + c8 * 8 + 4]);
        OutT* dst = orow + (long long)px * ld + c8 * 8;
        if constexpr (sizeof(OutT) == 2) {
          const float f8[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
          *reinterpret_cast<bf16x8*>(dst) = b3d_pack8(f8);
        } else {
          *reinterpret_cast<float4*>(dst) = lo;
          *reinterpret_cast<float4*>(dst + 4) = hi;
        }
      }
    } else {
      for (int i = threadIdx.x; i < 32 * 144; i += blockDim.x) {
        const int px = i / 144, c = i - px * 144;
        if (x0 + px < W) orow[(long long)px * ld + c] = (OutT)tile[px * kPitch + c];
      }
    }
  }
}

// d pos_emb[c,y,x] = sum_f dP[f,y,x,c];  d first/other[c] = sum_{y,x} dP[f,y,x,c]     (dP fp32 [2,H,W,ld])
__global__ void pose_embedding_bwd_kernel(const float* __restrict__ dP, long long ld, int H, int W, float* dpos, float* dfirst, float* dother) {
  const int c = blockIdx.x;  // one block per channel
  float s0 = 0.f, s1 = 0.f;
  for (int i = threadIdx.x; i < H * W; i += blockDim.x) {
    const float a = dP[(long long)i * ld + c], b = dP[((long long)H * W + i) * ld + c];
    if (dpos) dpos[(long long)c * H * W + i] = a + b;
    s0 += a; s1 += b;
  }
  __shared__ float r0[32], r1[32];
  s0 = b3d_warp_sum(s0); s1 = b3d_warp_sum(s1);
  if ((threadIdx.x & 31) == 0) { r0[threadIdx.x >> 5] = s0; r1[threadIdx.x >> 5] = s1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int w = 0; w < (int)blockDim.x / 32; ++w) { a += r0[w]; b += r1[w]; }
    if (dfirst) dfirst[c] = a;
    if (dother) dother[c] = b;
  }
}

// posenc_ddpm(clip(logsnr,-20,20), emb_ch, max_time=1): [n] -> [n, emb_ch] bf16   (xunet.py:32-46,305-307)
__global__ void ddpm_embedding_kernel(const float* __restrict__ logsnr, bf16* __restrict__ out, int n, int emb_ch) {
  const int half = emb_ch / 2;
  const float k = -logf(10000.f) / (float)(half - 1);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)n * emb_ch; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / emb_ch), c = (int)(i % emb_ch);
    const float tt = fminf(fmaxf(logsnr[r], -20.f), 20.f) * 1000.f;
    const int j = c < half ? c : c - half;
    const float arg = tt * expf((float)j * k);
    out[i] = __float2bfloat16(c < half ? sinf(arg) : cosf(arg));
  }
}

// stack([x,z],1) NCHW fp32 -> [2B,H,W,8] bf16 (channels 3..7 zero)            (xunet.py:493-494)
__global__ void pack_input_kernel(const float* __restrict__ x, const float* __restrict__ z, bf16* __restrict__ out, int B, int HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < 2LL * B * HW; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / HW), p = (int)(i % HW), b = n >> 1;
    const float* src = (n & 1) ? z : x;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int c = 0; c < 3; ++c) f[c] = src[((long long)b * 3 + c) * HW + p];
    *reinterpret_cast<bf16x8*>(out + i * 8) = b3d_pack8(f);
  }
}
// d(stem input) [2B,H,W,8] bf16 -> dx, dz NCHW fp32
__global__ void unpack_input_grad_kernel(const bf16* __restrict__ g, float* __restrict__ dx, float* __restrict__ dz, int B, int HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < 2LL * B * HW; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / HW), p = (int)(i % HW), b = n >> 1;
    float f[8]; b3d_unpack8(*reinterpret_cast<const bf16x8*>(g + i * 8), f);
    float* dst = (n & 1) ? dz : dx;
    if (dst) for (int c = 0; c < 3; ++c) dst[((long long)b * 3 + c) * HW + p] = f[c];
  }
}
// lastconv output [2B,H,W,ld] bf16 -> eps fp32 [B,3,H,W], frame 1 only          (xunet.py:536)
__global__ void unpack_output_kernel(const bf16* __restrict__ y, long long ld, float* __restrict__ eps, int B, int HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)B * 3 * HW; i += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(i % HW), c = (int)((i / HW) % 3), b = (int)(i / (3LL * HW));
    eps[i] = __bfloat162float(y[(((long long)(2 * b + 1)) * HW + p) * ld + c]);
  }
}
__global__ void pack_output_grad_kernel(const float* __restrict__ deps, bf16* __restrict__ dy, long long ld, int B, int HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < 2LL * B * HW; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / HW), p = (int)(i % HW), b = n >> 1;
    for (int c = 0; c < (int)ld; ++c) {
      float v = 0.f;
      if ((n & 1) && c < 3) v = deps[((long long)b * 3 + c) * HW + p];
      dy[i * ld + c] = __float2bfloat16(v);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// weight packers: fp32 PyTorch layout -> bf16 K-major GEMM operands, and packed fp32 grads back
//   w [Co, Ci, taps]   (taps = kh*kw row-major)
//   mode 0 (forward)  : out[co, k_off + tap*Cpad + (ci-ci0)] for ci in [ci0, ci0+cn)         ld = out_ld
//   mode 1 (dgrad)    : out[ci, k_off + tap'*Copad + co], tap' = taps-1-tap (spatial flip)
//   mode 2 (tap rows) : out[tap*Ci + ci, co]
// unpack (mode 0 inverse, fp32 -> fp32): dw[co,ci,tap] = g[co, k_off + tap*Cpad + (ci-ci0)]
// ---------------------------------------------------------------------------------------------
__global__ void pack_weight_kernel(const float* __restrict__ w, int Co, int Ci, int taps, int mode, int ci0, int cn, int Cpad,
                                   long long k_off, bf16* __restrict__ out, long long out_ld, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float v = 0.f;
    long long o;
    if (mode == 0) {
      const int cc = (int)(i % Cpad), tap = (int)((i / Cpad) % taps), co = (int)(i / ((long long)Cpad * taps));
      if (cc < cn) v = w[((long long)co * Ci + ci0 + cc) * taps + tap];
      o = (long long)co * out_ld + k_off + (long long)tap * Cpad + cc;
    } else if (mode == 1) {
      const int co = (int)(i % Cpad), tp = (int)((i / Cpad) % taps), ci = (int)(i / ((long long)Cpad * taps));
      if (co < Co) v = w[((long long)co * Ci + ci) * taps + (taps - 1 - tp)];
      o = (long long)ci * out_ld + k_off + (long long)tp * Cpad + co;
    } else {
      const int co = (int)(i % Co), ci = (int)((i / Co) % Ci), tap = (int)(i / ((long long)Co * Ci));
      v = w[((long long)co * Ci + ci) * taps + tap];
      o = ((long long)tap * Ci + ci) * out_ld + co;
    }
    out[o] = __float2bfloat16(v);
  }
}
__global__ void unpack_wgrad_kernel(const float* __restrict__ g, long long g_ld, long long k_off, int Cpad, int Co, int Ci, int taps, int ci0,
                                    int cn, float* __restrict__ dw, int accumulate) {
  const long long total = (long long)Co * cn * taps;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % taps), cc = (int)((i / taps) % cn), co = (int)(i / ((long long)taps * cn));
    const float v = g[(long long)co * g_ld + k_off + (long long)tap * Cpad + cc];
    float* d = dw + ((long long)co * Ci + ci0 + cc) * taps + tap;
    *d = accumulate ? *d + v : v;
  }
}
__global__ void cast_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = __float2bfloat16(in[i]);
}

// ---------------------------------------------------------------------------------------------
// resampling on NHWC bf16 (xunet.py:17-28) and their adjoints
// ---------------------------------------------------------------------------------------------
__global__ void avgpool2_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int N, int H, int W, int C, float scale) {
  b3d_pdl_sync();
  const int Ho = H / 2, Wo = W / 2, vecs = C / 8;
  const long long total = (long long)N * Ho * Wo * vecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs); long long q = i / vecs;
    const int x = (int)(q % Wo); q /= Wo; const int y = (int)(q % Ho); const int n = (int)(q / Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float f[8];
        b3d_unpack8(*reinterpret_cast<const bf16x8*>(in + ((((long long)n * H + 2 * y + dy) * W + 2 * x + dx) * C + v * 8)), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= scale;
    *reinterpret_cast<bf16x8*>(out + (((long long)n * Ho + y) * Wo + x) * C + v * 8) = b3d_pack8(acc);
  }
}
// out[n, y, x] = scale * in[n, y/2, x/2]  (nearest upsample; also the adjoint of avgpool with scale=1/4)
__global__ void upsample2_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int N, int Ho, int Wo, int C, float scale) {
  b3d_pdl_sync();
  const int H = Ho / 2, W = Wo / 2, vecs = C / 8;
  const long long total = (long long)N * Ho * Wo * vecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs); long long q = i / vecs;
    const int x = (int)(q % Wo); q /= Wo; const int y = (int)(q % Ho); const int n = (int)(q / Ho);
    float f[8];
    b3d_unpack8(*reinterpret_cast<const bf16x8*>(in + (((long long)n * H + y / 2) * W + x / 2) * C + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= scale;
    *reinterpret_cast<bf16x8*>(out + (((long long)n * Ho + y) * Wo + x) * C + v * 8) = b3d_pack8(f);
  }
}

// ---------------------------------------------------------------------------------------------
// elementwise helpers
// ---------------------------------------------------------------------------------------------
// dx = dy * silu'(x)
__global__ void silu_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, bf16* __restrict__ dx, long long nvec) {
  b3d_pdl_sync();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8];
    b3d_unpack8(reinterpret_cast<const bf16x8*>(dy)[i], a);
    b3d_unpack8(reinterpret_cast<const bf16x8*>(x)[i], b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] *= b3d_silu_grad(b[j]);
    reinterpret_cast<bf16x8*>(dx)[i] = b3d_pack8(a);
  }
}
// dx = dy * silu'(x) over contiguous [rows, C] AND, from the same pass, the column sums of dx per group of rows_per_group
// rows (group_out[g, c] +=, e.g. d(logsnr embedding) per image) and over all rows (total_out[c] +=, the bias gradient of
// the layer that produced x).  Replaces silu_bwd + two colsum passes over the 1024-channel conditioning tensors.
__global__ void __launch_bounds__(256) silu_bwd_colsum_kernel(const bf16* dy, const bf16* __restrict__ x, bf16* dx, int C, long long rows_per_group,
                                                              int rows_per_block, float* __restrict__ group_out, long long group_ld,
                                                              float* __restrict__ total_out) {
  b3d_pdl_sync();
  extern __shared__ float sm[];   // [rows_par, C]
  constexpr int U = 4;
  const int vecs = C / 8, g = blockIdx.y;
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(rows_per_group, r0 + rows_per_block);
  const long long base = (long long)g * rows_per_group * C + v * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (rsub < rows_par) {
    for (long long r = r0 + rsub; r < r1; r += (long long)rows_par * U) {
      bf16x8 a[U], b[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + (long long)u * rows_par;
        if (rr < r1) {
          a[u].u = reinterpret_cast<const bf16x8*>(dy + base + rr * C)->u;
          if (x != nullptr) b[u].u = reinterpret_cast<const bf16x8*>(x + base + rr * C)->u;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + (long long)u * rows_par;
        if (rr >= r1) break;
        float f[8], z[8];
        b3d_unpack8(a[u], f);
        if (x != nullptr) {
          b3d_unpack8(b[u], z);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] *= b3d_silu_grad(z[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
        if (dx != nullptr) *reinterpret_cast<bf16x8*>(dx + base + rr * C) = b3d_pack8(f);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[rsub * C + v * 8 + j] = acc[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s_ = 0.f;
    for (int r = 0; r < rows_par; ++r) s_ += sm[r * C + c];
    if (group_out) atomicAdd(group_out + (long long)g * group_ld + c, s_);
    if (total_out) atomicAdd(total_out + c, s_);
  }
}
// out = scale * (a + b)   (b optional)
__global__ void add_scale_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, bf16* __restrict__ out, float scale, long long nvec) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8];
    b3d_unpack8(reinterpret_cast<const bf16x8*>(a)[i], x);
    if (b) { b3d_unpack8(reinterpret_cast<const bf16x8*>(b)[i], y);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] += y[j]; }
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] *= scale;
    reinterpret_cast<bf16x8*>(out)[i] = b3d_pack8(x);
  }
}

// out[g, c] (+)= sum over rows of group g of in[row, c];  rows_per_group rows per group   (bias / logsnr-emb grads)
__global__ void colsum_kernel(const bf16* __restrict__ in, long long ld, int C, long long rows_per_group, int rows_per_block,
                              float* __restrict__ out, long long out_ld, float scale) {
  const int g = blockIdx.z;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (c >= C) return;
  const long long r0 = (long long)blockIdx.y * rows_per_block, r1 = min(rows_per_group, r0 + rows_per_block);
  float s0 = 0.f, s1 = 0.f;
  const bf16* base = in + (g * rows_per_group) * ld + c;
  for (long long r = r0; r < r1; ++r) {
    const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(base + r * ld));
    s0 += v.x; s1 += v.y;
  }
  atomicAdd(out + g * out_ld + c, s0 * scale);
  if (c + 1 < C) atomicAdd(out + g * out_ld + c + 1, s1 * scale);
}

// the same for C % 8 == 0, C <= 2048: 16-byte loads, 4 rows in flight per thread, one block-level reduction and C
// atomics per block (the 2-channel-per-thread loop above keeps one 4-byte load in flight: latency-bound)
__global__ void __launch_bounds__(256) colsum_vec_kernel(const bf16* __restrict__ in, long long ld, int C, long long rows_per_group,
                                                         int rows_per_block, float* __restrict__ out, long long out_ld, float scale) {
  b3d_pdl_sync();
  extern __shared__ float sm[];   // [rows_par, C]
  constexpr int U = 4;
  const int vecs = C / 8, g = blockIdx.y;
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(rows_per_group, r0 + rows_per_block);
  const bf16* base = in + (long long)g * rows_per_group * ld + v * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (rsub < rows_par) {
    for (long long r = r0 + rsub; r < r1; r += (long long)rows_par * U) {
      bf16x8 x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + (long long)u * rows_par;
        x[u].u = rr < r1 ? reinterpret_cast<const bf16x8*>(base + rr * ld)->u : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float f[8]; b3d_unpack8(x[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[rsub * C + v * 8 + j] = acc[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rows_par; ++r) s += sm[r * C + c];
    atomicAdd(out + (long long)g * out_ld + c, s * scale);
  }
}

// out[f, p, c] = sum_b in[b, f, p, c]   (in: [B,2,P,C] bf16)   -- batch reduction per frame
__global__ void batch_frame_sum_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int B, long long PCvec) {
  b3d_pdl_sync();
  const long long total = 2 * PCvec;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int b = 0; b < B; ++b) {
      float f[8];
      b3d_unpack8(reinterpret_cast<const bf16x8*>(in)[(long long)b * total + i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
    reinterpret_cast<bf16x8*>(out)[i] = b3d_pack8(acc);
  }
}

// adjoint of the strided 3x3 conditioning convs w.r.t. their input, from per-tap products:
//   G [2*h*w, ldg] fp32, column tap*C + ci  (= dEsum . W[:,tap,:]);   dP[f, y', x', ci] += sum over taps
//   with (y'+1-ky) % s == 0, (x'+1-kx) % s == 0                                  (xunet.py:340-349)
__global__ void conv_tap_scatter_kernel(const float* __restrict__ Gm, long long ldg, int C, int h, int w, int s, int H, int W,
                                        float* __restrict__ dP, long long ldp, int accumulate) {
  const long long total = 2LL * H * W * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C); long long q = i / C;
    const int xp = (int)(q % W); q /= W; const int yp = (int)(q % H); const int f = (int)(q / H);
    float acc = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int yy = yp + 1 - ky;
      if (yy < 0 || yy % s) continue;
      const int y = yy / s; if (y >= h) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int xx = xp + 1 - kx;
        if (xx < 0 || xx % s) continue;
        const int x = xx / s; if (x >= w) continue;
        acc += Gm[(((long long)f * h + y) * w + x) * ldg + (ky * 3 + kx) * C + c];
      }
    }
    float* d = dP + (((long long)f * H + yp) * W + xp) * ldp + c;
    *d = accumulate ? *d + acc : acc;
  }
}

// ---------------------------------------------------------------------------------------------
// attention softmax over rows of length L (logits fp32 or bf16 in, bf16 probabilities out, fp32 math); one warp
// per row   (inside nn.MultiheadAttention).  ld_s / ld_p: row strides of the two tensors.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const bf16* p) { return __bfloat162float(*p); }

template <typename TIn>
__global__ void softmax_kernel(const TIn* __restrict__ S, bf16* __restrict__ P, long long rows, int L, long long ld_s, long long ld_p) {
  const long long row = (long long)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const TIn* s = S + row * ld_s;
  float m = -INFINITY;
  for (int i = lane; i < L; i += 32) m = fmaxf(m, ldf(s + i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float sum = 0.f;
  for (int i = lane; i < L; i += 32) sum += __expf(ldf(s + i) - m);
  sum = b3d_warp_sum(sum);
  const float inv = 1.f / sum;
  for (int i = lane; i < L; i += 32) P[row * ld_p + i] = __float2bfloat16(__expf(ldf(s + i) - m) * inv);
}
// dS = P * (dP - sum(P*dP))
template <typename TIn>
__global__ void softmax_bwd_kernel(const bf16* __restrict__ P, const TIn* __restrict__ dP, bf16* __restrict__ dS, long long rows, int L,
                                   long long ld_p, long long ld_dp, long long ld_ds) {
  b3d_pdl_sync();
  const long long row = (long long)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float dot = 0.f;
  for (int i = lane; i < L; i += 32) dot += __bfloat162float(P[row * ld_p + i]) * ldf(dP + row * ld_dp + i);
  dot = b3d_warp_sum(dot);
  for (int i = lane; i < L; i += 32) {
    const float p = __bfloat162float(P[row * ld_p + i]);
    dS[row * ld_ds + i] = __float2bfloat16(p * (ldf(dP + row * ld_dp + i) - dot));
  }
}

// ---------------------------------------------------------------------------------------------
// diffusion glue
// ---------------------------------------------------------------------------------------------
// z_t = sqrt(sigmoid(l)) z + sqrt(sigmoid(-l)) eps ; x_cond = mask ? x : x_rand      (train.py:50-60,95-96)
__global__ void q_sample_kernel(const float* __restrict__ z, const float* __restrict__ noise, const float* __restrict__ logsnr,
                                const float* __restrict__ x, const float* __restrict__ xrand, const uint8_t* __restrict__ mask,
                                float* __restrict__ z_out, float* __restrict__ x_out, long long per, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / per);
    const float l = logsnr[b];
    const float alpha = sqrtf(1.f / (1.f + expf(-l))), sigma = sqrtf(1.f / (1.f + expf(l)));
    z_out[i] = alpha * z[i] + sigma * noise[i];
    if (x_out) x_out[i] = mask[b] ? x[i] : xrand[i];
  }
}

// loss = mean(rho(noise - pred)), d loss / d pred; deterministic two-stage reduction through `partials`
// (one float per block + a ticket counter that the kernel leaves at zero).           (train.py:102-112)
__global__ void loss_kernel(const float* __restrict__ noise, const float* __restrict__ pred, float* __restrict__ dpred, long long n, int type,
                            float grad_scale, float* __restrict__ partials, unsigned int* __restrict__ ticket, float* __restrict__ loss_out) {
  float acc = 0.f;
  const float inv_n = 1.f / (float)n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float d = pred[i] - noise[i];
    float v, g;
    if (type == 0) { v = d * d; g = 2.f * d; }
    else if (type == 1) { v = fabsf(d); g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); }
    else { const float a = fabsf(d); v = a < 1.f ? 0.5f * d * d : a - 0.5f; g = a < 1.f ? d : (d > 0.f ? 1.f : -1.f); }
    acc += v;
    if (dpred) dpred[i] = g * inv_n * grad_scale;
  }
  __shared__ float red[32];
  __shared__ bool last;
  acc = b3d_warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (int)blockDim.x / 32; ++w) s += red[w];
    partials[blockIdx.x] = s;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x < 32) {
    __threadfence();
    float s = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += 32) s += partials[i];   // fixed order per lane -> deterministic
    s = b3d_warp_sum(s);
    if (threadIdx.x == 0) { *loss_out = s * inv_n; *ticket = 0u; }
  }
}

// sampler update (sampling.py:79-127): classifier-free guidance mix, x0 clip, posterior mean, + sqrt(var) * noise
__global__ void sampler_update_kernel(const float* __restrict__ z, const float* __restrict__ eps_c, const float* __restrict__ eps_u,
                                      const float* __restrict__ w, const float* __restrict__ noise, float* __restrict__ out, float alpha,
                                      float sigma, float alpha_next, float c, float std_next, int add_noise, long long per, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float ww = w[i / per];
    const float e = (1.f + ww) * eps_c[i] - ww * eps_u[i];
    const float zz = z[i];
    float x0 = (zz - sigma * e) / alpha;
    x0 = fminf(fmaxf(x0, -1.f), 1.f);
    float m = alpha_next * (zz * (1.f - c) / alpha + c * x0);
    if (add_noise) m += std_next * noise[i];
    out[i] = m;
  }
}

// same update with the two schedule values read from device memory (sched[0] = logsnr, sched[1] = logsnr_next), so a
// captured CUDA graph of one denoise step can be replayed for all 256 steps
__global__ void sampler_update_dev_kernel(const float* __restrict__ z, const float* __restrict__ eps_c, const float* __restrict__ eps_u,
                                          const float* __restrict__ w, const float* __restrict__ noise, float* __restrict__ out,
                                          const float* __restrict__ sched, long long per, long long total) {
  const float l0 = sched[0], l1 = sched[1];
  const float c = -expm1f(l0 - l1);
  const float alpha = sqrtf(1.f / (1.f + expf(-l0))), sigma = sqrtf(1.f / (1.f + expf(l0)));
  const float alpha_next = sqrtf(1.f / (1.f + expf(-l1))), std_next = sqrtf(c / (1.f + expf(l1)));
  const bool add_noise = l1 != 0.f;          // sampling.py:124
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float ww = w[i / per];
    const float e = (1.f + ww) * eps_c[i] - ww * eps_u[i];
    const float zz = z[i];
    float x0 = (zz - sigma * e) / alpha;
    x0 = fminf(fmaxf(x0, -1.f), 1.f);
    float m = alpha_next * (zz * (1.f - c) / alpha + c * x0);
    if (add_noise) m += std_next * noise[i];
    out[i] = m;
  }
}

// Sampler: out[j, p, :] = silu(c[j, p, :] + lemb[row[j], :]) for the D distinct conditionings of a batch.  c is the cached,
// step-invariant part of the conditioning path (conv_l(pose_emb) + bias, xunet.py:340-349, fp32 so that the sum below is the
// same fp32 value the conditioning conv's epilogue forms), lemb the per-step logsnr embedding (xunet.py:502).
__global__ void __launch_bounds__(256) cond_silu_kernel(const float* __restrict__ c, const float* __restrict__ lemb, const int* __restrict__ row,
                                                        bf16* __restrict__ out, long long per_image, int E, long long total8) {
  b3d_pdl_sync();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total8; i += (long long)gridDim.x * blockDim.x) {
    const long long e0 = i * 8;
    const int j = (int)(e0 / per_image), ch = (int)(e0 % E);
    const float4 a0 = *reinterpret_cast<const float4*>(c + e0), a1 = *reinterpret_cast<const float4*>(c + e0 + 4);
    const float* lr = lemb + (long long)__ldg(row + j) * E + ch;
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(lr)), b1 = __ldg(reinterpret_cast<const float4*>(lr) + 1);
    const float f[8] = {b3d_silu(a0.x + b0.x), b3d_silu(a0.y + b0.y), b3d_silu(a0.z + b0.z), b3d_silu(a0.w + b0.w),
                        b3d_silu(a1.x + b1.x), b3d_silu(a1.y + b1.y), b3d_silu(a1.z + b1.z), b3d_silu(a1.w + b1.w)};
    *reinterpret_cast<bf16x8*>(out + e0) = b3d_pack8(f);
  }
}

// fused Adam over one flat fp32 parameter arena (train.py:235: lr 1e-4, betas (0.9, 0.99), eps 1e-8)
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr, float b1, float b2, float eps, float ibc1, float ibc2,
                                         float grad_scale) {
  const float gg = g * grad_scale;
  m = b1 * m + (1.f - b1) * gg;
  v = b2 * v + (1.f - b2) * gg * gg;
  p -= lr * (m * ibc1) / (sqrtf(v * ibc2) + eps);
}
// 128-bit accesses, two independent vectors in flight per thread (28 bytes of traffic per parameter: HBM-bound).
// kEma: also ema = decay * ema + (1 - decay) * p_new (the parameter average lightning/diff3d.py:11-19 describes), +8 bytes.
// kG16: the gradient is bf16 (the all-reduced communication buffer of a data-parallel step, read in place: 26 bytes per parameter)
template <bool kG16>
__device__ __forceinline__ float4 adam_load_g(const void* g, long long i4) {
  if (!kG16) return reinterpret_cast<const float4*>(g)[i4];
  const uint2 r = reinterpret_cast<const uint2*>(g)[i4];
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.x)), b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
template <bool kEma, bool kG16 = false>
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const void* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                                   float* __restrict__ ema, long long n, float lr, float b1, float b2, float eps, float ibc1,
                                                   float ibc2, float grad_scale, float decay) {
  const long long n4 = n / 4, stride = (long long)gridDim.x * blockDim.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* m4 = reinterpret_cast<float4*>(m); float4* v4 = reinterpret_cast<float4*>(v); float4* e4 = reinterpret_cast<float4*>(ema);
  const float omd = 1.f - decay;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += 2 * stride) {
    const long long i1 = i + stride;
    const bool two = i1 < n4;
    float4 P0 = p4[i], G0 = adam_load_g<kG16>(g, i), M0 = m4[i], V0 = v4[i], P1, G1, M1, V1, E0, E1;
    if (kEma) E0 = e4[i];
    if (two) { P1 = p4[i1]; G1 = adam_load_g<kG16>(g, i1); M1 = m4[i1]; V1 = v4[i1]; if (kEma) E1 = e4[i1]; }
    adam_one(P0.x, G0.x, M0.x, V0.x, lr, b1, b2, eps, ibc1, ibc2, grad_scale); adam_one(P0.y, G0.y, M0.y, V0.y, lr, b1, b2, eps, ibc1, ibc2, grad_scale);
    adam_one(P0.z, G0.z, M0.z, V0.z, lr, b1, b2, eps, ibc1, ibc2, grad_scale); adam_one(P0.w, G0.w, M0.w, V0.w, lr, b1, b2, eps, ibc1, ibc2, grad_scale);
    p4[i] = P0; m4[i] = M0; v4[i] = V0;
    if (kEma) e4[i] = make_float4(decay * E0.x + omd * P0.x, decay * E0.y + omd * P0.y, decay * E0.z + omd * P0.z, decay * E0.w + omd * P0.w);
    if (two) {
      adam_one(P1.x, G1.x, M1.x, V1.x, lr, b1, b2, eps, ibc1, ibc2, grad_scale); adam_one(P1.y, G1.y, M1.y, V1.y, lr, b1, b2, eps, ibc1, ibc2, grad_scale);
      adam_one(P1.z, G1.z, M1.z, V1.z, lr, b1, b2, eps, ibc1, ibc2, grad_scale); adam_one(P1.w, G1.w, M1.w, V1.w, lr, b1, b2, eps, ibc1, ibc2, grad_scale);
      p4[i1] = P1; m4[i1] = M1; v4[i1] = V1;
      if (kEma) e4[i1] = make_float4(decay * E1.x + omd * P1.x, decay * E1.y + omd * P1.y, decay * E1.z + omd * P1.z, decay * E1.w + omd * P1.w);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - n4 * 4)) {
    const long long i = n4 * 4 + threadIdx.x;
    const float gi = kG16 ? __bfloat162float(reinterpret_cast<const bf16*>(g)[i]) : reinterpret_cast<const float*>(g)[i];
    adam_one(p[i], gi, m[i], v[i], lr, b1, b2, eps, ibc1, ibc2, grad_scale);
    if (kEma) ema[i] = decay * ema[i] + omd * p[i];
  }
}


// out = scale * a (bf16 [rows, C]) and colsum[c] += sum_rows out[row, c]   (the residual-branch gradient and the bias
// gradients that equal its column sums: conv2.bias, dense.bias, linear.bias)
__global__ void __launch_bounds__(256) scale_colsum_kernel(const bf16* __restrict__ a, bf16* __restrict__ out, float scale, long long rows, int C,
                                                           int rows_per_block, float* __restrict__ colsum) {
  b3d_pdl_sync();
  extern __shared__ float sm[];   // red [256*8] (plain stores: shared fp32 atomics are CAS spin loops)
  constexpr int U = 4;
  const int vecs = C / 8;
  const int rows_par = max(1, (int)blockDim.x / vecs);
  const int v = threadIdx.x % vecs, rsub = threadIdx.x / vecs;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(rows, r0 + rows_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (rsub < rows_par) {
    for (long long r = r0 + rsub; r < r1; r += (long long)rows_par * U) {
      bf16x8 x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + (long long)u * rows_par;
        if (rr < r1) x[u] = *reinterpret_cast<const bf16x8*>(a + rr * C + v * 8);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + (long long)u * rows_par;
        if (rr >= r1) break;
        float f[8]; b3d_unpack8(x[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= scale;
        *reinterpret_cast<bf16x8*>(out + rr * C + v * 8) = b3d_pack8(f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[rsub * C + v * 8 + j] = acc[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rows_par; ++r) s += sm[r * C + c];
    atomicAdd(&colsum[c], s);
  }
}

// many weight tensors in ONE launch (the per-step repack of every parameter / the unpack of every packed gradient):
// blocks are distributed over the jobs through a prefix table; semantics of each job = pack_weight / unpack_wgrad above,
// plus mode 3 = plain fp32 copy (FiLM bias concatenation).
struct PackJob {
  const float* src; void* dst;
  int Co, Ci, taps, mode, ci0, cn, Cpad, accumulate;
  long long k_off, ld, total;
  long long block0;   // first block of this job
};

__device__ __forceinline__ int find_job(const PackJob* jobs, int n_jobs, long long blk) {
  int lo = 0, hi = n_jobs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].block0 <= blk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// The per-step repack of ALL parameters: one launch over a table of groups.  A group = one fp32 source tensor [Co, Ci, taps]
// and up to kMaxPackDst destinations built from it (forward operand, per-source column ranges of a concatenated input,
// the [conv2 | skip] operand, the transposed operand of the data gradient ...): each 32 x 32 (x taps) source tile is read
// ONCE, with the source's own contiguity, and written to every destination from shared memory as 4-byte bf16 pairs (128
// contiguous bytes per warp).  Pad columns / rows of the destinations are never written: the caller zeroes the operands once.
// block -> group through a table (a binary search over block0 cost ~9 dependent L2 reads per block, 40 % of a block's life).
constexpr int kMaxPackDst = 8;
struct PackDst { void* dst; int mode, ci0, cn, Cpad; long long k_off, ld; };
struct PackGroup {
  const float* src;
  int Co, Ci, taps, n_dst;
  long long total, block0;      // total: element count of a mode-3 copy group
  PackDst d[kMaxPackDst];
};

__device__ __forceinline__ void store_pair(bf16* o, float v0, float v1, bool ok0, bool ok1) {
  if (ok0 && ok1 && (reinterpret_cast<uintptr_t>(o) & 3) == 0) *reinterpret_cast<__nv_bfloat162*>(o) = __floats2bfloat162_rn(v0, v1);
  else { if (ok0) o[0] = __float2bfloat16(v0); if (ok1) o[1] = __float2bfloat16(v1); }
}

template <int TAPS>
__device__ __forceinline__ void pack_group_tile(const PackGroup& G, long long blk, float* tile) {
  const float* __restrict__ w = G.src;
  const int taps = TAPS ? TAPS : G.taps, rl = 32 * taps, rs = rl + 1, n = 32 * rl, Co = G.Co, Ci = G.Ci;
  const int tiles_in = (Ci + 31) / 32;
  const int c_in0 = (int)(blk % tiles_in) * 32, co0 = (int)(blk / tiles_in) * 32;
  constexpr int U = 8;
  for (int e0 = threadIdx.x; e0 < n; e0 += 256 * U) {
    float v[U]; int idx[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int e = e0 + u * 256;
      const int co_l = e / rl, r = e - co_l * rl, ci_l = r / taps;
      idx[u] = e < n ? e + co_l : -1;
      v[u] = (e < n && co0 + co_l < Co && c_in0 + ci_l < Ci) ? w[((long long)(co0 + co_l) * Ci + c_in0) * taps + r] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) if (idx[u] >= 0) tile[idx[u]] = v[u];
  }
  __syncthreads();
  const int hl = 16 * taps;            // pairs per tile row
  for (int k = 0; k < G.n_dst; ++k) {
    const PackDst D = G.d[k];
    bf16* __restrict__ out = reinterpret_cast<bf16*>(D.dst);
    if (D.mode == 0) {        // out[co, k_off + tap*Cpad + (ci - ci0)] for ci in [ci0, ci0 + cn)
      if (c_in0 + 32 <= D.ci0 || c_in0 >= D.ci0 + D.cn) continue;
#pragma unroll 2
      for (int e = threadIdx.x; e < 32 * hl; e += 256) {
        const int co_l = e / hl, r = e - co_l * hl, tap = r >> 4, cp = (r & 15) * 2;
        const int co = co0 + co_l, cc = c_in0 + cp - D.ci0;
        if (co >= Co) continue;
        const float* t = tile + co_l * rs + cp * taps + tap;
        store_pair(out + (long long)co * D.ld + D.k_off + (long long)tap * D.Cpad + cc, t[0], t[taps], cc >= 0 && cc < D.cn, cc + 1 >= 0 && cc + 1 < D.cn);
      }
    } else {                  // mode 1: out[ci, k_off + tp*Cpad + co], taps flipped;  mode 2: out[(tp*Ci + ci), co]
#pragma unroll 2
      for (int e = threadIdx.x; e < 32 * hl; e += 256) {
        const int ci_l = e / hl, r = e - ci_l * hl, tp = r >> 4, cp = (r & 15) * 2;
        const int ci = c_in0 + ci_l, co = co0 + cp;
        if (ci >= Ci) continue;
        const float* t = tile + cp * rs + ci_l * taps + (D.mode == 1 ? taps - 1 - tp : tp);
        bf16* o = D.mode == 1 ? out + (long long)ci * D.ld + D.k_off + (long long)tp * D.Cpad + co : out + ((long long)tp * Ci + ci) * D.ld + co;
        store_pair(o, t[0], t[rs], co < Co, co + 1 < Co);
      }
    }
  }
}

__global__ void __launch_bounds__(256) pack_groups_kernel(const PackGroup* __restrict__ groups, const int* __restrict__ block_group) {
  __shared__ float tile[32 * (32 * 9 + 1)];
  const PackGroup& G = groups[block_group[blockIdx.x]];
  const long long blk = (long long)blockIdx.x - G.block0;
  if (G.d[0].mode == 3) {     // flat fp32 copy, 1024 values per block
    float* dst = reinterpret_cast<float*>(G.d[0].dst);
    for (int k = 0; k < 4; ++k) {
      const long long i = blk * 1024 + k * 256 + threadIdx.x;
      if (i < G.total) dst[i] = G.src[i];
    }
    return;
  }
  if (G.taps == 9) pack_group_tile<9>(G, blk, tile);
  else if (G.taps == 1) pack_group_tile<1>(G, blk, tile);
  else pack_group_tile<0>(G, blk, tile);
}

// Packed fp32 weight gradients [Co, k_off + tap*Cpad + cc] -> PyTorch layout [Co, Ci, taps], one 32 x 32 (x taps) tile per
// block through shared memory: 128-byte runs on the read side, 32*taps consecutive floats per output channel on the write
// side (indexing by destination element read with a 4*Cpad-byte stride between neighbouring threads).
__global__ void __launch_bounds__(256) unpack_multi_kernel(const PackJob* __restrict__ jobs, int n_jobs, const int* __restrict__ block_job) {
  __shared__ float tile[32 * (32 * 9 + 1)];
  const int j = block_job ? block_job[blockIdx.x] : find_job(jobs, n_jobs, blockIdx.x);
  const PackJob J = jobs[j];
  const long long blk = (long long)blockIdx.x - J.block0;
  const float* __restrict__ g = J.src;
  float* __restrict__ dw = reinterpret_cast<float*>(J.dst);
  const int taps = J.taps, rl = 32 * taps, rs = rl + 1, n = 32 * rl;
  const int tiles_in = (J.cn + 31) / 32;
  const int cc0 = (int)(blk % tiles_in) * 32, co0 = (int)(blk / tiles_in) * 32;
  constexpr int U = 8;
  for (int e0 = threadIdx.x; e0 < n; e0 += 256 * U) {
    float v[U]; int idx[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int e = e0 + u * 256;
      const int co_l = e / rl, r = e - co_l * rl, tap = r >> 5, cc_l = r & 31;
      const int co = co0 + co_l, cc = cc0 + cc_l;
      idx[u] = e < n ? co_l * rs + cc_l * taps + tap : -1;
      v[u] = (e < n && co < J.Co && cc < J.cn) ? g[(long long)co * J.ld + J.k_off + (long long)tap * J.Cpad + cc] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) if (idx[u] >= 0) tile[idx[u]] = v[u];
  }
  __syncthreads();
#pragma unroll 4
  for (int e = threadIdx.x; e < n; e += 256) {
    const int co_l = e / rl, r = e - co_l * rl, cc_l = r / taps;
    const int co = co0 + co_l, cc = cc0 + cc_l;
    if (co >= J.Co || cc >= J.cn) continue;
    float* d = dw + ((long long)co * J.Ci + J.ci0 + cc0) * taps + r;
    const float v = tile[co_l * rs + r];
    *d = J.accumulate ? *d + v : v;
  }
}

// y = W v (+ add) for fp32 W [R, C] (transpose = 0: one warp per row) or y = W^T v (+ add) (transpose = 1: a block covers 32
// columns, its 8 warps stride the rows -- 128-byte coalesced reads, 4 independent loads in flight -- and are summed through
// shared memory).  Weight-space helper of the composed attention output projection.
__global__ void __launch_bounds__(256) matvec_kernel(const float* __restrict__ W, int R, int C, const float* __restrict__ v,
                                                     const float* __restrict__ add, float* __restrict__ y, int transpose) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (!transpose) {
    const int r = blockIdx.x * 8 + warp;
    if (r >= R) return;
    float s0 = 0.f, s1 = 0.f;
    int c = lane;
    for (; c + 32 < C; c += 64) { s0 += W[(long long)r * C + c] * v[c]; s1 += W[(long long)r * C + c + 32] * v[c + 32]; }
    if (c < C) s0 += W[(long long)r * C + c] * v[c];
    const float s = b3d_warp_sum(s0 + s1);
    if (lane == 0) y[r] = s + (add ? add[r] : 0.f);
  } else {
    __shared__ float part[8][33];
    const int c = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c < C) {
      int r = warp;
      for (; r + 24 < R; r += 32) {
        s0 += W[(long long)r * C + c] * v[r]; s1 += W[(long long)(r + 8) * C + c] * v[r + 8];
        s2 += W[(long long)(r + 16) * C + c] * v[r + 16]; s3 += W[(long long)(r + 24) * C + c] * v[r + 24];
      }
      for (; r < R; r += 8) s0 += W[(long long)r * C + c] * v[r];
    }
    part[warp][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (warp == 0 && c < C) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += part[k][lane];
      y[c] = s + (add ? add[c] : 0.f);
    }
  }
}

// A[r, c] += u[r] * v[c]   (fp32 [R, C]; the b_out term of d W_lin in the composed attention output projection)
__global__ void __launch_bounds__(256) rank1_add_kernel(float* __restrict__ A, int R, int C, const float* __restrict__ u, const float* __restrict__ v) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)R * C) return;
  const int r = (int)(i / C), c = (int)(i - (long long)r * C);
  A[i] += u[r] * v[c];
}

}  // namespace

#define STREAM ((cudaStream_t)stream)
#define DONE(n) do { B3D_LAUNCH_CHECK(); b3d_count_launch(n); return B3D_OK; } while (0)

extern "C" int b3d_pose_embedding(const double* K, const double* R, const double* t, const uint8_t* cond_mask, const float* pos_emb,
                                  const float* first_emb, const float* other_emb, void* out, int out_fp32, int64_t ld, int B, int H, int W,
                                  void* stream) {
  B3D_CHECK_ARG(K && R && t && cond_mask && out, "pose_embedding: null pointer");
  B3D_CHECK_ARG(ld >= 144 && B > 0 && H > 0 && W > 0, "pose_embedding: bad sizes");
  dim3 grid(cdiv(W, 32), cdiv(H, kPoseRows), 2 * B);
  const int vec_ok = (ld % 8 == 0 && ((uintptr_t)out & 15) == 0) ? 1 : 0;      // 16-byte stores of 8 channels
  if (out_fp32) pose_embedding_kernel<float><<<grid, 256, 0, STREAM>>>(K, R, t, cond_mask, pos_emb, first_emb, other_emb, (float*)out, ld, H, W, vec_ok);
  else pose_embedding_kernel<bf16><<<grid, 256, 0, STREAM>>>(K, R, t, cond_mask, pos_emb, first_emb, other_emb, (bf16*)out, ld, H, W, vec_ok);
  DONE(1);
}
extern "C" int b3d_pose_embedding_bwd(const float* dP, int64_t ld, int H, int W, float* dpos_emb, float* dfirst, float* dother, void* stream) {
  B3D_CHECK_ARG(dP, "pose_embedding_bwd: null");
  pose_embedding_bwd_kernel<<<144, kT, 0, STREAM>>>(dP, ld, H, W, dpos_emb, dfirst, dother);
  DONE(1);
}
extern "C" int b3d_ddpm_embedding(const float* logsnr, void* out, int n, int emb_ch, void* stream) {
  B3D_CHECK_ARG(logsnr && out && emb_ch % 2 == 0 && emb_ch >= 4, "ddpm_embedding: args");
  ddpm_embedding_kernel<<<grid_for((long long)n * emb_ch), kT, 0, STREAM>>>(logsnr, (bf16*)out, n, emb_ch);
  DONE(1);
}
extern "C" int b3d_pack_input(const float* x, const float* z, void* out, int B, int HW, void* stream) {
  B3D_CHECK_ARG(x && z && out, "pack_input: null");
  pack_input_kernel<<<grid_for(2LL * B * HW), kT, 0, STREAM>>>(x, z, (bf16*)out, B, HW);
  DONE(1);
}
extern "C" int b3d_unpack_input_grad(const void* g, float* dx, float* dz, int B, int HW, void* stream) {
  unpack_input_grad_kernel<<<grid_for(2LL * B * HW), kT, 0, STREAM>>>((const bf16*)g, dx, dz, B, HW);
  DONE(1);
}
extern "C" int b3d_unpack_output(const void* y, int64_t ld, float* eps, int B, int HW, void* stream) {
  unpack_output_kernel<<<grid_for(3LL * B * HW), kT, 0, STREAM>>>((const bf16*)y, ld, eps, B, HW);
  DONE(1);
}
extern "C" int b3d_pack_output_grad(const float* deps, void* dy, int64_t ld, int B, int HW, void* stream) {
  pack_output_grad_kernel<<<grid_for(2LL * B * HW), kT, 0, STREAM>>>(deps, (bf16*)dy, ld, B, HW);
  DONE(1);
}
extern "C" int b3d_pack_weight(const float* w, int Co, int Ci, int taps, int mode, int ci0, int cn, int Cpad, int64_t k_off, void* out,
                               int64_t out_ld, void* stream) {
  B3D_CHECK_ARG(w && out && mode >= 0 && mode <= 2, "pack_weight: args");
  long long total = mode == 0 ? (long long)Co * taps * Cpad : mode == 1 ? (long long)Ci * taps * Cpad : (long long)Co * Ci * taps;
  pack_weight_kernel<<<grid_for(total), kT, 0, STREAM>>>(w, Co, Ci, taps, mode, ci0, cn, Cpad, k_off, (bf16*)out, out_ld, total);
  DONE(1);
}
extern "C" int b3d_unpack_wgrad(const float* g, int64_t g_ld, int64_t k_off, int Cpad, int Co, int Ci, int taps, int ci0, int cn, float* dw,
                                int accumulate, void* stream) {
  unpack_wgrad_kernel<<<grid_for((long long)Co * cn * taps), kT, 0, STREAM>>>(g, g_ld, k_off, Cpad, Co, Ci, taps, ci0, cn, dw, accumulate);
  DONE(1);
}
extern "C" int b3d_cast_bf16(const float* in, void* out, int64_t n, void* stream) {
  cast_kernel<<<grid_for(n), kT, 0, STREAM>>>(in, (bf16*)out, n);
  DONE(1);
}
extern "C" int b3d_matvec(const float* W, int R, int C, const float* v, const float* add, float* y, int transpose, void* stream) {
  B3D_CHECK_ARG(W && v && y && R > 0 && C > 0, "matvec: args");
  matvec_kernel<<<transpose ? cdiv(C, 32) : cdiv(R, 8), 256, 0, STREAM>>>(W, R, C, v, add, y, transpose);
  DONE(1);
}
extern "C" int b3d_rank1_add(float* A, int R, int C, const float* u, const float* v, void* stream) {
  B3D_CHECK_ARG(A && u && v && R > 0 && C > 0, "rank1_add: args");
  rank1_add_kernel<<<cdiv((long long)R * C, 256), 256, 0, STREAM>>>(A, R, C, u, v);
  DONE(1);
}
extern "C" int b3d_avgpool2(const void* in, void* out, int N, int H, int W, int C, float scale, void* stream) {
  B3D_CHECK_ARG(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "avgpool2: C%%8, even H/W");
  B3D_CUDA(b3d_launch_pdl(avgpool2_kernel, dim3(grid_for((long long)N * (H / 2) * (W / 2) * (C / 8))), dim3(kT), 0, STREAM, (const bf16*)in, (bf16*)out, N, H, W, C, scale));
  DONE(1);
}
extern "C" int b3d_upsample2(const void* in, void* out, int N, int Ho, int Wo, int C, float scale, void* stream) {
  B3D_CHECK_ARG(C % 8 == 0 && Ho % 2 == 0 && Wo % 2 == 0, "upsample2: C%%8, even H/W");
  B3D_CUDA(b3d_launch_pdl(upsample2_kernel, dim3(grid_for((long long)N * Ho * Wo * (C / 8))), dim3(kT), 0, STREAM, (const bf16*)in, (bf16*)out, N, Ho, Wo, C, scale));
  DONE(1);
}
extern "C" int b3d_silu_bwd(const void* dy, const void* x, void* dx, int64_t n, void* stream) {
  B3D_CHECK_ARG(n % 8 == 0, "silu_bwd: n%%8");
  B3D_CUDA(b3d_launch_pdl(silu_bwd_kernel, dim3(grid_for(n / 8)), dim3(kT), 0, STREAM, (const bf16*)dy, (const bf16*)x, (bf16*)dx, n / 8));
  DONE(1);
}
extern "C" int b3d_silu_bwd_colsum(const void* dy, const void* x, void* dx, int C, int64_t rows_per_group, int groups, float* group_out,
                                   int64_t group_ld, float* total_out, void* stream) {
  B3D_CHECK_ARG(dy && ((x && dx) || (!x && !dx)) && C % 8 == 0 && C <= 2048 && rows_per_group > 0 && groups > 0,
                "silu_bwd_colsum: args (C %% 8, C <= 2048; x and dx both or neither)");
  const int rows_par = 256 / (C / 8) > 0 ? 256 / (C / 8) : 1;
  const int min_rows = rows_par * 4;
  long long rpb = cdiv(rows_per_group * groups, 148LL * 8);
  const long long amort = cdiv(32768, C);
  if (rpb < amort) rpb = amort;
  if (rpb < min_rows) rpb = min_rows;
  rpb = (rpb + min_rows - 1) / min_rows * min_rows;
  dim3 grid((unsigned)cdiv(rows_per_group, rpb), groups);
  B3D_CUDA(b3d_launch_pdl(silu_bwd_colsum_kernel, grid, dim3(256), (size_t)rows_par * C * sizeof(float), STREAM, (const bf16*)dy, (const bf16*)x,
                          (bf16*)dx, C, (long long)rows_per_group, (int)rpb, group_out, (long long)group_ld, total_out));
  DONE(1);
}
extern "C" int b3d_add_scale(const void* a, const void* b, void* out, float scale, int64_t n, void* stream) {
  B3D_CHECK_ARG(n % 8 == 0, "add_scale: n%%8");
  add_scale_kernel<<<grid_for(n / 8), kT, 0, STREAM>>>((const bf16*)a, (const bf16*)b, (bf16*)out, scale, n / 8);
  DONE(1);
}
extern "C" int b3d_colsum(const void* in, int64_t ld, int C, int64_t rows_per_group, int groups, float* out, int64_t out_ld, float scale,
                          void* stream) {
  B3D_CHECK_ARG(C % 2 == 0 && ld % 2 == 0, "colsum: even C/ld");
  if (C % 8 == 0 && ld % 8 == 0 && C <= 2048 && (((uintptr_t)in) & 15) == 0) {
    const int rows_par = 256 / (C / 8) > 0 ? 256 / (C / 8) : 1;
    const int min_rows = rows_par * 4;
    long long rpb = cdiv(rows_per_group * groups, 148LL * 8);      // about 8 blocks per SM over the launch
    const long long amort = cdiv(32768, C);                        // >= 64 KB streamed per block of C atomics
    if (rpb < amort) rpb = amort;
    if (rpb < min_rows) rpb = min_rows;
    rpb = (rpb + min_rows - 1) / min_rows * min_rows;
    dim3 grid((unsigned)cdiv(rows_per_group, rpb), groups);
    B3D_CUDA(b3d_launch_pdl(colsum_vec_kernel, dim3(grid), dim3(256), (size_t)rows_par * C * sizeof(float), STREAM, (const bf16*)in, ld, C, rows_per_group, (int)rpb, out, out_ld, scale));
    DONE(1);
  }
  int rpb = 256;
  while ((long long)cdiv(rows_per_group, rpb) * groups * cdiv(C, 256) > 148LL * 16 && rpb < rows_per_group) rpb *= 2;
  dim3 grid(cdiv(C / 2, 128), cdiv(rows_per_group, rpb), groups);
  colsum_kernel<<<grid, 128, 0, STREAM>>>((const bf16*)in, ld, C, rows_per_group, rpb, out, out_ld, scale);
  DONE(1);
}
extern "C" int b3d_batch_frame_sum(const void* in, void* out, int B, int64_t per_frame_elems, void* stream) {
  B3D_CHECK_ARG(per_frame_elems % 8 == 0, "batch_frame_sum: %%8");
  B3D_CUDA(b3d_launch_pdl(batch_frame_sum_kernel, dim3(grid_for(2 * per_frame_elems / 8)), dim3(kT), 0, STREAM, (const bf16*)in, (bf16*)out, B, per_frame_elems / 8));
  DONE(1);
}
extern "C" int b3d_conv_tap_scatter(const float* G, int64_t ldg, int C, int h, int w, int s, int H, int W, float* dP, int64_t ldp, int accumulate,
                                    void* stream) {
  conv_tap_scatter_kernel<<<grid_for(2LL * H * W * C), kT, 0, STREAM>>>(G, ldg, C, h, w, s, H, W, dP, ldp, accumulate);
  DONE(1);
}
extern "C" int b3d_softmax(const void* S, int s_fp32, int64_t ld_s, void* P, int64_t ld_p, int64_t rows, int L, void* stream) {
  if (s_fp32) softmax_kernel<float><<<cdiv(rows, 8), 256, 0, STREAM>>>((const float*)S, (bf16*)P, rows, L, ld_s, ld_p);
  else softmax_kernel<bf16><<<cdiv(rows, 8), 256, 0, STREAM>>>((const bf16*)S, (bf16*)P, rows, L, ld_s, ld_p);
  DONE(1);
}
extern "C" int b3d_softmax_bwd(const void* P, int64_t ld_p, const void* dP, int dp_fp32, int64_t ld_dp, void* dS, int64_t ld_ds, int64_t rows,
                               int L, void* stream) {
  if (dp_fp32) B3D_CUDA(b3d_launch_pdl(softmax_bwd_kernel<float>, dim3(cdiv(rows, 8)), dim3(256), 0, STREAM, (const bf16*)P, (const float*)dP, (bf16*)dS, rows, L, ld_p, ld_dp, ld_ds));
  else B3D_CUDA(b3d_launch_pdl(softmax_bwd_kernel<bf16>, dim3(cdiv(rows, 8)), dim3(256), 0, STREAM, (const bf16*)P, (const bf16*)dP, (bf16*)dS, rows, L, ld_p, ld_dp, ld_ds));
  DONE(1);
}
extern "C" int b3d_q_sample(const float* z, const float* noise, const float* logsnr, const float* x, const float* xrand, const uint8_t* mask,
                            float* z_out, float* x_out, int B, int64_t per_sample, void* stream) {
  B3D_CHECK_ARG(z && noise && logsnr && z_out, "q_sample: null");
  B3D_CHECK_ARG(!x_out || (x && xrand && mask), "q_sample: cond replacement needs x, xrand, mask");
  q_sample_kernel<<<grid_for((long long)B * per_sample), kT, 0, STREAM>>>(z, noise, logsnr, x, xrand, mask, z_out, x_out, per_sample,
                                                                        (long long)B * per_sample);
  DONE(1);
}
#define B3D_LOSS_MAX_BLOCKS 1024
extern "C" int b3d_loss_scratch_bytes(void) { return (B3D_LOSS_MAX_BLOCKS + 1) * 4; }
extern "C" int b3d_loss(const float* noise, const float* pred, float* dpred, int64_t n, int loss_type, float grad_scale, void* scratch,
                        float* loss_out, void* stream) {
  B3D_CHECK_ARG(noise && pred && scratch && loss_out && n > 0, "loss: null/empty");
  B3D_CHECK_ARG(loss_type >= 0 && loss_type <= 2, "loss: type must be 0 (l2), 1 (l1) or 2 (huber)");
  const int blocks = grid_for(n, kT * 4, B3D_LOSS_MAX_BLOCKS);
  float* partials = (float*)scratch;
  unsigned int* ticket = (unsigned int*)scratch + B3D_LOSS_MAX_BLOCKS;   // must be zero on entry; left at zero
  loss_kernel<<<blocks, kT, 0, STREAM>>>(noise, pred, dpred, n, loss_type, grad_scale, partials, ticket, loss_out);
  DONE(1);
}
// ---------------------------------------------------------------------------------------------
// b3d_loss_static: the same loss kernel for callers that have no device memory to hand in as scratch.
//
// Who that is: code that runs inside someone else's memory budget.  A trainer that carved the whole HBM
// into its own arenas at start-up, or a service that forbids driver allocations after initialisation
// because cudaMalloc/cudaFree stall every stream of the device, cannot produce the small scratch buffer
// b3d_loss asks for at an arbitrary call site.  Here the block partials and the ticket live in a
// fixed array that is part of the library image: every GPU that loads libb3d.so carries these
// (B3D_LOSS_STATIC_MAX_BLOCKS + 1) * 4 = 4100 bytes until the process exits.
//
// Limits and rules:
//  * n <= B3D_LOSS_STATIC_MAX_ELEMS (4 194 304 values = 1024 blocks x 4096); larger inputs are refused.
//  * The array is shared by every call on the device and still holds the previous call's partials; the
//    ticket word is cleared on `stream` before each launch.  Two calls that overlap on the GPU (different
//    streams, no ordering) race on it -- serialise them.
//  * The first call on each device looks up the array's device address (host code cannot use the symbol
//    name as a pointer).  Make that first call outside any stream capture.
//  * No cudaMalloc / cudaFree / allocating library call is made here.
// ---------------------------------------------------------------------------------------------
#define B3D_LOSS_STATIC_MAX_BLOCKS 1024
#define B3D_LOSS_STATIC_ELEMS_PER_BLOCK 4096
#define B3D_LOSS_STATIC_MAX_ELEMS ((int64_t)B3D_LOSS_STATIC_MAX_BLOCKS * B3D_LOSS_STATIC_ELEMS_PER_BLOCK)
static __device__ __align__(256) unsigned int g_loss_static_scratch[B3D_LOSS_STATIC_MAX_BLOCKS + 1];
static void* g_loss_static_addr[64] = {nullptr};   // per device, filled by the first call on that device

extern "C" int64_t b3d_loss_static_max_elems(void) { return B3D_LOSS_STATIC_MAX_ELEMS; }
extern "C" int b3d_loss_static(const float* noise, const float* pred, float* dpred, int64_t n, int loss_type, float grad_scale,
                               float* loss_out, void* stream) {
  B3D_CHECK_ARG(noise && pred && loss_out && n > 0, "loss_static: null/empty");
  B3D_CHECK_ARG(loss_type >= 0 && loss_type <= 2, "loss_static: type must be 0 (l2), 1 (l1) or 2 (huber)");
  B3D_CHECK_ARG(n <= B3D_LOSS_STATIC_MAX_ELEMS,
                "loss_static: n = %lld exceeds the built-in scratch (maximum %lld elements); use b3d_loss with caller scratch",
                (long long)n, (long long)B3D_LOSS_STATIC_MAX_ELEMS);
  int dev = 0;
  B3D_CUDA(cudaGetDevice(&dev));
  B3D_CHECK_ARG(dev >= 0 && dev < 64, "loss_static: device index %d out of range", dev);
  if (g_loss_static_addr[dev] == nullptr) {
    void* p = nullptr;
    B3D_CUDA(cudaGetSymbolAddress((void**)&p, g_loss_static_scratch));   // with `dev` current: this device's copy
    g_loss_static_addr[dev] = p;
  }
  float* partials = (float*)g_loss_static_addr[dev];
  unsigned int* ticket = (unsigned int*)g_loss_static_addr[dev] + B3D_LOSS_STATIC_MAX_BLOCKS;
  B3D_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), STREAM));       // must start at zero; partials are overwritten
  const int blocks = (int)((n + B3D_LOSS_STATIC_ELEMS_PER_BLOCK - 1) / B3D_LOSS_STATIC_ELEMS_PER_BLOCK);
  loss_kernel<<<blocks, kT, 0, STREAM>>>(noise, pred, dpred, n, loss_type, grad_scale, partials, ticket, loss_out);
  DONE(1);
}
extern "C" int b3d_sampler_update(const float* z, const float* eps_c, const float* eps_u, const float* w, const float* noise, float* out,
                                  float logsnr, float logsnr_next, int B, int64_t per_sample, void* stream) {
  B3D_CHECK_ARG(z && eps_c && eps_u && w && out, "sampler_update: null");
  // scalar schedule terms in fp32 like the reference's 0-d tensors (sampling.py:84-90)
  const float c = -expm1f(logsnr - logsnr_next);
  const float sa = 1.f / (1.f + expf(-logsnr)), san = 1.f / (1.f + expf(-logsnr_next));
  const float ss = 1.f / (1.f + expf(logsnr)), ssn = 1.f / (1.f + expf(logsnr_next));
  const int add_noise = (logsnr_next != 0.f) ? 1 : 0;   // sampling.py:124
  B3D_CHECK_ARG(!add_noise || noise, "sampler_update: noise required");
  sampler_update_kernel<<<grid_for((long long)B * per_sample), kT, 0, STREAM>>>(z, eps_c, eps_u, w, noise, out, sqrtf(sa), sqrtf(ss), sqrtf(san), c,
                                                                              sqrtf(ssn * c), add_noise, per_sample, (long long)B * per_sample);
  DONE(1);
}
extern "C" int b3d_sampler_update_dev(const float* z, const float* eps_c, const float* eps_u, const float* w, const float* noise, float* out,
                                      const float* sched_dev, int B, int64_t per_sample, void* stream) {
  B3D_CHECK_ARG(z && eps_c && eps_u && w && noise && out && sched_dev, "sampler_update_dev: null");
  sampler_update_dev_kernel<<<grid_for((long long)B * per_sample), kT, 0, STREAM>>>(z, eps_c, eps_u, w, noise, out, sched_dev, per_sample,
                                                                                  (long long)B * per_sample);
  DONE(1);
}
extern "C" int b3d_cond_silu(const float* c, const float* lemb, const int32_t* lemb_row, void* out, int D, int64_t rows_per_image, int E,
                             void* stream) {
  B3D_CHECK_ARG(c && lemb && lemb_row && out && D > 0 && rows_per_image > 0 && E % 8 == 0, "cond_silu: args (E %% 8)");
  B3D_CHECK_ARG((((uintptr_t)c | (uintptr_t)lemb | (uintptr_t)out) & 15) == 0, "cond_silu: 16-byte aligned tensors");
  const long long per_image = rows_per_image * E, total8 = (long long)D * per_image / 8;
  B3D_CUDA(b3d_launch_pdl(cond_silu_kernel, dim3(grid_for(total8)), dim3(256), 0, STREAM, c, lemb, (const int*)lemb_row, (bf16*)out, per_image, E, total8));
  DONE(1);
}
extern "C" int b3d_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, int step,
                        float grad_scale, void* stream) {
  B3D_CHECK_ARG(p && g && m && v && step >= 1, "adam: args");
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
  B3D_CHECK_ARG((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adam: arenas must be 16-byte aligned");
  adam_kernel<false><<<148 * 8, 256, 0, STREAM>>>(p, g, m, v, nullptr, n, lr, beta1, beta2, eps, 1.f / bc1, 1.f / bc2, grad_scale, 0.f);
  DONE(1);
}
extern "C" int b3d_adam_ema(float* p, const float* g, float* m, float* v, float* ema, int64_t n, float lr, float beta1, float beta2, float eps,
                            int step, float grad_scale, float ema_decay, void* stream) {
  B3D_CHECK_ARG(p && g && m && v && ema && step >= 1 && ema_decay >= 0.f && ema_decay <= 1.f, "adam_ema: args");
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
  B3D_CHECK_ARG((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v | (uintptr_t)ema) & 15) == 0, "adam_ema: arenas must be 16-byte aligned");
  adam_kernel<true><<<148 * 8, 256, 0, STREAM>>>(p, g, m, v, ema, n, lr, beta1, beta2, eps, 1.f / bc1, 1.f / bc2, grad_scale, ema_decay);
  DONE(1);
}

extern "C" int b3d_adam_bf16grad(float* p, const void* g_bf16, float* m, float* v, float* ema, int64_t n, float lr, float beta1, float beta2,
                                 float eps, int step, float grad_scale, float ema_decay, void* stream) {
  B3D_CHECK_ARG(p && g_bf16 && m && v && step >= 1 && ema_decay >= 0.f && ema_decay <= 1.f, "adam_bf16grad: args");
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
  B3D_CHECK_ARG((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)ema) & 15) == 0 && ((uintptr_t)g_bf16 & 7) == 0,
                "adam_bf16grad: fp32 arenas must be 16-byte aligned, the gradient 8-byte aligned");
  if (ema) adam_kernel<true, true><<<148 * 8, 256, 0, STREAM>>>(p, g_bf16, m, v, ema, n, lr, beta1, beta2, eps, 1.f / bc1, 1.f / bc2, grad_scale, ema_decay);
  else adam_kernel<false, true><<<148 * 8, 256, 0, STREAM>>>(p, g_bf16, m, v, nullptr, n, lr, beta1, beta2, eps, 1.f / bc1, 1.f / bc2, grad_scale, 0.f);
  DONE(1);
}

extern "C" int b3d_scale_colsum(const void* a, void* out, float scale, int64_t rows, int C, float* colsum, void* stream) {
  B3D_CHECK_ARG(a && out && colsum && C % 8 == 0 && C <= 2048, "scale_colsum: args (C %% 8, C <= 2048)");
  const int rows_par = 256 / (C / 8) > 0 ? 256 / (C / 8) : 1;
  long long rpb = (rows + 148 * 8 - 1) / (148 * 8);
  const int min_rows = rows_par * 4;
  if (rpb < min_rows) rpb = min_rows;
  rpb = (rpb + min_rows - 1) / min_rows * min_rows;
  B3D_CUDA(b3d_launch_pdl(scale_colsum_kernel, dim3(cdiv(rows, rpb)), dim3(256), 2048 * sizeof(float), STREAM, (const bf16*)a, (bf16*)out, scale, rows, C, (int)rpb, colsum));
  DONE(1);
}
// groups: device array of b3d_pack_group (include/b3d.h); block_group: device array, group index of every block
extern "C" int b3d_pack_weights_multi(const void* groups, int n_groups, const int32_t* block_group, int64_t total_blocks, void* stream) {
  B3D_CHECK_ARG(groups && block_group && n_groups > 0 && total_blocks > 0, "pack_weights_multi: args");
  static_assert(sizeof(PackGroup) == sizeof(b3d_pack_group), "b3d_pack_group layout");
  pack_groups_kernel<<<(unsigned)total_blocks, 256, 0, STREAM>>>((const PackGroup*)groups, block_group);
  DONE(1);
}
extern "C" int b3d_unpack_wgrads_multi(const void* jobs, int n_jobs, const int32_t* block_job, int64_t total_blocks, void* stream) {
  B3D_CHECK_ARG(jobs && n_jobs > 0 && total_blocks > 0, "unpack_wgrads_multi: args");
  unpack_multi_kernel<<<(unsigned)total_blocks, 256, 0, STREAM>>>((const PackJob*)jobs, n_jobs, block_job);
  DONE(1);
}
