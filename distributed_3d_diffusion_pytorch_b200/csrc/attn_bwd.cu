// Fused self / cross-frame attention backward (autograd of the core of nn.MultiheadAttention inside AttnBlock,
// xunet.py:161,175,193-211) for the token counts of the CFG-64 attention levels:
//
//   per work item (image n, head), for each 128-query tile t:
//       dP = dO_t V^T                     (tcgen05, accumulator in TMEM)
//       dV += P_t^T dO_t                  (TMEM, accumulated over the tiles)
//       dS = P_t o (dP - D),  D_i = dO_i . O_i   (four warps: TMEM + the P, dO, O tiles in shared memory -> bf16 dS IN PLACE of P)
//       dQ_t = scale dS K                 (TMEM, aliases the dP columns once dS is out)
//       dK += scale dS^T Q_t              (TMEM, accumulated over the tiles)
//
// Nothing of size L x L leaves the SM: the unfused path wrote fp32 dP and bf16 dS to HBM and launched five GEMMs and a
// softmax-backward kernel per block.  K and V stay in shared memory for the whole item; P (saved by the forward), Q_t, O_t and
// dO_t arrive by TMA straight from the packed [N, L, 3C] / [N, L, C] activations; the cross-frame variant reads K, V of
// image n ^ 1 and writes dK, dV there.  The same 128B-swizzled tiles serve as K-major and as MN-major UMMA operands
// (P_t^T and dS^T are the MN-major views of the row-major tile), so no transposed copy exists.
//
//   warp 0 (one lane) TMA producer | warp 1 (one lane) MMA issuer | warps 2..5: dS + epilogues (thread = row)
#include "../../include/b3d.h"
#include "b3d_common.cuh"
#include "tc_ptx.cuh"

void b3d_count_launch(int n);

namespace {
using namespace b3d_tc;

constexpr int kThreads = 192;

struct alignas(64) BP {
  CUtensorMap tmQ, tmK, tmV, tmDO, tmO, tmP;
  bf16* dqkv;
  int N, L, Lb2, heads, d, C, cross, qtiles, items, dch, kch2;
  uint32_t off_k, off_v, off_q, off_do, off_o, off_p;    // shared-memory offsets of the operand tiles
  uint32_t col_k, col_v;                           // TMEM columns of the dK / dV accumulators (dP and dQ start at 0)
  float scale;
  int nbuf;              // 1 or 2 sets of per-tile operand buffers (Q, dO, O): with 2 the next tile's loads overlap this tile's work
  uint32_t tile_bytes;   // bytes of one set
  const float* lse;      // != nullptr: P is recomputed from Q, K and these row statistics (no P tensor is read)
  float scale_log2e;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float dot8(const uint4& a, const uint4& b) {
  const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 u = __bfloat1622float2(x[i]), v = __bfloat1622float2(y[i]); s = fmaf(u.x, v.x, s); s = fmaf(u.y, v.y, s); }
  return s;
}

__global__ void __launch_bounds__(kThreads, 1) attn_bwd_kernel(const __grid_constant__ BP p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[10];
  __shared__ uint32_t tmem_base_smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // t_full[b] / post_full[b]: per operand-buffer set b (a tile T uses set T % nbuf, phase (T / nbuf) & 1); the others are per tile
  const uint32_t kv_full = smem_u32(&bars[0]), pre_full = smem_u32(&bars[2]), ds_ready = smem_u32(&bars[3]),
                 acc_read = smem_u32(&bars[5]), s_full = smem_u32(&bars[6]), p_ready = smem_u32(&bars[7]);
  auto t_full = [&](uint32_t T) { return smem_u32(&bars[T % p.nbuf == 0 ? 1 : 8]); };
  auto post_full = [&](uint32_t T) { return smem_u32(&bars[T % p.nbuf == 0 ? 4 : 9]); };
  auto tpar = [&](uint32_t T) { return (T / p.nbuf) & 1u; };
  const bool recompute = p.lse != nullptr;
  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1); mbar_init(smem_u32(&bars[1]), 1); mbar_init(smem_u32(&bars[8]), 1); mbar_init(pre_full, 1); mbar_init(ds_ready, 4);
    mbar_init(smem_u32(&bars[4]), 1); mbar_init(smem_u32(&bars[9]), 1); mbar_init(acc_read, 4);
    mbar_init(s_full, 1); mbar_init(p_ready, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), 512u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  b3d_pdl_sync();   // nothing above touches global memory
  const uint32_t sK = smem_base + p.off_k, sV = smem_base + p.off_v, sP = smem_base + p.off_p;
  const uint32_t sQ0 = smem_base + p.off_q, sDO0 = smem_base + p.off_do, sO0 = smem_base + p.off_o;
  const uint32_t kv_chunk = (uint32_t)p.Lb2 * 128u;      // bytes of one 64-wide head_dim chunk of K or V (all keys)
  const int halves = p.Lb2 / 128;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t tiles = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
        const int head = it % p.heads, n = it / p.heads, nk = n ^ p.cross;
        for (int t = 0; t < p.qtiles; ++t, ++tiles) {
          // this tile's operand set is free once the MMAs of the tile that used it last have finished reading shared memory;
          // K and V of a new item additionally wait for the previous item's last tile
          if (tiles >= (uint32_t)p.nbuf) mbar_wait(post_full(tiles - p.nbuf), tpar(tiles - p.nbuf));
          const uint32_t bo = (tiles % p.nbuf) * p.tile_bytes, sQ = sQ0 + bo, sDO = sDO0 + bo, sO = sO0 + bo;
          const uint32_t tf = t_full(tiles);
          mbar_expect_tx(tf, (uint32_t)(3 * p.dch + (recompute ? 0 : p.kch2)) * 16384u);
          for (int j = 0; j < p.dch; ++j) {
            tma_load_4d(sQ + j * 16384, &p.tmQ, tf, j * 64, t * 128, head, n);
            tma_load_4d(sDO + j * 16384, &p.tmDO, tf, j * 64, t * 128, head, n);
            tma_load_4d(sO + j * 16384, &p.tmO, tf, j * 64, t * 128, head, n);
          }
          if (!recompute)
            for (int c = 0; c < p.kch2; ++c) tma_load_4d(sP + c * 16384, &p.tmP, tf, c * 64, t * 128, head, n);
          if (t == 0) {      // (after the tile's own loads: with two operand sets those are already in flight while this waits)
            if (tiles > 0 && p.nbuf > 1) mbar_wait(post_full(tiles - 1), tpar(tiles - 1));
            mbar_expect_tx(kv_full, 2u * (uint32_t)p.dch * kv_chunk);
            for (int j = 0; j < p.dch; ++j) {
              tma_load_4d(sK + j * kv_chunk, &p.tmK, kv_full, j * 64, 0, head, nk);
              tma_load_4d(sV + j * kv_chunk, &p.tmV, kv_full, j * 64, 0, head, nk);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptors: D fp32, A/B bf16, M = 128; bit 15 / 16 = A / B MN-major
      const uint32_t ibase = (1u << 4) | (1u << 7) | (1u << 10) | (8u << 24);
      const uint32_t idesc_dp = ibase | ((uint32_t)(p.Lb2 >> 3) << 17);
      const uint32_t idesc_dq = ibase | (1u << 16) | ((uint32_t)(p.d >> 3) << 17);
      const uint32_t idesc_tt = ibase | (1u << 15) | (1u << 16) | ((uint32_t)(p.d >> 3) << 17);
      uint32_t tiles = 0, items = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++items) {
        mbar_wait(kv_full, items & 1);
        for (int t = 0; t < p.qtiles; ++t, ++tiles) {
          mbar_wait(t_full(tiles), tpar(tiles));
          if (tiles > 0) mbar_wait(acc_read, (tiles - 1) & 1);      // dQ (and, across items, dK / dV) drained from TMEM
          tc_fence_after();
          const uint32_t bo = (tiles % p.nbuf) * p.tile_bytes, sQ = sQ0 + bo, sDO = sDO0 + bo;
          if (recompute) {
            // S = Q_t K^T into the dP columns; the four warps turn it into the bf16 P tile in shared memory, then dP may overwrite it
            for (int j = 0; j < p.dch; ++j) {
              const int ksteps = min(4, (p.d - j * 64 + 15) / 16);
              for (int k = 0; k < ksteps; ++k)
                umma_bf16(tmem_base, umma_desc(sQ + j * 16384 + k * 32, 16, 1024), umma_desc(sK + j * kv_chunk + k * 32, 16, 1024), idesc_dp,
                          (j | k) != 0 ? 1u : 0u);
            }
            umma_commit(s_full);
            mbar_wait(p_ready, tiles & 1);
            tc_fence_after();
          }
          // dP = dO_t V^T : M = 128 queries, N = Lb2 keys, K = head_dim
          for (int j = 0; j < p.dch; ++j) {
            const int ksteps = min(4, (p.d - j * 64 + 15) / 16);
            for (int k = 0; k < ksteps; ++k)
              umma_bf16(tmem_base, umma_desc(sDO + j * 16384 + k * 32, 16, 1024), umma_desc(sV + j * kv_chunk + k * 32, 16, 1024), idesc_dp,
                        (j | k) != 0 ? 1u : 0u);
          }
          // dV[h] += P_t[:, h]^T dO_t : M = 128 keys, N = head_dim, K = 128 queries
          for (int h = 0; h < halves; ++h)
            for (int kq = 0; kq < 8; ++kq)
              umma_bf16(tmem_base + p.col_v + h * p.d, umma_desc(sP + (2 * h) * 16384 + kq * 2048, 16384, 1024),
                        umma_desc(sDO + kq * 2048, 16384, 1024), idesc_tt, (t | kq) != 0 ? 1u : 0u);
          umma_commit(pre_full);
          mbar_wait(ds_ready, tiles & 1);
          tc_fence_after();
          // dQ_t = dS K : M = 128 queries, N = head_dim, K = Lb2 keys   (into the dP columns: dP has been consumed)
          for (int c = 0; c < p.kch2; ++c)
            for (int k = 0; k < 4; ++k)
              umma_bf16(tmem_base, umma_desc(sP + c * 16384 + k * 32, 16, 1024), umma_desc(sK + (c * 64 + k * 16) * 128, kv_chunk, 1024), idesc_dq,
                        (c | k) != 0 ? 1u : 0u);
          // dK[h] += dS[:, h]^T Q_t
          for (int h = 0; h < halves; ++h)
            for (int kq = 0; kq < 8; ++kq)
              umma_bf16(tmem_base + p.col_k + h * p.d, umma_desc(sP + (2 * h) * 16384 + kq * 2048, 16384, 1024),
                        umma_desc(sQ + kq * 2048, 16384, 1024), idesc_tt, (t | kq) != 0 ? 1u : 0u);
          umma_commit(post_full(tiles));
        }
      }
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t swz = (uint32_t)(r & 7);
    const uint32_t lane_addr = ((uint32_t)(quad * 32) << 16);
    const int C3 = 3 * p.C;
    uint32_t tiles = 0;
    for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
      const int head = it % p.heads, n = it / p.heads, nk = n ^ p.cross;
      for (int t = 0; t < p.qtiles; ++t, ++tiles) {
        const int q = t * 128 + r;
        const bool valid = q < p.L;
        // D_i = dO_i . O_i (fp32) from the two tiles in shared memory (rows of 128 B, same swizzle, so 16-byte units pair up);
        // reading them from global memory here put a full load latency on the serial chain of every tile
        mbar_wait(t_full(tiles), tpar(tiles));
        const uint32_t bo = (tiles % p.nbuf) * p.tile_bytes, sDO = sDO0 + bo, sO = sO0 + bo;
        float D = 0.f;
        for (int j = 0; j < p.dch; ++j) {
          const uint32_t ro = sO + (uint32_t)j * 16384u + (uint32_t)r * 128u, rd = sDO + (uint32_t)j * 16384u + (uint32_t)r * 128u;
#pragma unroll
          for (int u = 0; u < 8; ++u) D += dot8(lds128(ro + ((((uint32_t)u) ^ swz) << 4)), lds128(rd + ((((uint32_t)u) ^ swz) << 4)));   // conflict-free order
        }
        if (recompute) {
          // P_ij = 2^(s_ij * scale * log2e - lse_i) as bf16 into the P tile (the layout a TMA load of the saved P would give);
          // rows past L and keys past L are zero, like the zero fill of that load
          const float l2 = valid ? __ldg(p.lse + ((long long)n * p.heads + head) * p.L + q) : 0.f;
          mbar_wait(s_full, tiles & 1);
          tc_fence_after();
          for (int c = 0; c < p.Lb2 / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(tmem_base + lane_addr + c * 32, v);
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i)
              f[i] = (valid && c * 32 + i < p.L) ? ex2_approx(fmaf(__uint_as_float(v[i]), p.scale_log2e, -l2)) : 0.f;
            const uint32_t row = sP + (uint32_t)(c >> 1) * 16384u + (uint32_t)r * 128u;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 o4;
              o4.x = pack_bf16x2(f[8 * g], f[8 * g + 1]); o4.y = pack_bf16x2(f[8 * g + 2], f[8 * g + 3]);
              o4.z = pack_bf16x2(f[8 * g + 4], f[8 * g + 5]); o4.w = pack_bf16x2(f[8 * g + 6], f[8 * g + 7]);
              sts128(row + ((((uint32_t)((c & 1) * 4 + g)) ^ swz) << 4), o4);
            }
          }
          tc_fence_before();
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(p_ready);
        }
        mbar_wait(pre_full, tiles & 1);
        tc_fence_after();
        // dS = P o (dP - D), written over P (the dV MMAs that read P have completed: pre_full)
        for (int c = 0; c < p.Lb2 / 32; ++c) {
          uint32_t v[32];
          tmem_ld32(tmem_base + lane_addr + c * 32, v);
          const uint32_t row = sP + (uint32_t)(c >> 1) * 16384u + (uint32_t)r * 128u;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t a = row + ((((uint32_t)((c & 1) * 4 + g)) ^ swz) << 4);
            const uint4 pq = lds128(a);
            const __nv_bfloat162* ph = reinterpret_cast<const __nv_bfloat162*>(&pq);
            float f[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 pp = __bfloat1622float2(ph[i]);
              f[2 * i] = pp.x * (__uint_as_float(v[8 * g + 2 * i]) - D);
              f[2 * i + 1] = pp.y * (__uint_as_float(v[8 * g + 2 * i + 1]) - D);
            }
            uint4 o4;
            o4.x = pack_bf16x2(f[0], f[1]); o4.y = pack_bf16x2(f[2], f[3]); o4.z = pack_bf16x2(f[4], f[5]); o4.w = pack_bf16x2(f[6], f[7]);
            sts128(a, o4);
          }
        }
        tc_fence_before();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(ds_ready);
        mbar_wait(post_full(tiles), tpar(tiles));
        tc_fence_after();
        // dQ rows of this tile
        for (int c = 0; c < p.d / 32 + ((p.d & 31) ? 1 : 0); ++c) {
          uint32_t v[32];
          tmem_ld32(tmem_base + lane_addr + c * 32, v);
          if (valid) {
            bf16* dst = p.dqkv + ((long long)n * p.L + q) * C3 + head * p.d + c * 32;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (c * 32 + g * 8 >= p.d) break;
              uint4 o4;
              o4.x = pack_bf16x2(__uint_as_float(v[8 * g]) * p.scale, __uint_as_float(v[8 * g + 1]) * p.scale);
              o4.y = pack_bf16x2(__uint_as_float(v[8 * g + 2]) * p.scale, __uint_as_float(v[8 * g + 3]) * p.scale);
              o4.z = pack_bf16x2(__uint_as_float(v[8 * g + 4]) * p.scale, __uint_as_float(v[8 * g + 5]) * p.scale);
              o4.w = pack_bf16x2(__uint_as_float(v[8 * g + 6]) * p.scale, __uint_as_float(v[8 * g + 7]) * p.scale);
              *reinterpret_cast<uint4*>(dst + g * 8) = o4;
            }
          }
        }
        if (t == p.qtiles - 1) {        // dK, dV rows (keys h*128 + r) of the key/value image
          for (int h = 0; h < halves; ++h) {
            const int key = h * 128 + r;
            for (int which = 0; which < 2; ++which) {
              const uint32_t col = (which == 0 ? p.col_k : p.col_v) + (uint32_t)(h * p.d);
              const float sc = which == 0 ? p.scale : 1.f;
              for (int c = 0; c < p.d / 32 + ((p.d & 31) ? 1 : 0); ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + lane_addr + col + c * 32, v);
                if (key < p.L) {
                  bf16* dst = p.dqkv + ((long long)nk * p.L + key) * C3 + (1 + which) * p.C + head * p.d + c * 32;
#pragma unroll
                  for (int g = 0; g < 4; ++g) {
                    if (c * 32 + g * 8 >= p.d) break;
                    uint4 o4;
                    o4.x = pack_bf16x2(__uint_as_float(v[8 * g]) * sc, __uint_as_float(v[8 * g + 1]) * sc);
                    o4.y = pack_bf16x2(__uint_as_float(v[8 * g + 2]) * sc, __uint_as_float(v[8 * g + 3]) * sc);
                    o4.z = pack_bf16x2(__uint_as_float(v[8 * g + 4]) * sc, __uint_as_float(v[8 * g + 5]) * sc);
                    o4.w = pack_bf16x2(__uint_as_float(v[8 * g + 6]) * sc, __uint_as_float(v[8 * g + 7]) * sc);
                    *reinterpret_cast<uint4*>(dst + g * 8) = o4;
                  }
                }
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_read);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace

// 1 when b3d_attn_bwd covers the shape (the host falls back to the b3d_gemm path otherwise)
extern "C" int b3d_attn_bwd_supported(int L, int d) {
  if (L < 1 || L > 256 || d < 16 || d > 128 || d % 16 != 0) return 0;
  const int Lb2 = (L + 127) / 128 * 128, halves = Lb2 / 128, dch = (d + 63) / 64;
  if (Lb2 + 2 * halves * d > 512 || d > Lb2) return 0;
  const size_t smem = (size_t)2 * dch * Lb2 * 128 + (size_t)3 * dch * 16384 + (size_t)(Lb2 / 64) * 16384 + 1024;
  return smem <= 220 * 1024 ? 1 : 0;
}

extern "C" int b3d_attn_bwd(const void* qkv, const void* probs, int64_t ld_probs, const float* lse, const void* o, const void* dout, void* dqkv,
                            int N, int L, int heads, int d, int cross, float scale, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B3D_CHECK_ARG(qkv && o && dout && dqkv && N > 0 && heads > 0, "attn_bwd: null / empty");
  B3D_CHECK_ARG((probs != nullptr) != (lse != nullptr), "attn_bwd: pass either the saved probabilities or the saved row statistics (lse)");
  B3D_CHECK_ARG(b3d_attn_bwd_supported(L, d), "attn_bwd: L = %d, head_dim = %d is outside the fused path (use the b3d_gemm path)", L, d);
  B3D_CHECK_ARG(!cross || N % 2 == 0, "attn_bwd: cross-frame attention needs an even number of images");
  B3D_CHECK_ARG(probs == nullptr || (ld_probs % 8 == 0 && ld_probs >= L), "attn_bwd: probs row stride must be a multiple of 8 and >= L");
  const int C = heads * d;
  B3D_CHECK_ARG(((((uintptr_t)o) | ((uintptr_t)dout) | ((uintptr_t)dqkv)) & 15) == 0, "attn_bwd: o / dout / dqkv must be 16-byte aligned");
  BP p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.L = L; p.heads = heads; p.d = d; p.C = C; p.cross = cross ? 1 : 0; p.scale = scale;
  p.lse = lse; p.scale_log2e = scale * 1.4426950408889634f;
  p.Lb2 = (L + 127) / 128 * 128;
  p.qtiles = (L + 127) / 128;
  p.items = N * heads;
  p.dch = (d + 63) / 64; p.kch2 = p.Lb2 / 64;
  p.dqkv = reinterpret_cast<bf16*>(dqkv);
  const uint32_t kv_bytes = (uint32_t)p.dch * (uint32_t)p.Lb2 * 128u, qd_bytes = (uint32_t)p.dch * 16384u;
  // two sets of per-tile operand buffers when P is recomputed (no P load to wait for) and they fit beside K, V and the P / dS tile
  p.tile_bytes = 3 * qd_bytes;
  p.nbuf = (lse != nullptr && (size_t)2 * kv_bytes + 2 * (size_t)p.tile_bytes + (size_t)p.kch2 * 16384 + 1024 <= 227 * 1024) ? 2 : 1;
  p.off_k = 0; p.off_v = kv_bytes; p.off_q = 2 * kv_bytes; p.off_do = p.off_q + qd_bytes; p.off_o = p.off_do + qd_bytes;
  p.off_p = p.off_q + (uint32_t)p.nbuf * p.tile_bytes;
  p.col_k = (uint32_t)p.Lb2; p.col_v = (uint32_t)(p.Lb2 + (p.Lb2 / 128) * d);
  const uint32_t ones[4] = {1, 1, 1, 1};
  b3d_view4 v;
  v.dim[0] = d; v.dim[1] = L; v.dim[2] = heads; v.dim[3] = N;
  v.stride[0] = 3 * C; v.stride[1] = d; v.stride[2] = (int64_t)L * 3 * C;
  int rc;
  const uint32_t boxQ[4] = {64, 128, 1, 1}, boxK[4] = {64, (uint32_t)p.Lb2, 1, 1};
  v.ptr = qkv;
  rc = make_map(&p.tmQ, v, boxQ, ones); if (rc) return rc;
  v.ptr = reinterpret_cast<const bf16*>(qkv) + C;
  rc = make_map(&p.tmK, v, boxK, ones); if (rc) return rc;
  v.ptr = reinterpret_cast<const bf16*>(qkv) + 2 * C;
  rc = make_map(&p.tmV, v, boxK, ones); if (rc) return rc;
  v.ptr = dout; v.stride[0] = C; v.stride[2] = (int64_t)L * C;
  rc = make_map(&p.tmDO, v, boxQ, ones); if (rc) return rc;
  v.ptr = o;
  rc = make_map(&p.tmO, v, boxQ, ones); if (rc) return rc;
  if (probs != nullptr) {
    b3d_view4 pv;
    pv.ptr = probs; pv.dim[0] = L; pv.dim[1] = L; pv.dim[2] = heads; pv.dim[3] = N;
    pv.stride[0] = ld_probs; pv.stride[1] = (int64_t)L * ld_probs; pv.stride[2] = (int64_t)heads * L * ld_probs;
    rc = make_map(&p.tmP, pv, boxQ, ones); if (rc) return rc;
  }
  const size_t smem = (size_t)p.off_p + (size_t)p.kch2 * 16384 + 1024;
  static size_t attr_smem = 0;
  if (smem > attr_smem) {
    B3D_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem = smem;
  }
  const int grid = p.items < b3d_num_sms() ? p.items : b3d_num_sms();
  B3D_CUDA(b3d_launch_pdl(attn_bwd_kernel, dim3(grid), dim3(kThreads), smem, stream, p));
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}
