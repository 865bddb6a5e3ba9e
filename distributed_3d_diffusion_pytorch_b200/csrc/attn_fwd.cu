// Fused self / cross-frame attention forward (the core of nn.MultiheadAttention inside AttnBlock, xunet.py:161,175,
// 193-211) for the token counts of the X-UNet attention levels (L <= 256, head_dim <= 256):
//
//   per work item (image n, head, 128-query tile):   S = Q K^T   ->   P = softmax(S / sqrt(d))   ->   O = P V
//
// S never leaves the SM: it is accumulated in TMEM by tcgen05.mma, read back by the softmax warps with tcgen05.ld,
// exponentiated once in fp32 registers, written as bf16 into a 128B-swizzled shared-memory tile (normalised there in a
// second sweep) that is directly the A
// operand of the second tcgen05.mma (O in TMEM), and O leaves through shared memory + TMA store.  Q, K, V arrive by TMA
// straight from the packed in_proj output [N, L, 3C] (no head split / transpose copies); the cross-frame variant reads
// K, V of image n ^ 1.  When the backward pass will follow, P (bf16, normalised) is also written out by TMA.
//
//   warp 0 (one lane) TMA producer | warp 1 (one lane) MMA issuer | warps 2..5: softmax + epilogue (thread = query row)
#include "../../include/b3d.h"
#include "b3d_common.cuh"
#include "tc_ptx.cuh"

void b3d_count_launch(int n);

namespace {
using namespace b3d_tc;

constexpr int kStages = 3;
constexpr uint32_t kStageBytes = 16384 + 32768;   // A: 128 x 64 bf16 | B: up to 256 rows x 64 (K chunk) or 4 x (64 x 64) (V chunks)
constexpr uint32_t kPBytes = 65536;               // P tile 128 x 256 bf16 = 4 chunks of 16 KB; reused as the O staging tile
constexpr int kThreads = 192;
constexpr uint32_t kTmemS = 0, kTmemO = 256;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct alignas(64) AP {
  CUtensorMap tmQ, tmK, tmV, tmO, tmP;
  int N, L, Lb, heads, d, cross, qtiles, items, write_p;
  float* lse;           // optional fp32 [N, heads, L]: log2-domain log-sum-exp of the scaled scores (the backward recomputes P from it)
  int v_one_stage;      // all V chunks of an item fit the B part of ONE pipeline stage (kch * dch * 8 KB <= 32 KB)
  float scale_log2e;
};

__global__ void __launch_bounds__(kThreads, 1) attn_fwd_kernel(const __grid_constant__ AP p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * kStages + 3];
  __shared__ uint32_t tmem_base_smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t p_base = smem_base + kStages * kStageBytes;
  auto full_bar = [&](int s) { return smem_u32(&bars[s]); };
  auto empty_bar = [&](int s) { return smem_u32(&bars[kStages + s]); };
  const uint32_t s_full = smem_u32(&bars[2 * kStages]), p_ready = smem_u32(&bars[2 * kStages + 1]), o_full = smem_u32(&bars[2 * kStages + 2]);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(s_full, 1); mbar_init(p_ready, 4); mbar_init(o_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), 512u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  b3d_pdl_sync();   // launched with programmatic stream serialization: nothing above touches global memory
  const int dch = (p.d + 63) / 64, kch = (p.L + 63) / 64;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
        const int qt = it % p.qtiles, head = (it / p.qtiles) % p.heads, n = it / (p.qtiles * p.heads);
        const int nk = n ^ p.cross;
        for (int dc = 0; dc < dch; ++dc) {            // S = Q K^T, reduction over head_dim in 64-wide chunks
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_expect_tx(full_bar(stage), 16384u + (uint32_t)p.Lb * 128u);
          const uint32_t sa = smem_base + stage * kStageBytes;
          tma_load_4d(sa, &p.tmQ, full_bar(stage), dc * 64, qt * 128, head, n);
          tma_load_4d(sa + 16384, &p.tmK, full_bar(stage), dc * 64, 0, head, nk);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (p.v_one_stage) {
          // O = P V: the whole V of this (image, head) in one stage.  With one 64-key chunk per 48 KB stage (the general form
          // below) the three stages ran out inside an item: the last V chunk and the next item's Q / K were requested only
          // after PV MMAs had freed a stage, and their load latency sat on the serial chain (ncu: the softmax warps spent
          // 39 % of their time waiting for O).
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_expect_tx(full_bar(stage), (uint32_t)(kch * dch) * 8192u);
          const uint32_t sb = smem_base + stage * kStageBytes + 16384;
          for (int kc = 0; kc < kch; ++kc)
            for (int j = 0; j < dch; ++j) tma_load_4d(sb + (kc * dch + j) * 8192, &p.tmV, full_bar(stage), j * 64, kc * 64, head, nk);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        } else {
          for (int kc = 0; kc < kch; ++kc) {            // O = P V, reduction over keys in 64-key chunks
            mbar_wait(empty_bar(stage), phase ^ 1);
            mbar_expect_tx(full_bar(stage), (uint32_t)dch * 8192u);
            const uint32_t sb = smem_base + stage * kStageBytes + 16384;
            for (int j = 0; j < dch; ++j) tma_load_4d(sb + j * 8192, &p.tmV, full_bar(stage), j * 64, kc * 64, head, nk);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0, iph = 0;
      // instruction descriptors: D fp32, A/B bf16; QK^T: both K-major, N = Lb; PV: B MN-major, N = d
      const uint32_t idesc_qk = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.Lb >> 3) << 17) | (8u << 24);
      const uint32_t idesc_pv = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(p.d >> 3) << 17) | (8u << 24);
      for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
        for (int dc = 0; dc < dch; ++dc) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * kStageBytes, sb = sa + 16384;
          const int ksteps = min(4, (p.d - dc * 64 + 15) / 16);
          for (int k = 0; k < ksteps; ++k)
            umma_bf16(tmem_base + kTmemS, umma_desc(sa + k * 32, 16, 1024), umma_desc(sb + k * 32, 16, 1024), idesc_qk, (dc | k) != 0 ? 1u : 0u);
          umma_commit(empty_bar(stage));
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(s_full);
        mbar_wait(p_ready, iph);                       // P tile written (and S fully read) by the softmax warps
        tc_fence_after();
        for (int kc = 0; kc < kch; ++kc) {
          if (!p.v_one_stage || kc == 0) {
            mbar_wait(full_bar(stage), phase);
            tc_fence_after();
          }
          const uint32_t sb = smem_base + stage * kStageBytes + 16384 + (p.v_one_stage ? (uint32_t)(kc * dch) * 8192u : 0u);
          const int ksteps = min(4, (p.L - kc * 64 + 15) / 16);
          for (int k = 0; k < ksteps; ++k)
            umma_bf16(tmem_base + kTmemO, umma_desc(p_base + kc * 16384 + k * 32, 16, 1024), umma_desc(sb + k * 2048, 8192, 1024), idesc_pv,
                      (kc | k) != 0 ? 1u : 0u);
          if (!p.v_one_stage || kc == kch - 1) {
            umma_commit(empty_bar(stage));
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
        umma_commit(o_full);
        iph ^= 1;
      }
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const bool leader = threadIdx.x == 64;
    const uint32_t swz = (uint32_t)(r & 7);
    const uint32_t lane_addr = ((uint32_t)(quad * 32) << 16);
    uint32_t iph = 0;
    for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
      const int qt = it % p.qtiles, head = (it / p.qtiles) % p.heads, n = it / (p.qtiles * p.heads);
      mbar_wait(s_full, iph);
      tc_fence_after();
      // ---- softmax over the L keys of this query row: max sweep over TMEM, exp sweep TMEM -> smem, normalisation sweep over smem
      const int nch32 = (p.L + 31) / 32;
      float m = -INFINITY;
      for (int c = 0; c < nch32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem_base + kTmemS + lane_addr + c * 32, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) if (c * 32 + i < p.L) m = fmaxf(m, __uint_as_float(v[i]));
      }
      const float mb = m * p.scale_log2e;
      if (leader) tma_store_wait_read<0>();             // previous item's O / P stores have finished reading the tile
      epi_bar_sync();
      // second sweep: e = 2^(s*scale*log2e - mb) ONCE per element (one MUFU.EX2; exp2f()'s range handling and a second
      // exponential pass made this loop the kernel's bottleneck), written unnormalised as bf16 into the P tile, row sum in fp32
      float l = 0.f;
      for (int c = 0; c < 2 * kch; ++c) {               // all 64-key chunks the PV MMAs will read (keys >= L are zero)
        uint32_t v[32];
        if (c < nch32) tmem_ld32(tmem_base + kTmemS + lane_addr + c * 32, v);
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          f[i] = (c < nch32 && c * 32 + i < p.L) ? ex2_approx(fmaf(__uint_as_float(v[i]), p.scale_log2e, -mb)) : 0.f;
          l += f[i];
        }
        const uint32_t row = p_base + (uint32_t)(c >> 1) * 16384u + (uint32_t)r * 128u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 q;
          q.x = pack_bf16x2(f[8 * g], f[8 * g + 1]); q.y = pack_bf16x2(f[8 * g + 2], f[8 * g + 3]);
          q.z = pack_bf16x2(f[8 * g + 4], f[8 * g + 5]); q.w = pack_bf16x2(f[8 * g + 6], f[8 * g + 7]);
          sts128(row + ((((uint32_t)((c & 1) * 4 + g)) ^ swz) << 4), q);
        }
      }
      // row statistic for a backward pass that recomputes P instead of reading it back: P_ij = 2^(s_ij*scale*log2e - lse_i)
      if (p.lse != nullptr && qt * 128 + r < p.L) p.lse[((long long)n * p.heads + head) * p.L + qt * 128 + r] = mb + __log2f(l);
      // third sweep, over shared memory only: normalise the row in place (each thread re-reads what it wrote)
      const float inv_l = 1.f / l;
      for (int c = 0; c < kch; ++c) {                    // one 64-key chunk (a 128-byte row segment) per iteration: 8 loads in flight
        const uint32_t row = p_base + (uint32_t)c * 16384u + (uint32_t)r * 128u;
        uint4 q[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) q[u] = lds128(row + ((((uint32_t)u) ^ swz) << 4));      // (u ^ swz: conflict-free across the rows of a quarter warp)
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          uint32_t* w = reinterpret_cast<uint32_t*>(&q[u]);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            w[i] = pack_bf16x2(__uint_as_float(w[i] << 16) * inv_l, __uint_as_float(w[i] & 0xffff0000u) * inv_l);
          sts128(row + ((((uint32_t)u) ^ swz) << 4), q[u]);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready);
      if (p.write_p) {                                  // probabilities for the backward pass
        epi_bar_sync();
        if (leader) {
          for (int kc = 0; kc < kch; ++kc) tma_store_4d(&p.tmP, p_base + kc * 16384, kc * 64, qt * 128, head, n);
          tma_store_commit();
        }
      }
      // ---- O: TMEM -> bf16 -> swizzled staging (the P tile, free once the PV MMAs have completed) -> TMA store
      mbar_wait(o_full, iph);
      tc_fence_after();
      if (leader) tma_store_wait_read<0>();
      epi_bar_sync();
      for (int c = 0; c < (p.d + 31) / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem_base + kTmemO + lane_addr + c * 32, v);
        const uint32_t row = p_base + (uint32_t)(c >> 1) * 16384u + (uint32_t)r * 128u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 q;
          q.x = pack_bf16x2(__uint_as_float(v[8 * g]), __uint_as_float(v[8 * g + 1]));
          q.y = pack_bf16x2(__uint_as_float(v[8 * g + 2]), __uint_as_float(v[8 * g + 3]));
          q.z = pack_bf16x2(__uint_as_float(v[8 * g + 4]), __uint_as_float(v[8 * g + 5]));
          q.w = pack_bf16x2(__uint_as_float(v[8 * g + 6]), __uint_as_float(v[8 * g + 7]));
          sts128(row + ((((uint32_t)((c & 1) * 4 + g)) ^ swz) << 4), q);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      epi_bar_sync();
      if (leader) {
        for (int j = 0; j < dch; ++j) tma_store_4d(&p.tmO, p_base + j * 16384, j * 64, qt * 128, head, n);
        tma_store_commit();
      }
      iph ^= 1;
    }
    if (leader) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace

extern "C" int b3d_attn_fwd(const void* qkv, int N, int L, int heads, int d, int cross, float scale, void* out, void* probs, int64_t ld_probs,
                            float* lse, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B3D_CHECK_ARG(qkv && out && N > 0 && heads > 0, "attn_fwd: null / empty");
  B3D_CHECK_ARG(L >= 1 && L <= 256, "attn_fwd: L = %d tokens (fused path covers L <= 256; longer sequences use the b3d_gemm path)", L);
  B3D_CHECK_ARG(d % 16 == 0 && d >= 16 && d <= 256, "attn_fwd: head_dim %d must be a multiple of 16 in [16,256]", d);
  B3D_CHECK_ARG(!cross || N % 2 == 0, "attn_fwd: cross-frame attention needs an even number of images");
  B3D_CHECK_ARG(probs == nullptr || (ld_probs % 8 == 0 && ld_probs >= L), "attn_fwd: probs row stride must be a multiple of 8 and >= L");
  const int C = heads * d;
  AP p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.L = L; p.heads = heads; p.d = d; p.cross = cross ? 1 : 0;
  p.Lb = (L + 15) / 16 * 16;
  p.qtiles = (L + 127) / 128;
  p.items = N * heads * p.qtiles;
  p.write_p = probs != nullptr;
  p.lse = lse;
  p.v_one_stage = ((L + 63) / 64) * ((d + 63) / 64) * 8192 <= 32768 ? 1 : 0;
  p.scale_log2e = scale * 1.4426950408889634f;
  const uint32_t ones[4] = {1, 1, 1, 1};
  b3d_view4 v;
  v.dim[0] = d; v.dim[1] = L; v.dim[2] = heads; v.dim[3] = N;
  v.stride[0] = 3 * C; v.stride[1] = d; v.stride[2] = (int64_t)L * 3 * C;
  int rc;
  const uint32_t boxQ[4] = {64, 128, 1, 1}, boxK[4] = {64, (uint32_t)p.Lb, 1, 1}, boxV[4] = {64, 64, 1, 1};
  v.ptr = qkv;
  rc = make_map(&p.tmQ, v, boxQ, ones); if (rc) return rc;
  v.ptr = reinterpret_cast<const bf16*>(qkv) + C;
  rc = make_map(&p.tmK, v, boxK, ones); if (rc) return rc;
  v.ptr = reinterpret_cast<const bf16*>(qkv) + 2 * C;
  rc = make_map(&p.tmV, v, boxV, ones); if (rc) return rc;
  v.ptr = out; v.stride[0] = C; v.stride[2] = (int64_t)L * C;
  rc = make_map(&p.tmO, v, boxQ, ones); if (rc) return rc;
  if (probs) {
    b3d_view4 pv;
    pv.ptr = probs; pv.dim[0] = L; pv.dim[1] = L; pv.dim[2] = heads; pv.dim[3] = N;
    pv.stride[0] = ld_probs; pv.stride[1] = (int64_t)L * ld_probs; pv.stride[2] = (int64_t)heads * L * ld_probs;
    rc = make_map(&p.tmP, pv, boxQ, ones); if (rc) return rc;
  }
  const size_t smem = (size_t)kStages * kStageBytes + kPBytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    B3D_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int grid = p.items < b3d_num_sms() ? p.items : b3d_num_sms();
  B3D_CUDA(b3d_launch_pdl(attn_fwd_kernel, dim3(grid), dim3(kThreads), smem, stream, p));
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}
