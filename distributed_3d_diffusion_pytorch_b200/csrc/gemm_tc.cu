// b3d tensor-core path: ONE warp-specialised, persistent tcgen05 kernel for every GEMM-shaped op of
// the X-UNet (3x3/1x1 convolutions as implicit GEMM, FiLM / linear layers, attention matmuls),
// forward, data-gradient and weight-gradient.
//
//   warp 0  TMA producer : 4-D tiled tensor maps, SWIZZLE_128B, zero fill out of bounds (3x3 halo = shifted box
//                          coordinates, strided conv = traversal strides; no im2col buffer ever exists)
//   warp 1  MMA issuer   : tcgen05.mma kind::f16, M = 128 (cta_group::1) or 256 across a CTA pair (cta_group::2),
//                          N = 16..256, K = 16, bf16 x bf16 -> fp32 accumulators in TMEM (double buffered)
//   warps 2..5 epilogue  : tcgen05.ld 32x32b.x32 -> registers -> bias / per-image vector / residual / scale / SiLU ->
//                          swizzled smem tile -> TMA store (bf16, fp32) or TMA reduce-add (split-K weight gradients)
//   Warps 0 and 1 run warp-uniform loops; one elected lane issues the TMA / MMA instructions.
//
// Tile shapes: 128 x BN; 256 x BN with two accumulators per B tile (BN <= 128); 256 x 256 over a CTA pair; weight-gradient
// tiles spanning several taps of a narrow layer.  b3d_gemm() picks per launch (see the comments there).
//
// Operands may be K-major (K = channels, contiguous) or MN-major (K = pixels); see include/b3d.h.
#include <cuda.h>
#include <stdarg.h>

#include <cstdlib>
#include "../../include/b3d.h"
#include "b3d_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2..5 and 6..9: two epilogue groups
constexpr int kMaxSeg = 27;
constexpr uint32_t kABytes = BM * BK * 2;  // 16 KB in both majors
constexpr int kMaxDynSmem = 227 * 1024 - 3072;  // 227 KB per CTA minus the static barriers and the two bias slices

struct Seg {
  int32_t src, dx, dy, ncb, bk0;
  int32_t klast;   // K = 16 steps of the LAST 64-channel block of the segment that hold real channels (1..4): the zero-filled rest is not multiplied
};

struct alignas(64) KP {
  CUtensorMap tmA[3];
  CUtensorMap tmB;
  CUtensorMap tmO, tmO2, tmR;   // epilogue: TMA store of the bf16 output(s), TMA load of the residual
  Seg seg[kMaxSeg];
  int nseg;
  int a_major, b_major, batched, b_xor, o_xor;
  int BN, b_chunks;
  int tiles_m, tiles_n, tiles_b, tiles_z, tiles_total;
  int bw, bh, bnn, tiles_x, tiles_y;
  int c2;                            // CTA pair (cta_group::2): the tile's two 128-row halves live in the two CTAs of a cluster
  int m2, m2_dim, tbw, tbh, tbnn;   // two 128-row accumulators per tile sharing one B tile (see b3d_gemm); tile-origin strides
  uint32_t a_bytes, acc_half;       // bytes of the A tile of a stage; TMEM column distance of the second accumulator
  int W, H, N;
  int stride, pad;
  int kbw, kbh, kbn, kblk_x, kblk_y, kblk_total;
  int split_k, b_taps;
  int tpg, tap_groups, cpt;          // weight gradients: taps per tile (their Ci-wide slabs side by side in one N tile), tile groups, 64-chunks per tap
  int kchan_blocks;
  int m_rows, n_cols;
  int num_stages;
  uint32_t a_tx, b_tx, b_bytes, stage_bytes;
  uint32_t idesc;
  int tmem_cols, acc_stride;
  void* out;
  void* out2;
  const float* bias;
  const float* rowvec;
  const bf16* resid;
  long long rowvec_ld;
  long long os_x, os_y, os_n, os_tap;
  float alpha, out_scale;
  int act, out_fp32, out_atomic;
  float* gn_sums;       // optional: GroupNorm (sum, sumsq) of the stored bf16 output, fp32 [images, 32, 2], accumulated
  int gn_cpg, gn_rpi;   // channels per group (4 / 8 / 16 / 32); rows per image of the dense form (0: conv form)
  // Halo-patch form of a stride-1 3x3 convolution: per 64-channel block ONE (8+2) x (rows+2)-pixel patch of the input is loaded
  // (rows of 10 pixels at a 16-pixel pitch) and the nine taps' MMAs read it through shifted A descriptors -- 6x less
  // A traffic from L2 than nine shifted tile loads.  Separate rings for the patches (a_stages) and the weight tiles.
  struct HSrc { int32_t src, taps, ncb, bk0, klast; } hsrc[3];
  int halo, n_hsrc, h_accs, h_rows, a_stages, b_stages;
  uint32_t h_stage_bytes, b_ring_off, h_pitch;   // h_pitch: bytes between patch rows (1280 = packed, one TMA load per patch; 2048 = one load per row)
  int resid2;           // two residual tiles per epilogue group (the next round's residual loads while this round computes)
  int epi_tma;          // bf16 outputs leave through shared memory + TMA store (coalesced, no LSU global traffic)
  uint32_t epi_off;     // byte offset of the epilogue staging buffers inside dynamic smem
};

using namespace b3d_tc;

struct TileCoord {
  int tn, tm, tb, tz;
  int x0, y0, n0;  // a_major = 0: first output pixel; a_major = 1: batch (y0,n0)
  int m0;          // a_major = 1: first channel row
  int tap, split;
};

__device__ __forceinline__ TileCoord decode_tile(const KP& p, int t) {
  TileCoord c;
  c.tn = t % p.tiles_n; t /= p.tiles_n;
  c.tm = t % p.tiles_m; t /= p.tiles_m;
  c.tb = t % p.tiles_b; c.tz = t / p.tiles_b;
  c.tap = c.tz % p.tap_groups;       // tap, or first-tap group when several taps share a tile (tpg > 1)
  c.split = c.tz / p.tap_groups;
  c.m0 = 0; c.x0 = 0; c.y0 = 0; c.n0 = 0;
  if (p.a_major == 0) {
    int tx = c.tm % p.tiles_x;
    int r = c.tm / p.tiles_x;
    c.x0 = tx * p.tbw;
    c.y0 = (r % p.tiles_y) * p.tbh;
    c.n0 = (r / p.tiles_y) * p.tbnn;
  } else {
    c.m0 = c.tm * BM;
    if (p.batched) { c.y0 = c.tb % p.H; c.n0 = c.tb / p.H; }
  }
  return c;
}

// number of K blocks of a tile and (for pixel-K without batching) the first pixel block
__device__ __forceinline__ void k_range(const KP& p, const TileCoord& c, int& kb0, int& nkb) {
  if (p.a_major == 0) {
    kb0 = 0;
    if (p.b_major == 0) { nkb = 0; for (int s = 0; s < p.nseg; ++s) nkb += p.seg[s].ncb; }
    else nkb = p.kchan_blocks;
  } else if (p.batched) {
    kb0 = 0; nkb = p.kblk_x;
  } else {
    long long tot = p.kblk_total;
    kb0 = (int)(tot * c.split / p.split_k);
    nkb = (int)(tot * (c.split + 1) / p.split_k) - kb0;
  }
}

// kPair = true: the CTA-pair (cta_group::2) instantiation -- it must be launched as clusters of two (a kernel containing
// cta_group::2 instructions is rejected at launch without them), so the single-CTA path is a separate instantiation.
template <bool kPair>
__global__ void __launch_bounds__(kThreads, 1) b3d_gemm_kernel(const __grid_constant__ KP p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * 8 + 4 + 4];
  __shared__ uint64_t abars[2 * 3];                  // halo form: full / empty barriers of the patch ring
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) float s_bias2[2][256];   // bias slice of the current N tile, one copy per epilogue group (bf16 TMA epilogue)

  // warp index through a shuffle so the compiler knows it is warp-uniform: the producer and MMA loops below then run on
  // uniform registers (a `lane == 0` region made every TMA / MMA operand go through an ELECT + R2UR.BROADCAST loop, and
  // the ~130 instructions per K block of the single issuing thread, not the tensor pipe, paced the kernel).
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int S = p.num_stages;
  auto full_bar = [&](int s) { return smem_u32(&bars[s]); };
  auto empty_bar = [&](int s) { return smem_u32(&bars[8 + s]); };
  auto tfull_bar = [&](int s) { return smem_u32(&bars[16 + s]); };
  auto tempty_bar = [&](int s) { return smem_u32(&bars[18 + s]); };
  auto resid_bar = [&](int s) { return smem_u32(&bars[20 + s]); };
  auto afull_bar = [&](int s) { return smem_u32(&abars[s]); };
  auto aempty_bar = [&](int s) { return smem_u32(&abars[3 + s]); };

  if (threadIdx.x == 0) {
    for (int s = 0; s < (p.halo ? p.b_stages : S); ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 3; ++s) { mbar_init(afull_bar(s), 1); mbar_init(aempty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), (kPair ? 2u : 1u) * ((p.epi_tma && !p.out_fp32) ? 8u : 4u)); mbar_init(resid_bar(s), 1); mbar_init(resid_bar(2 + s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x == 64) {      // descriptor fetches overlap the barrier init / TMEM allocation (short launches are latency-bound)
    prefetch_tensormap(&p.tmA[0]);
    prefetch_tensormap(&p.tmB);
    if (p.epi_tma) prefetch_tensormap(&p.tmO);
  }
  if (warp == 1) {
    if constexpr (kPair) tmem_alloc2(smem_u32(&tmem_base_smem), (uint32_t)p.tmem_cols);
    else tmem_alloc(smem_u32(&tmem_base_smem), (uint32_t)p.tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (kPair) cluster_sync_all();     // the peer's barriers are initialised before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  b3d_pdl_sync();   // everything above overlapped the previous kernel's tail; global memory is touched only below
  // CTA pair: both CTAs walk the same tile sequence; rank 1 takes the second 128 rows (offset along m2_dim) and the
  // second half of the B tile's rows, rank 0 issues the M = 256 MMAs for both
  const uint32_t crank = kPair ? __shfl_sync(0xffffffffu, cluster_ctarank(), 0) : 0u;    // (shuffle: provably warp-uniform)
  const int t_first = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, t_step = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int rstep = (int)crank * (p.m2 ? 2 : 1);      // boxes this CTA's rows start after (a CTA of a pair may itself hold two)
  const int half_x = kPair && p.m2_dim == 0 ? rstep * p.bw : 0, half_y = kPair && p.m2_dim == 1 ? rstep * p.bh : 0,
            half_n = kPair && p.m2_dim == 2 ? rstep * p.bnn : 0;

  if (warp == 0) {
    // ===================================================================== TMA producer
    // The whole warp walks the (warp-uniform) loops and polls the barriers; one elected lane issues the copies.
    {
      int stage = 0; uint32_t phase = 0;
      int hs = 0; uint32_t hphase = 0;
      for (int t = t_first; t < p.tiles_total; t += t_step) {
        const TileCoord c = decode_tile(p, t);
        int kb0, nkb; k_range(p, c, kb0, nkb);
        const int ncol0 = c.tn * p.BN;
        auto acquire = [&]() -> uint32_t {
          mbar_wait(empty_bar(stage), phase ^ 1);
          return smem_base + (uint32_t)stage * p.stage_bytes;
        };
        auto advance = [&]() { if (++stage == S) { stage = 0; phase ^= 1; } };
        if (p.halo) {
          // (stage / phase above are the weight-tile ring's; hs / hphase the patch ring's -- both persist across tiles)
          const int px = c.x0 - 1, py = c.y0 + half_y - 1, pn = c.n0 + half_n;
          const uint32_t a_bytes_patch = (uint32_t)p.h_rows * 1280u, a_bytes_center = (uint32_t)(p.h_rows - 2) * 1280u;
          for (int q = 0; q < p.n_hsrc; ++q) {
            const auto hq = p.hsrc[q];
            for (int cb = 0; cb < hq.ncb; ++cb) {
              mbar_wait(aempty_bar(hs), hphase ^ 1);
              const uint32_t da = smem_base + (uint32_t)hs * p.h_stage_bytes;
              if (elect_one()) {
                const uint32_t nb = hq.taps == 9 ? a_bytes_patch : a_bytes_center;
                const int r0 = hq.taps == 9 ? 0 : 1, r1 = hq.taps == 9 ? p.h_rows : p.h_rows - 1;
                if (p.h_pitch == 1280u) {       // packed rows: the whole patch is one box (1-tap sources load it too: same map)
                  if constexpr (kPair) {
                    const uint32_t lead_bar = mapa_shared(afull_bar(hs), 0u);
                    if (crank == 0) mbar_expect_tx(afull_bar(hs), 2u * a_bytes_patch);
                    tma_load_4d_pair(da, &p.tmA[hq.src], lead_bar, cb * BK, px, py, pn);
                  } else {
                    mbar_expect_tx(afull_bar(hs), a_bytes_patch);
                    tma_load_4d(da, &p.tmA[hq.src], afull_bar(hs), cb * BK, px, py, pn);
                  }
                } else if constexpr (kPair) {
                  const uint32_t lead_bar = mapa_shared(afull_bar(hs), 0u);
                  if (crank == 0) mbar_expect_tx(afull_bar(hs), 2u * nb);
                  for (int r = r0; r < r1; ++r) tma_load_4d_pair(da + (uint32_t)r * 2048u, &p.tmA[hq.src], lead_bar, cb * BK, px, py + r, pn);
                } else {
                  mbar_expect_tx(afull_bar(hs), nb);
                  for (int r = r0; r < r1; ++r) tma_load_4d(da + (uint32_t)r * 2048u, &p.tmA[hq.src], afull_bar(hs), cb * BK, px, py + r, pn);
                }
              }
              for (int tp = 0; tp < hq.taps; ++tp) {
                mbar_wait(empty_bar(stage), phase ^ 1);
                const uint32_t db = smem_base + p.b_ring_off + (uint32_t)stage * p.b_bytes;
                if (elect_one()) {
                  const int bk = hq.bk0 + (tp * hq.ncb + cb) * BK;
                  if constexpr (kPair) {
                    const uint32_t lead_bar = mapa_shared(full_bar(stage), 0u);
                    if (crank == 0) mbar_expect_tx(full_bar(stage), 2u * p.b_tx);
                    tma_load_4d_pair(db, &p.tmB, lead_bar, bk, ncol0 + (int)crank * (p.BN / 2), 0, 0);
                  } else {
                    mbar_expect_tx(full_bar(stage), p.b_tx);
                    tma_load_4d(db, &p.tmB, full_bar(stage), bk, ncol0, 0, 0);
                  }
                }
                if (++stage == p.b_stages) { stage = 0; phase ^= 1; }
              }
              if (++hs == p.a_stages) { hs = 0; hphase ^= 1; }
            }
          }
        } else if (p.a_major == 0 && p.b_major == 0) {
          const int by = p.batched ? c.y0 : 0, bn = p.batched ? (c.n0 ^ p.b_xor) : 0;
          for (int s = 0; s < p.nseg; ++s) {
            const Seg sg = p.seg[s];
            const int ax = (c.x0 + half_x) * p.stride + sg.dx, ay = (c.y0 + half_y) * p.stride + sg.dy, an = c.n0 + half_n;
            for (int cb = 0; cb < sg.ncb; ++cb) {
              const uint32_t sa = acquire();
              if (elect_one()) {
                if constexpr (kPair) {    // both CTAs' bytes are counted on the leader's barrier, armed by the leader for the pair
                  const uint32_t lead_bar = mapa_shared(full_bar(stage), 0u);
                  if (crank == 0) mbar_expect_tx(full_bar(stage), 2u * (p.a_tx + p.b_tx));
                  tma_load_4d_pair(sa, &p.tmA[sg.src], lead_bar, cb * BK, ax, ay, an);
                  tma_load_4d_pair(sa + p.a_bytes, &p.tmB, lead_bar, sg.bk0 + cb * BK, ncol0 + (int)crank * (p.BN / 2), by, bn);
                } else {
                  mbar_expect_tx(full_bar(stage), p.a_tx + p.b_tx);
                  tma_load_4d(sa, &p.tmA[sg.src], full_bar(stage), cb * BK, ax, ay, an);
                  tma_load_4d(sa + p.a_bytes, &p.tmB, full_bar(stage), sg.bk0 + cb * BK, ncol0, by, bn);
                }
              }
              advance();
            }
          }
        } else if (p.a_major == 0) {  // (0,1): K = channels of A = pixels (x) of B, batched
          for (int cb = 0; cb < nkb; ++cb) {
            const uint32_t sa = acquire();
            if (elect_one()) {
              mbar_expect_tx(full_bar(stage), p.a_tx + p.b_tx);
              tma_load_4d(sa, &p.tmA[0], full_bar(stage), cb * BK, c.x0, c.y0, c.n0);
              for (int j = 0; j < p.b_chunks; ++j)
                tma_load_4d(sa + kABytes + j * 8192, &p.tmB, full_bar(stage), ncol0 + j * 64, cb * BK, c.y0, c.n0 ^ p.b_xor);
            }
            advance();
          }
        } else {  // (1,1): K = pixels of both
          const int tap0 = c.tap * p.tpg;
          // several taps per tile: the B tile is the taps' shifted pixel windows side by side (cpt 64-channel chunks each)
          const int nchunks = p.tpg > 1 ? min(p.b_chunks, (p.b_taps - tap0) * p.cpt) : p.b_chunks;
          // per-chunk channel offset and tap shift, hoisted out of the K loop (max 4 chunks per tile)
          int ch_c[4], ch_dx[4], ch_dy[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int tp = p.tpg > 1 ? tap0 + j / p.cpt : tap0;
            ch_c[j] = p.tpg > 1 ? (j % p.cpt) * 64 : ncol0 + j * 64;
            ch_dx[j] = p.b_taps == 9 ? (tp % 3) - p.pad : 0;
            ch_dy[j] = p.b_taps == 9 ? (tp / 3) - p.pad : 0;
          }
          // pixel-block counters advanced incrementally (a div/mod pair per K block made this warp the bottleneck)
          int bx = 0, by = 0, bn = 0;
          if (!p.batched) { int pb = kb0; bx = pb % p.kblk_x; pb /= p.kblk_x; by = pb % p.kblk_y; bn = pb / p.kblk_y; }
          for (int i = 0; i < nkb; ++i) {
            int kx, ky, kn;
            if (p.batched) { kx = i * p.kbw; ky = c.y0; kn = c.n0; }
            else {
              kx = bx * p.kbw; ky = by * p.kbh; kn = bn * p.kbn;
              if (++bx == p.kblk_x) { bx = 0; if (++by == p.kblk_y) { by = 0; ++bn; } }
            }
            const uint32_t sa = acquire();
            if (elect_one()) {
              mbar_expect_tx(full_bar(stage), p.a_tx + (uint32_t)nchunks * 8192u);
              tma_load_4d(sa, &p.tmA[0], full_bar(stage), c.m0, kx, ky, kn);
              tma_load_4d(sa + 8192, &p.tmA[0], full_bar(stage), c.m0 + 64, kx, ky, kn);
              const int sx = kx * p.stride, sy = ky * p.stride, nb = kn ^ p.b_xor;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (j < nchunks) tma_load_4d(sa + kABytes + j * 8192, &p.tmB, full_bar(stage), ch_c[j], sx + ch_dx[j], sy + ch_dy[j], nb);
            }
            advance();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    // Warp-uniform loop; the same elected lane issues every tcgen05.mma and the commits that track them.
    // In a CTA pair only the leader issues (cta_group::2); its commits arrive on both CTAs' barriers.
    if (crank == 0) {
      int stage = 0; uint32_t phase = 0;
      int hs = 0; uint32_t hphase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      const uint32_t a_kstep = p.a_major ? (2048u >> 4) : (32u >> 4), b_kstep = p.b_major ? (2048u >> 4) : (32u >> 4);
      const uint32_t a_lbo = p.a_major ? 8192u : 16u, b_lbo = p.b_major ? 8192u : 16u;
      const uint64_t a_desc0 = umma_desc(0u, a_lbo, 1024u), b_desc0 = umma_desc(0u, b_lbo, 1024u);
      const uint32_t idesc = p.idesc;
      for (int t = t_first; t < p.tiles_total; t += t_step) {
        const TileCoord c = decode_tile(p, t);
        int kb0, nkb; k_range(p, c, kb0, nkb);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_stride);
        // convolution / dense form: the last 64-channel block of a K segment may hold fewer than 64 real channels (the
        // 144-channel conditioning input, the 8-channel stem input): only its real 16-channel steps are issued
        const bool segk = p.a_major == 0 && p.b_major == 0;
        int sidx = 0, cbi = 0, ncb_cur = segk ? p.seg[0].ncb : 0x7fffffff;
        if (p.halo) {
          // per 64-channel block: one patch, nine (or one) weight tiles; tap (dx, dy) of accumulator a reads the patch from row
          // a*16 + dy + 1, pixel dx + 1 on: groups of 8 pixels (one image row of the tile) 2048 bytes apart
          bool first = true;
          for (int q = 0; q < p.n_hsrc; ++q) {
            const auto hq = p.hsrc[q];
            for (int cb = 0; cb < hq.ncb; ++cb) {
              const int ksteps = cb == hq.ncb - 1 ? hq.klast : BK / 16;
              mbar_wait(afull_bar(hs), hphase);
              tc_fence_after();
              const uint32_t pa = smem_base + (uint32_t)hs * p.h_stage_bytes;
              for (int tp = 0; tp < hq.taps; ++tp) {
                const int dx = hq.taps == 9 ? tp % 3 - 1 : 0, dy = hq.taps == 9 ? tp / 3 - 1 : 0;
                mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                const uint32_t sb = smem_base + p.b_ring_off + (uint32_t)stage * p.b_bytes;
                const uint64_t bd = b_desc0 + (uint64_t)(sb >> 4);
                if (elect_one()) {
                  for (int a = 0; a < p.h_accs; ++a) {
                    const uint32_t sa = pa + (uint32_t)(a * 16 + dy + 1) * p.h_pitch + (uint32_t)(dx + 1) * 128u;
                    // (no base offset: the 128-byte swizzle of TMA and UMMA is a function of the absolute shared-memory address
                    // bits, measured -- a start shifted by whole rows, and groups 1280 bytes apart, read what TMA wrote there; with
                    // the descriptor's base_offset set to the start row mod 8 the results were wrong)
                    const uint64_t ad = umma_desc(sa, 16u, p.h_pitch);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                      if (k < ksteps) {
                        if constexpr (kPair) umma_bf16_pair(d_tmem + (uint32_t)a * p.acc_half, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc, (first && k == 0) ? 0u : 1u);
                        else umma_bf16(d_tmem + (uint32_t)a * p.acc_half, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc, (first && k == 0) ? 0u : 1u);
                      }
                  }
                  if constexpr (kPair) umma_commit_pair(empty_bar(stage)); else umma_commit(empty_bar(stage));
                }
                first = false;
                if (++stage == p.b_stages) { stage = 0; phase ^= 1; }
              }
              if (elect_one()) { if constexpr (kPair) umma_commit_pair(aempty_bar(hs)); else umma_commit(aempty_bar(hs)); }
              if (++hs == p.a_stages) { hs = 0; hphase ^= 1; }
            }
          }
        }
        for (int kb = 0; kb < (p.halo ? 0 : nkb); ++kb) {
          int ksteps = BK / 16;
          if (segk) {
            if (cbi == ncb_cur - 1) ksteps = p.seg[sidx].klast;
            if (++cbi == ncb_cur) { cbi = 0; ++sidx; ncb_cur = sidx < p.nseg ? p.seg[sidx].ncb : 0x7fffffff; }
          }
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + (uint32_t)stage * p.stage_bytes;
          const uint64_t ad = a_desc0 + (uint64_t)(sa >> 4), bd = b_desc0 + (uint64_t)((sa + p.a_bytes) >> 4);
          if constexpr (kPair) {
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                if (k < ksteps) umma_bf16_pair(d_tmem, ad + (uint64_t)(k * a_kstep), bd + (uint64_t)(k * b_kstep), idesc, (kb | k) != 0 ? 1u : 0u);
              if (p.m2) {                   // second 128 rows of each CTA's A tile, second accumulator
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)
                  if (k < ksteps)
                    umma_bf16_pair(d_tmem + p.acc_half, ad + (uint64_t)((kABytes >> 4) + k * a_kstep), bd + (uint64_t)(k * b_kstep), idesc,
                                   (kb | k) != 0 ? 1u : 0u);
              }
              umma_commit_pair(empty_bar(stage));
            }
          } else if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              if (k < ksteps) umma_bf16(d_tmem, ad + (uint64_t)(k * a_kstep), bd + (uint64_t)(k * b_kstep), idesc, (kb | k) != 0 ? 1u : 0u);
            if (p.m2) {                     // rows 128..255 of the A tile against the same B tile, into the second accumulator
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                if (k < ksteps)
                  umma_bf16(d_tmem + p.acc_half, ad + (uint64_t)((kABytes >> 4) + k * a_kstep), bd + (uint64_t)(k * b_kstep), idesc,
                            (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit(empty_bar(stage));  // frees the smem slot once these MMAs have read it
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) { if constexpr (kPair) umma_commit_pair(tfull_bar(acc)); else umma_commit(tfull_bar(acc)); }   // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 6 && !(p.epi_tma && !p.out_fp32)) {
    // (the second epilogue group only works in the bf16 TMA-store epilogue)
  } else if (p.epi_tma && p.out_fp32) {
    // ===================================================================== epilogue, fp32 outputs through TMA (warps 2..5)
    // 32 fp32 columns (one 128-byte row) per round: TMEM -> registers -> epilogue math -> swizzled smem tile -> TMA store,
    // or TMA reduce-add for split-K weight gradients (the fp32 adds happen in L2, asynchronously, with full-line bursts;
    // per-thread RED instructions to 32 different rows made the LSU the bottleneck of short-K splits).
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const bool leader = threadIdx.x == 64;
    const uint32_t so_base = smem_base + p.epi_off;
    int acc = 0; uint32_t acc_phase = 0;
    int buf = 0;
    const uint32_t swz = (uint32_t)(r & 7);
    for (int t = t_first; t < p.tiles_total; t += t_step) {
      const TileCoord c = decode_tile(p, t);
      int cx, cy, cn, img;
      if (p.a_major == 0) {
        cx = c.x0; cy = c.y0; cn = c.n0 ^ p.o_xor;
        img = c.n0 + r / (p.bw * p.bh);
        if (img >= p.N) img = p.N - 1;
      } else {
        cx = c.m0; cy = c.y0; cn = c.n0 ^ p.o_xor; img = c.n0;
      }
      const bool grouped = p.tpg > 1;    // the tile spans several whole taps: every column below ncols_tile is real
      const int ncols_tile = grouped ? min(p.BN, (p.b_taps - c.tap * p.tpg) * p.n_cols) : min(p.BN, p.n_cols - c.tn * p.BN);
      const int nch = (ncols_tile + 31) / 32;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(acc * p.acc_stride) + ((uint32_t)(quad * 32) << 16);
      for (int ch = 0; ch < nch; ++ch) {
        const int cc = c.tn * p.BN + ch * 32;
        if (leader) tma_store_wait_read<1>();
        epi_bar_sync();
        uint32_t v[32];
        tmem_ld32(taddr + ch * 32, v);
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
        const bool full = grouped || cc + 32 <= p.n_cols;
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 32; ++i) if (full || cc + i < p.n_cols) f[i] += __ldg(p.bias + cc + i);
        }
        if (p.rowvec) {
          const float* rv = p.rowvec + (long long)img * p.rowvec_ld + cc;
#pragma unroll
          for (int i = 0; i < 32; ++i) if (full || cc + i < p.n_cols) f[i] += __ldg(rv + i);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] *= p.out_scale;
        if (p.act == 1) {
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = b3d_silu(f[i]);
        }
        if (!full) {     // columns past n_cols land in the padding of a packed gradient slab: add exact zeros there
#pragma unroll
          for (int i = 0; i < 32; ++i) if (cc + i >= p.n_cols) f[i] = 0.f;
        }
        const uint32_t orow = so_base + buf * 16384 + (uint32_t)r * 128u;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint4 q;
          q.x = __float_as_uint(f[4 * g]); q.y = __float_as_uint(f[4 * g + 1]); q.z = __float_as_uint(f[4 * g + 2]); q.w = __float_as_uint(f[4 * g + 3]);
          sts128(orow + ((((uint32_t)g) ^ swz) << 4), q);
        }
        if (ch == nch - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        fence_proxy_async();
        epi_bar_sync();
        if (leader) {
          const int col = cc + c.tap * p.tpg * (int)p.os_tap;     // grouped: os_tap == n_cols, the taps' slabs are contiguous
          if (p.out_atomic) tma_reduce_add_4d(&p.tmO, so_base + buf * 16384, col, cx, cy, cn);
          else tma_store_4d(&p.tmO, so_base + buf * 16384, col, cx, cy, cn);
          tma_store_commit();
        }
        buf ^= 1;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (leader) tma_store_wait_all();
  } else if (p.epi_tma) {
    // ===================================================================== epilogue, bf16 outputs (warps 2..5)
    // TMEM -> registers -> (bias / per-image vector / residual / scale / SiLU) -> bf16 -> 128B-swizzled smem tile
    // -> TMA store.  The residual tile arrives by TMA load.  64 output columns (one 128-byte row) per round,
    // double buffered, so global memory sees only full-line bulk transfers and the LSU only shared memory.
    // Two groups of four warps (2..5 and 6..9; a warp reads the TMEM lanes of its quadrant warp & 3): the 64-column rounds of a
    // tile alternate between them, each group with its own staging tile, residual tile, barrier and TMA-issuing leader.  With
    // one group the epilogue of the 128-column tiles was as long as their mainloop; a single warp per scheduler cannot hide
    // its own TMEM / shared-memory latencies.
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const int grp = (warp - 2) >> 2;
    const bool leader = threadIdx.x == 64 + grp * 128;
    float* s_bias = s_bias2[grp];
    auto grp_sync = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory"); };
    const uint32_t so_base = smem_base + p.epi_off + (uint32_t)grp * 16384u;
    const uint32_t sx_base = smem_base + p.epi_off + 2 * 16384 + (uint32_t)grp * 16384u;
    const uint32_t rows_bytes = p.a_major == 0 ? (uint32_t)(p.bw * p.bh * p.bnn) * 128u : 128u * 128u;
    int acc = 0; uint32_t acc_phase = 0;
    uint32_t rph0 = 0, rph1 = 0;
    int rb = 0;                                  // residual buffer of this group's current round
    const uint32_t rstride = p.resid2 ? 2u * 16384u : 0u;
    int bias_tn = -1;
    const uint32_t swz = (uint32_t)(r & 7);
    auto arrive_tempty = [&]() {       // this warp has read all it needs of the accumulator(s)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (kPair) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0u));
        else mbar_arrive(tempty_bar(acc));
      }
    };
    for (int t = t_first; t < p.tiles_total; t += t_step) {
      const TileCoord c = decode_tile(p, t);
      int cx, cy, cn, img;
      if (p.a_major == 0) {
        cx = c.x0 + half_x; cy = c.y0 + half_y; cn = (c.n0 + half_n) ^ p.o_xor;
        img = c.n0 + half_n + r / (p.bw * p.bh);
        if (img >= p.N) img = p.N - 1;
      } else {
        cx = c.m0; cy = c.y0; cn = c.n0 ^ p.o_xor; img = c.n0;
      }
      const int nch = (min(p.BN, p.n_cols - c.tn * p.BN) + 63) / 64;
      const int rounds = (1 + p.m2) * nch;      // (accumulator half, 64-column chunk), half-major; round rd belongs to group rd & 1
      if (p.bias && c.tn != bias_tn) {            // stage this N tile's bias slice in shared memory (zeros past n_cols): the
        grp_sync();                               // per-round __ldg's of it sat on the critical path of short-K tiles
        for (int i = r; i < 256; i += 128) s_bias[i] = (i < p.BN && c.tn * p.BN + i < p.n_cols) ? __ldg(p.bias + c.tn * p.BN + i) : 0.f;
        bias_tn = c.tn;
        grp_sync();
      }
      // output-box origin of a half: the second 128 rows of a doubled tile continue along m2_dim (x / y / image)
      const int cx0 = cx, cy0 = cy, cn0 = cn, img0 = img;
      const int hx = p.m2_dim == 0 ? p.bw : 0, hy = p.m2_dim == 1 ? p.bh : 0, hn = p.m2_dim == 2 ? p.bnn : 0;
      auto load_resid = [&](int rd, int b) {      // (leader) residual tile of round rd into buffer b of this group
        const int nh = rd >= nch ? 1 : 0, nc = rd - nh * nch;
        mbar_expect_tx(resid_bar(grp * 2 + b), rows_bytes);
        tma_load_4d(sx_base + (uint32_t)b * rstride, &p.tmR, resid_bar(grp * 2 + b), c.tn * p.BN + nc * 64, cx0 + nh * hx, cy0 + nh * hy,
                    cn0 + nh * hn);
      };
      if (p.resid != nullptr && leader && grp < rounds) load_resid(grp, rb);
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + (uint32_t)(acc * p.acc_stride) + ((uint32_t)(quad * 32) << 16);
      if (grp >= rounds) arrive_tempty();         // a one-round tile: nothing to read for the second group
      for (int rd = grp; rd < rounds; rd += 2) {
        const int half = rd >= nch ? 1 : 0, ch = rd - half * nch;
        cx = cx0 + half * hx; cy = cy0 + half * hy; cn = cn0 + half * hn;
        img = img0 + half * hn; if (img >= p.N) img = p.N - 1;
        const uint32_t taddr = taddr0 + (uint32_t)half * p.acc_half;
        const int col0 = c.tn * p.BN + ch * 64;
        if (leader) {
          tma_store_wait_read<0>();                // the store that last used this group's tile(s) has finished reading
          if (p.resid2 && rd + 2 < rounds) load_resid(rd + 2, rb ^ 1);      // next round's residual, into the other buffer
        }
        grp_sync();
        if (p.resid != nullptr) {
          mbar_wait(resid_bar(grp * 2 + rb), rb ? rph1 : rph0);
          if (rb) rph1 ^= 1; else rph0 ^= 1;
        }
        constexpr int buf = 0;
        const uint32_t orow = so_base + buf * 16384 + (uint32_t)r * 128u;
        const uint32_t xrow = sx_base + (uint32_t)rb * rstride + (uint32_t)r * 128u;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          const int cc = col0 + hf * 32;
          if (ch * 64 + hf * 32 >= p.BN) break;
          uint32_t v[32];
          tmem_ld32(taddr + ch * 64 + hf * 32, v);
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
          const bool full = cc + 32 <= p.n_cols;
          if (p.bias) {
            const float4* sb4 = reinterpret_cast<const float4*>(s_bias + ch * 64 + hf * 32);     // broadcast reads
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const float4 b4 = sb4[g];
              f[4 * g] += b4.x; f[4 * g + 1] += b4.y; f[4 * g + 2] += b4.z; f[4 * g + 3] += b4.w;
            }
          }
          if (p.rowvec) {
            const float* rv = p.rowvec + (long long)img * p.rowvec_ld + cc;
            if (full) {
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(rv) + g);
                f[4 * g] += b4.x; f[4 * g + 1] += b4.y; f[4 * g + 2] += b4.z; f[4 * g + 3] += b4.w;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (cc + i < p.n_cols) f[i] += __ldg(rv + i);
            }
          }
          if (p.resid != nullptr) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint4 q = lds128(xrow + ((((uint32_t)(hf * 4 + g)) ^ swz) << 4));
              const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float2 t2 = __bfloat1622float2(h2[i]);
                if (p.act == 2) { f[8 * g + 2 * i] *= b3d_silu_grad(t2.x); f[8 * g + 2 * i + 1] *= b3d_silu_grad(t2.y); }   // resid = pre-activation
                else { f[8 * g + 2 * i] += t2.x; f[8 * g + 2 * i + 1] += t2.y; }
              }
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] *= p.out_scale;
          if (p.act == 1 && p.out2 == nullptr) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = b3d_silu(f[i]);
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 q;
            q.x = pack_bf16x2(f[8 * g], f[8 * g + 1]); q.y = pack_bf16x2(f[8 * g + 2], f[8 * g + 3]);
            q.z = pack_bf16x2(f[8 * g + 4], f[8 * g + 5]); q.w = pack_bf16x2(f[8 * g + 6], f[8 * g + 7]);
            sts128(orow + ((((uint32_t)(hf * 4 + g)) ^ swz) << 4), q);
          }
          if (p.out2) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 q;
              q.x = pack_bf16x2(b3d_silu(f[8 * g]), b3d_silu(f[8 * g + 1])); q.y = pack_bf16x2(b3d_silu(f[8 * g + 2]), b3d_silu(f[8 * g + 3]));
              q.z = pack_bf16x2(b3d_silu(f[8 * g + 4]), b3d_silu(f[8 * g + 5])); q.w = pack_bf16x2(b3d_silu(f[8 * g + 6]), b3d_silu(f[8 * g + 7]));
              sts128(xrow + ((((uint32_t)(hf * 4 + g)) ^ swz) << 4), q);
            }
          }
        }
        if (rd + 2 >= rounds) arrive_tempty();      // this group's last round of the tile: hand the TMEM buffer back to the MMA warp
        fence_proxy_async();
        grp_sync();
        if (leader) {
          tma_store_4d(&p.tmO, so_base + buf * 16384, col0, cx, cy, cn);
          if (p.out2) tma_store_4d(&p.tmO2, sx_base + buf * 16384, col0, cx, cy, cn);
          tma_store_commit();
          if (p.resid != nullptr && !p.resid2 && rd + 2 < rounds) load_resid(rd + 2, 0);     // (the group has finished reading its residual tile)
        }
        if (p.resid2) rb ^= 1;
        if (p.gn_sums != nullptr) {
          // GroupNorm statistics of the tile just staged (the consumer's gn_stats pass over HBM disappears): thread t sums
          // the cpg columns of group g over its slice of rows straight from the swizzled tile, slices of one (group, image)
          // are adjacent lanes and combine by shuffles, one lane per (group, image) adds to the global sums.
          const int cpg = p.gn_cpg, ng = 64 / cpg, lpg = 128 / ng, g = r / lpg, sl = r % lpg, rps = ng;
          const uint32_t tile = so_base + (uint32_t)buf * 16384u;
          float s1 = 0.f, s2 = 0.f;
          for (int i = 0; i < rps; ++i) {
            const int row = sl * rps + i;
            const uint32_t rb = tile + (uint32_t)row * 128u, rz = (uint32_t)(row & 7);
            if (cpg == 4) {
              const uint32_t a = rb + ((((uint32_t)(g >> 1)) ^ rz) << 4) + (uint32_t)(g & 1) * 8u;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(a));
              const float v0 = __uint_as_float(w0 << 16), v1 = __uint_as_float(w0 & 0xffff0000u), v2 = __uint_as_float(w1 << 16),
                          v3 = __uint_as_float(w1 & 0xffff0000u);
              s1 += (v0 + v1) + (v2 + v3);
              s2 = fmaf(v0, v0, fmaf(v1, v1, fmaf(v2, v2, fmaf(v3, v3, s2))));
            } else {
              const int upr = cpg >> 3;                        // 16-byte units of this group per row
              for (int u = 0; u < upr; ++u) {
                const uint4 q = lds128(rb + ((((uint32_t)(g * upr + u)) ^ rz) << 4));
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float v0 = __uint_as_float(w[k] << 16), v1 = __uint_as_float(w[k] & 0xffff0000u);
                  s1 += v0 + v1;
                  s2 = fmaf(v0, v0, fmaf(v1, v1, s2));
                }
              }
            }
          }
          const int rpib = p.gn_rpi > 0 ? min(128, p.gn_rpi) : min(128, p.bw * p.bh);     // rows of one image in this box
          const int spi = rpib / rps;                                                      // slices (lanes) per image
          const int seg = min(spi, 32);
          for (int o = seg >> 1; o > 0; o >>= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
          }
          if ((sl & (seg - 1)) == 0) {
            const int row0 = sl * rps;
            const int im = p.gn_rpi > 0 ? (cx + row0) / p.gn_rpi : cn + row0 / (p.bw * p.bh);
            float* dst = p.gn_sums + ((long long)im * 32 + (col0 / cpg + g)) * 2;
            atomicAdd(dst, s1);
            atomicAdd(dst + 1, s2);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (leader) tma_store_wait_all();
  } else {
    // ===================================================================== epilogue, direct stores (fp32 / atomic / odd shapes)
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int r = quad * 32 + lane;
    int acc = 0; uint32_t acc_phase = 0;
    for (int t = t_first; t < p.tiles_total; t += t_step) {
      const TileCoord c = decode_tile(p, t);
      bool row_ok;
      long long off;
      int img = 0;
      if (p.a_major == 0) {
        const int x = c.x0 + r % p.bw;
        const int q = r / p.bw;
        const int y = c.y0 + q % p.bh;
        const int n = c.n0 + q / p.bh;
        row_ok = (r < p.bw * p.bh * p.bnn) && x < p.W && y < p.H && n < p.N;
        img = n;
        off = (long long)(n ^ p.o_xor) * p.os_n + (long long)y * p.os_y + (long long)x * p.os_x;
      } else {
        const int row = c.m0 + r;
        row_ok = row < p.m_rows;
        img = c.n0;
        off = (long long)(c.n0 ^ p.o_xor) * p.os_n + (long long)c.y0 * p.os_y + (long long)row * p.os_x + (long long)c.tap * p.os_tap;
      }
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(acc * p.acc_stride) + ((uint32_t)(quad * 32) << 16);
      const int nchunks = (p.BN + 31) / 32;
      for (int ch = 0; ch < nchunks; ++ch) {
        uint32_t v[32];
        tmem_ld32(taddr + ch * 32, v);
        const int col0 = c.tn * p.BN + ch * 32;
        if (row_ok && col0 < p.n_cols) {
          const int ncol = min(32, min(p.n_cols - col0, p.BN - ch * 32));
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
          if (p.bias) {
#pragma unroll
            for (int i = 0; i < 32; ++i) if (i < ncol) f[i] += __ldg(p.bias + col0 + i);
          }
          if (p.rowvec) {
            const float* rv = p.rowvec + (long long)img * p.rowvec_ld + col0;
#pragma unroll
            for (int i = 0; i < 32; ++i) if (i < ncol) f[i] += __ldg(rv + i);
          }
          const long long o = off + col0;
          const bool vec_ok = (ncol == 32) && ((o & 7) == 0);
          if (p.resid) {
            const bf16* rp = p.resid + o;
            if (vec_ok) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                bf16x8 pk = *reinterpret_cast<const bf16x8*>(rp + g * 8);
                float t8[8]; b3d_unpack8(pk, t8);
#pragma unroll
                for (int i = 0; i < 8; ++i) { if (p.act == 2) f[g * 8 + i] *= b3d_silu_grad(t8[i]); else f[g * 8 + i] += t8[i]; }
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (i < ncol) { if (p.act == 2) f[i] *= b3d_silu_grad(__bfloat162float(rp[i])); else f[i] += __bfloat162float(rp[i]); }
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] *= p.out_scale;
          if (p.act == 1 && p.out2 == nullptr) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = b3d_silu(f[i]);
          }
          if (p.out_fp32) {
            float* op = reinterpret_cast<float*>(p.out) + o;
            if (p.out_atomic) {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (i < ncol) atomicAdd(op + i, f[i]);
            } else if (vec_ok) {
#pragma unroll
              for (int g = 0; g < 8; ++g)
                *reinterpret_cast<float4*>(op + g * 4) = make_float4(f[g * 4], f[g * 4 + 1], f[g * 4 + 2], f[g * 4 + 3]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (i < ncol) op[i] = f[i];
            }
          } else {
            bf16* op = reinterpret_cast<bf16*>(p.out) + o;
            if (vec_ok) {
#pragma unroll
              for (int g = 0; g < 4; ++g) *reinterpret_cast<bf16x8*>(op + g * 8) = b3d_pack8(f + g * 8);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (i < ncol) op[i] = __float2bfloat16(f[i]);
            }
          }
          if (p.out2) {
            bf16* op2 = reinterpret_cast<bf16*>(p.out2) + o;
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = b3d_silu(f[i]);
            if (vec_ok) {
#pragma unroll
              for (int g = 0; g < 4; ++g) *reinterpret_cast<bf16x8*>(op2 + g * 8) = b3d_pack8(f + g * 8);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (i < ncol) op2[i] = __float2bfloat16(f[i]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (kPair) cluster_sync_all();     // neither CTA of a pair frees TMEM or exits while the other may still signal it
  if (warp == 1) {
    tc_fence_after();
    if constexpr (kPair) tmem_dealloc2(tmem_base, (uint32_t)p.tmem_cols);
    else tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}


int choose_bn(int n_cols, int b_major, long long tiles_other) {
  // widest tile that wastes the fewest padded columns ...
  int best = 128;
  if (n_cols <= 256) best = (n_cols + 15) / 16 * 16;
  else {
    const int cands[] = {256, 192, 128, 64};
    long long best_cost = -1;
    for (int c : cands) {
      long long tiles = (n_cols + c - 1) / c;
      long long cost = tiles * c;  // padded columns computed
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = c; }
    }
  }
  // ... but a launch that cannot fill the 148 SMs is narrowed (down to 128, then 64 columns) while that adds CTAs:
  // a half-empty GPU costs more than the lower MMA/shared-memory efficiency of a narrower tile
  const long long sms = b3d_num_sms();
  // (tiles stay multiples of 64 columns: the TMA-store epilogue writes 64-column boxes)
  // -- but only while the narrower tiles still fit ONE wave: crossing the SM count turns one wave of wide tiles into two
  // of narrow ones (measured, 32 images of 8x8, 1024 -> 1024: 128-wide 32 us, 64-wide 51 us, 256-wide 48 us)
  while (best >= 128 && (best / 2) % 64 == 0 && tiles_other * ((n_cols + best - 1) / best) < sms &&
         tiles_other * ((n_cols + best / 2 - 1) / (best / 2)) <= sms)
    best /= 2;
  (void)b_major;
  return best;
}

}  // namespace

static int g_halo = -1;
extern "C" void b3d_set_conv_halo(int on) { g_halo = on; }      // 1: patch rows at a 2048-byte pitch, 2: packed rows (1280 bytes)
static int64_t g_launches = 0;
extern "C" int64_t b3d_launch_count(void) { return g_launches; }
void b3d_count_launch(int n) { g_launches += n; }

// GroupNorm statistics in the epilogue: convolution / dense form, bf16 TMA-store output, n_cols = 32 * cpg, every
// 128-row box full and made of whole row slices of single images.
static bool gn_stats_ok(const b3d_gemm_desc* d) {
  if (d->a_major != 0 || d->b_major != 0 || d->batched || d->out_fp32 || d->out_atomic || d->out_tap_stride != 0) return false;
  const int cpg = d->gn_cpg;
  if (!(cpg == 4 || cpg == 8 || cpg == 16 || cpg == 32) || d->n_cols != 32 * cpg) return false;
  if (d->out_stride[0] % 8 || d->out_stride[1] % 8 || d->out_stride[2] % 8 || ((uintptr_t)d->out & 15)) return false;
  if (d->resid && d->out2) return false;
  const int bw = d->W < BM ? d->W : BM;
  int bh = BM / bw; if (bh > d->H) bh = d->H; if (bh < 1) bh = 1;
  int bnn = (bh == d->H) ? BM / (bw * bh) : 1; if (bnn > d->N) bnn = d->N; if (bnn < 1) bnn = 1;
  if (bw * bh * bnn != BM || d->W % bw || d->H % bh || d->N % bnn) return false;
  const int rps = 64 / cpg;                                      // rows per thread slice
  const int rpi = d->gn_rows_per_image > 0 ? d->gn_rows_per_image : bw * bh;
  if (d->gn_rows_per_image > 0 && !(d->H == 1 && d->N == 1)) return false;
  const int rpib = rpi < BM ? rpi : BM;                          // rows of one image inside a box
  if (rpib % rps || BM % rpib || (rpi > BM && rpi % BM)) return false;
  const int spi = rpib / rps;
  return (spi & (spi - 1)) == 0;
}
extern "C" int b3d_gemm_stats_supported(const b3d_gemm_desc* d) { return d && gn_stats_ok(d) ? 1 : 0; }

extern "C" int b3d_gemm(const b3d_gemm_desc* d, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B3D_CHECK_ARG(d != nullptr && d->out != nullptr, "b3d_gemm: null descriptor/output");
  B3D_CHECK_ARG(d->a_major == 0 || d->a_major == 1, "b3d_gemm: a_major");
  B3D_CHECK_ARG(!(d->a_major == 1 && d->b_major == 0), "b3d_gemm: (a_major=1, b_major=0) is not used by the model");
  B3D_CHECK_ARG(d->n_cols > 0, "b3d_gemm: n_cols");
  const int s = d->conv_stride > 0 ? d->conv_stride : 1;
  B3D_CHECK_ARG(s <= 8, "b3d_gemm: conv_stride > 8 (TMA traversal stride limit)");

  KP p;
  memset(&p, 0, sizeof(p));
  p.a_major = d->a_major; p.b_major = d->b_major; p.batched = d->batched;
  p.b_xor = d->b_batch_xor; p.o_xor = d->out_batch_xor;
  p.W = d->W; p.H = d->H; p.N = d->N; p.stride = s; p.pad = d->pad;
  p.n_cols = d->n_cols; p.m_rows = d->m_rows;
  p.b_taps = d->b_taps > 0 ? d->b_taps : 1;
  p.split_k = d->split_k > 0 ? d->split_k : 1;
  B3D_CHECK_ARG(p.W > 0 && p.H > 0 && p.N > 0, "b3d_gemm: W/H/N");

  // bf16 outputs whose rows are 16-byte aligned leave through smem + TMA store (2 x 16 KB staging, + 2 x 16 KB for
  // the residual tile or the second output)
  const bool al16 = ((uintptr_t)d->out & 15) == 0;
  if (!d->out_fp32)
    p.epi_tma = (!d->out_atomic && d->n_cols % 8 == 0 && d->out_stride[0] % 8 == 0 && d->out_stride[1] % 8 == 0 && d->out_stride[2] % 8 == 0 &&
                 d->out_tap_stride == 0 && al16 && (d->resid == nullptr || d->out2 == nullptr) &&
                 (d->resid == nullptr || ((uintptr_t)d->resid & 15) == 0) && (d->out2 == nullptr || ((uintptr_t)d->out2 & 15) == 0)) ? 1 : 0;
  else   // fp32: plain TMA store, or TMA reduce-add for the atomically accumulated (split-K) weight gradients
    p.epi_tma = (d->n_cols % 4 == 0 && d->out_stride[0] % 4 == 0 && d->out_stride[1] % 4 == 0 && d->out_stride[2] % 4 == 0 &&
                 d->out_tap_stride % 4 == 0 && al16 && d->resid == nullptr && d->out2 == nullptr) ? 1 : 0;

  // Per-tap weight gradients of narrow layers (Ci = 64 / 128 with unpadded slabs): 256 / Ci taps share one tile -- the
  // dy (A) tile is loaded once for all of them and the MMA runs at N = 256 (a 128 x 128 tile per tap was bound by the
  // L2 -> SM operand path at 0.73 PFLOP/s).  Their slabs are adjacent in the packed gradient, so the epilogue is unchanged.
  p.tpg = 1; p.cpt = 1;
  static const bool no_tpg = getenv("B3D_NO_TAPGROUP") != nullptr;     // measurement switch
  if (!no_tpg && d->a_major == 1 && d->b_major == 1 && !d->batched && p.b_taps == 9 && p.epi_tma && d->out_fp32 && d->n_cols % 64 == 0 &&
      d->n_cols <= 128 && d->out_tap_stride == d->n_cols) {
    p.tpg = 256 / d->n_cols; p.cpt = d->n_cols / 64;
  }
  p.tap_groups = (p.b_taps + p.tpg - 1) / p.tpg;

  int BN = p.tpg > 1 ? p.tpg * d->n_cols : d->bn_tile;
  if (BN <= 0) {
    // number of tiles in all other grid dimensions, to see whether the launch can fill the GPU
    long long other;
    if (d->a_major == 0) {
      const int bw = d->W < BM ? d->W : BM;
      int bh = 1, bnn = 1;
      if (!d->batched) {
        bh = BM / bw; if (bh > d->H) bh = d->H; if (bh < 1) bh = 1;
        bnn = (bh == d->H) ? BM / (bw * bh) : 1; if (bnn > d->N) bnn = d->N; if (bnn < 1) bnn = 1;
      }
      other = (long long)((d->W + bw - 1) / bw) * ((d->H + bh - 1) / bh) * ((d->N + bnn - 1) / bnn);
    } else {
      other = (long long)((d->m_rows + BM - 1) / BM) * (d->batched ? (long long)d->H * d->N : 1) * p.b_taps * p.split_k;
    }
    BN = choose_bn(d->n_cols, d->b_major, other);
  }
  B3D_CHECK_ARG(BN % 16 == 0 && BN >= 16 && BN <= 256, "b3d_gemm: tile N %d must be a multiple of 16 in [16,256]", BN);
  p.BN = BN;
  p.tiles_n = p.tpg > 1 ? 1 : (d->n_cols + BN - 1) / BN;
  p.b_chunks = (BN + 63) / 64;
  p.b_bytes = d->b_major ? (uint32_t)p.b_chunks * 8192u : (uint32_t)((BN * 128 + 1023) / 1024 * 1024);
  p.b_tx = d->b_major ? (uint32_t)p.b_chunks * 8192u : (uint32_t)BN * 128u;
  uint32_t epi_bytes = p.epi_tma ? ((d->resid || d->out2) ? 65536u : 32768u) : 0u;
  // Narrow tiles (BN <= 128) of the convolution / dense form: 256 output rows per tile -- two 128-row accumulators fed
  // by ONE B tile.  A 128 x 128 tile loads 32 KB of operands per 2.1 MFLOP and is bound by the L2 -> SM operand path
  // (measured 1.0 PFLOP/s); 256 x 128 loads 48 KB per 4.2 MFLOP.  The doubled tile continues along the outermost
  // dimension of the pixel box (rows, else images, else x for the dense form), so each half is an ordinary box for the
  // epilogue's TMA stores.  Needs 4 * BN TMEM columns (double-buffered) and enough tiles to keep every SM busy.
  p.a_bytes = kABytes; p.acc_half = 0;
  int m2_bw = 0, m2_bh = 0, m2_bnn = 0;
  static const bool no_m2 = getenv("B3D_NO_M2") != nullptr;      // measurement switch
  if (!no_m2 && d->a_major == 0 && d->b_major == 0 && !d->batched && s == 1 && p.epi_tma && !d->out_fp32 && BN <= 128) {
    const int bw = d->W < BM ? d->W : BM;
    int bh = BM / bw; if (bh > d->H) bh = d->H; if (bh < 1) bh = 1;
    int bnn = (bh == d->H) ? BM / (bw * bh) : 1; if (bnn > d->N) bnn = d->N; if (bnn < 1) bnn = 1;
    int dim = -1;
    if (bw * bh * bnn == BM) {
      if (bh < d->H) dim = 1;
      else if (2 * bnn <= d->N) dim = 2;
      else if (d->H == 1 && d->N == 1 && d->W >= 2 * BM) dim = 0;
    }
    if (dim >= 0) {
      const int tbw = dim == 0 ? 2 * bw : bw, tbh = dim == 1 ? 2 * bh : bh, tbnn = dim == 2 ? 2 * bnn : bnn;
      const long long tiles = (long long)((d->W + tbw - 1) / tbw) * ((d->H + tbh - 1) / tbh) * ((d->N + tbnn - 1) / tbnn) * p.tiles_n;
      const bool fits = d->gn_sums == nullptr || (d->W % tbw == 0 && d->H % tbh == 0 && d->N % tbnn == 0);   // statistics need full tiles
      if (fits && tiles >= 2LL * b3d_num_sms()) { p.m2 = 1; p.m2_dim = dim; m2_bw = tbw; m2_bh = tbh; m2_bnn = tbnn; p.a_bytes = 2 * kABytes; }
    }
  }
  // CTA pairs (cta_group::2): the same 256-row tile, but its halves live in the two CTAs of a cluster and ONE M = 256 MMA
  // reads A from both and half of the B tile from each -- per SM half the B shared-memory traffic of a single-CTA tile
  // (the 128 B/clk shared-memory read port is what keeps a lone CTA's M = 128 MMAs below the tensor pipe's rate).
  // Measured (B200, batch 128): +7.6% on the 256 -> 256 convolutions (1.43 -> 1.54 PFLOP/s), +2% on the FiLM GEMM; for
  // BN = 128 the single-CTA 256-row tile above is faster (1.12 vs 0.99 PFLOP/s) and the dual-output epilogue loses 5%,
  // so pairs are used for full-width tiles with one output.  B3D_NO_C2=1 turns them off (measurement switch).
  static const bool use_c2 = getenv("B3D_NO_C2") == nullptr;
  int c2_kblocks = 0;       // K blocks per tile: short reductions (the K = 256 / 512 dense layers of the attention blocks) lose to
  if (d->a_major == 0 && d->b_major == 0)   // the pair's extra synchronisation (23.5 -> 26.3 us at [65536,256]x[256,256])
    for (int i = 0; i < d->n_src && i < 3; ++i) c2_kblocks += (d->a_taps[i] == 9 ? 9 : 1) * ((d->a[i].dim[0] + BK - 1) / BK);
  // Narrow tiles (BN <= 128) combine both: each CTA of the pair keeps two accumulators (512 rows per pair tile).
  static const bool use_c2m2 = getenv("B3D_NO_C2M2") == nullptr;
  const bool pair_wide = BN == 256, pair_quad = use_c2m2 && BN <= 128 && BN % 32 == 0;
  if (use_c2 && d->a_major == 0 && d->b_major == 0 && !d->batched && s == 1 && p.epi_tma && !d->out_fp32 && (pair_wide || pair_quad) &&
      d->out2 == nullptr && c2_kblocks >= 16 && b3d_num_sms() % 2 == 0) {
    const int bw = d->W < BM ? d->W : BM;
    int bh = BM / bw; if (bh > d->H) bh = d->H; if (bh < 1) bh = 1;
    int bnn = (bh == d->H) ? BM / (bw * bh) : 1; if (bnn > d->N) bnn = d->N; if (bnn < 1) bnn = 1;
    int dim = -1;
    if (bw * bh * bnn == BM) {
      if (bh < d->H) dim = 1;
      else if (2 * bnn <= d->N) dim = 2;
      else if (d->H == 1 && d->N == 1 && d->W >= 2 * BM) dim = 0;
    }
    if (dim >= 0) {
      const int mul = pair_wide ? 2 : 4;            // 128-row boxes per pair tile
      const int tbw = dim == 0 ? mul * bw : bw, tbh = dim == 1 ? mul * bh : bh, tbnn = dim == 2 ? mul * bnn : bnn;
      const long long tiles = (long long)((d->W + tbw - 1) / tbw) * ((d->H + tbh - 1) / tbh) * ((d->N + tbnn - 1) / tbnn) * p.tiles_n;
      const bool fits = d->gn_sums == nullptr || (d->W % tbw == 0 && d->H % tbh == 0 && d->N % tbnn == 0);
      if (fits && tiles >= (long long)b3d_num_sms() && (dim != 0 || tbw <= 512)) {      // >= 2 tiles per pair
        p.c2 = 1; p.m2 = pair_wide ? 0 : 1; p.m2_dim = dim; m2_bw = tbw; m2_bh = tbh; m2_bnn = tbnn;
        p.a_bytes = pair_wide ? kABytes : 2 * kABytes;
        p.b_bytes = (uint32_t)((BN / 2 * 128 + 1023) / 1024 * 1024);
        p.b_tx = (uint32_t)(BN / 2) * 128u;
      }
    }
  }
  // Halo-patch form (see KP): stride-1 3x3 convolutions whose image tiles into 8 x (16 * accumulators * CTAs) pixel tiles
  if (g_halo < 0) g_halo = getenv("B3D_HALO") ? atoi(getenv("B3D_HALO")) : 2;
  if (g_halo && d->a_major == 0 && d->b_major == 0 && !d->batched && s == 1 && d->pad == 1 && p.epi_tma && !d->out_fp32) {
    bool any9 = false, ok = d->n_src >= 1 && d->n_src <= 3;
    for (int i = 0; ok && i < d->n_src; ++i) { if (d->a_taps[i] == 9) any9 = true; else if (d->a_taps[i] != 1) ok = false; }
    const int accs = p.m2 ? 2 : 1, th = 16 * accs * (p.c2 ? 2 : 1);
    const uint32_t hpitch = g_halo == 2 ? 1280u : 2048u;
    const uint32_t hbytes = ((uint32_t)(16 * accs + 2) * hpitch + 1023u) / 1024u * 1024u;
    // two residual tiles per epilogue group when the B ring keeps at least 4 stages with them (measurement switch B3D_NO_RESID2)
    static const bool no_r2 = getenv("B3D_NO_RESID2") != nullptr;
    const uint32_t r2_bytes = (d->resid && !d->out2 && !no_r2 &&
                               (long long)kMaxDynSmem - 1024 - epi_bytes - 32768 - 2LL * hbytes >= 4LL * p.b_bytes) ? 32768u : 0u;
    const long long room = (long long)kMaxDynSmem - 1024 - epi_bytes - r2_bytes - 2LL * hbytes;
    // a CTA pair whose tile would be taller than the image takes two images instead (16 x 16 level): rank 1 = image n + 1
    const bool pair_images = p.c2 && !p.m2 && d->H % th != 0 && d->H % 16 == 0 && d->N % 2 == 0;
    if (ok && any9 && d->W % 8 == 0 && (d->H % th == 0 || pair_images) && room >= 2LL * p.b_bytes) {
      if (r2_bytes) { p.resid2 = 1; epi_bytes += r2_bytes; }
      p.halo = 1; p.h_accs = accs; p.h_rows = 16 * accs + 2; p.h_stage_bytes = hbytes; p.a_stages = 2; p.h_pitch = hpitch;
      p.b_stages = (int)(room / p.b_bytes); if (p.b_stages > 8) p.b_stages = 8;
      if (room - (long long)p.b_stages * p.b_bytes >= (long long)hbytes && p.b_stages >= 4) { p.a_stages = 3; }      // a third patch when it is free
      else if (p.b_stages > 4 && room - 4LL * p.b_bytes >= (long long)hbytes) { p.a_stages = 3; p.b_stages = (int)((room - hbytes) / p.b_bytes); if (p.b_stages > 8) p.b_stages = 8; }
      m2_bw = 8; m2_bh = th; m2_bnn = 1; p.m2_dim = 1;
      if (pair_images) { m2_bh = 16; m2_bnn = 2; p.m2_dim = 2; }
    }
  }
  p.stage_bytes = p.a_bytes + p.b_bytes;
  p.num_stages = (int)((kMaxDynSmem - 1024 - epi_bytes) / p.stage_bytes);
  if (p.num_stages > 8) p.num_stages = 8;
  B3D_CHECK_ARG(p.num_stages >= 2, "b3d_gemm: not enough shared memory for a 2-stage pipeline");
  p.epi_off = (uint32_t)p.num_stages * p.stage_bytes;
  if (p.halo) {
    p.b_ring_off = (uint32_t)p.a_stages * p.h_stage_bytes;
    p.epi_off = p.b_ring_off + (uint32_t)p.b_stages * p.b_bytes;
  }
  p.acc_stride = (BN + 31) / 32 * 32;
  if (p.m2) { p.acc_half = (uint32_t)p.acc_stride; p.acc_stride *= 2; }
  p.tmem_cols = pow2_ceil(2 * p.acc_stride);
  if (p.tmem_cols < 32) p.tmem_cols = 32;
  // instruction descriptor: D fp32 (bit 4), A/B bf16 (bits 7,10), majors (15,16), N>>3 at 17, M>>4 at 24
  p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)d->a_major << 15) | ((uint32_t)d->b_major << 16) | ((uint32_t)(BN >> 3) << 17) |
            ((uint32_t)((p.c2 ? 2 * BM : BM) >> 4) << 24);

  const uint32_t ones[4] = {1, 1, 1, 1};
  int rc;
  if (d->a_major == 0) {
    if (d->batched) { p.bw = d->W < BM ? d->W : BM; p.bh = 1; p.bnn = 1; }
    else {
      p.bw = d->W < BM ? d->W : BM;
      p.bh = BM / p.bw; if (p.bh > d->H) p.bh = d->H; if (p.bh < 1) p.bh = 1;
      p.bnn = (p.bh == d->H) ? BM / (p.bw * p.bh) : 1; if (p.bnn > d->N) p.bnn = d->N; if (p.bnn < 1) p.bnn = 1;
    }
    if (p.halo) { p.bw = 8; p.bh = 16; p.bnn = 1; }
    p.tbw = (p.m2 || p.c2) ? m2_bw : p.bw; p.tbh = (p.m2 || p.c2) ? m2_bh : p.bh; p.tbnn = (p.m2 || p.c2) ? m2_bnn : p.bnn;
    p.tiles_x = (d->W + p.tbw - 1) / p.tbw;
    p.tiles_y = (d->H + p.tbh - 1) / p.tbh;
    const int tiles_img = (d->N + p.tbnn - 1) / p.tbnn;
    p.tiles_m = p.tiles_x * p.tiles_y * tiles_img;
    p.tiles_b = 1; p.tiles_z = 1;
    // rows one CTA loads: the whole tile, or its half of a pair tile
    const int adiv = p.c2 ? 2 : 1;
    const int abw = p.m2_dim == 0 ? p.tbw / adiv : p.tbw, abh = p.m2_dim == 1 ? p.tbh / adiv : p.tbh, abn = p.m2_dim == 2 ? p.tbnn / adiv : p.tbnn;
    p.a_tx = (uint32_t)(abw * abh * abn) * 128u;
    uint32_t boxA[4] = {64, (uint32_t)((abw - 1) * s + 1), (uint32_t)((abh - 1) * s + 1), (uint32_t)abn};
    if (p.halo) { boxA[1] = 10; boxA[2] = p.h_pitch == 1280u ? (uint32_t)p.h_rows : 1u; boxA[3] = 1; }      // a whole patch, or one 10-pixel row of it, per TMA load
    const uint32_t esA[4] = {1, (uint32_t)s, (uint32_t)s, 1};
    if (d->b_major == 0) {
      B3D_CHECK_ARG(d->n_src >= 1 && d->n_src <= 3, "b3d_gemm: n_src");
      int bk = 0;
      for (int i = 0; i < d->n_src; ++i) {
        rc = make_map(&p.tmA[i], d->a[i], boxA, esA); if (rc) return rc;
        const int taps = d->a_taps[i] == 9 ? 9 : 1;
        const int ncb = (d->a[i].dim[0] + BK - 1) / BK;
        if (p.halo) {
          auto& hq = p.hsrc[p.n_hsrc++];
          hq.src = i; hq.taps = taps; hq.ncb = ncb; hq.bk0 = bk;
          const int rem = d->a[i].dim[0] - (ncb - 1) * BK;
          hq.klast = (rem + 15) / 16; if (hq.klast < 1) hq.klast = 1; if (hq.klast > BK / 16) hq.klast = BK / 16;
        }
        for (int tp = 0; tp < taps; ++tp) {
          B3D_CHECK_ARG(p.nseg < kMaxSeg, "b3d_gemm: too many K segments");
          Seg& sg = p.seg[p.nseg++];
          sg.src = i; sg.ncb = ncb; sg.bk0 = bk;
          { const int rem = d->a[i].dim[0] - (ncb - 1) * BK; sg.klast = (rem + 15) / 16; if (sg.klast < 1) sg.klast = 1; if (sg.klast > BK / 16) sg.klast = BK / 16; }
          { static const bool no_ks = getenv("B3D_NO_KSTEPS") != nullptr; if (no_ks) sg.klast = BK / 16; }      // measurement switch
          sg.dx = taps == 9 ? (tp % 3) - d->pad : 0;
          sg.dy = taps == 9 ? (tp / 3) - d->pad : 0;
          bk += ncb * BK;
        }
      }
      B3D_CHECK_ARG(bk <= d->b.dim[0] + BK - 1, "b3d_gemm: B has K=%d but A enumerates %d", d->b.dim[0], bk);
      const uint32_t boxB[4] = {64, (uint32_t)(p.c2 ? BN / 2 : BN), 1, 1};
      rc = make_map(&p.tmB, d->b, boxB, ones); if (rc) return rc;
    } else {
      B3D_CHECK_ARG(d->batched && s == 1, "b3d_gemm: (a_major=0,b_major=1) is the batched attention form");
      rc = make_map(&p.tmA[0], d->a[0], boxA, esA); if (rc) return rc;
      p.kchan_blocks = (d->a[0].dim[0] + BK - 1) / BK;
      const uint32_t boxB[4] = {64, 64, 1, 1};
      rc = make_map(&p.tmB, d->b, boxB, ones); if (rc) return rc;
    }
  } else {
    B3D_CHECK_ARG(d->m_rows > 0, "b3d_gemm: m_rows");
    p.tiles_m = (d->m_rows + BM - 1) / BM;
    if (d->batched) { p.kbw = 64; p.kbh = 1; p.kbn = 1; p.tiles_b = d->H * d->N; }
    else {
      p.kbw = pow2_ceil(d->W); if (p.kbw > 64) p.kbw = 64;
      p.kbh = 64 / p.kbw; { int ph = pow2_ceil(d->H); if (p.kbh > ph) p.kbh = ph; }
      p.kbn = 64 / (p.kbw * p.kbh);
      p.tiles_b = 1;
    }
    p.kblk_x = (d->W + p.kbw - 1) / p.kbw;
    p.kblk_y = (d->H + p.kbh - 1) / p.kbh;
    p.kblk_total = d->batched ? p.kblk_x : p.kblk_x * p.kblk_y * ((d->N + p.kbn - 1) / p.kbn);
    if (d->batched) p.split_k = 1;
    if (p.split_k > p.kblk_total) p.split_k = p.kblk_total;
    B3D_CHECK_ARG(p.split_k == 1 || d->out_atomic, "b3d_gemm: split_k > 1 needs out_atomic");
    p.tiles_z = p.tap_groups * p.split_k;
    p.a_tx = 2 * 8192u;
    const uint32_t boxA[4] = {64, (uint32_t)p.kbw, (uint32_t)p.kbh, (uint32_t)p.kbn};
    rc = make_map(&p.tmA[0], d->a[0], boxA, ones); if (rc) return rc;
    const uint32_t boxB[4] = {64, (uint32_t)((p.kbw - 1) * s + 1), (uint32_t)((p.kbh - 1) * s + 1), (uint32_t)p.kbn};
    const uint32_t esB[4] = {1, (uint32_t)s, (uint32_t)s, 1};
    rc = make_map(&p.tmB, d->b, boxB, esB); if (rc) return rc;
  }
  if (p.epi_tma && d->out_fp32) {
    b3d_view4 ov;
    ov.ptr = d->out;
    // columns: n_cols, or (for per-tap weight gradients) every tap slab of the packed matrix up to the last one
    ov.dim[0] = p.b_taps > 1 ? (int)((p.b_taps - 1) * d->out_tap_stride) + d->n_cols : d->n_cols;
    ov.dim[1] = d->a_major == 0 ? d->W : d->m_rows; ov.dim[2] = d->H; ov.dim[3] = d->N;
    ov.stride[0] = d->out_stride[0]; ov.stride[1] = d->out_stride[1]; ov.stride[2] = d->out_stride[2];
    uint32_t boxO[4] = {32, 128, 1, 1};
    if (d->a_major == 0) { boxO[1] = (uint32_t)p.bw; boxO[2] = (uint32_t)p.bh; boxO[3] = (uint32_t)p.bnn; }
    rc = make_map(&p.tmO, ov, boxO, ones, 4); if (rc) return rc;
  } else if (p.epi_tma) {
    b3d_view4 ov;
    ov.ptr = d->out;
    ov.dim[0] = d->n_cols; ov.dim[1] = d->a_major == 0 ? d->W : d->m_rows; ov.dim[2] = d->H; ov.dim[3] = d->N;
    ov.stride[0] = d->out_stride[0]; ov.stride[1] = d->out_stride[1]; ov.stride[2] = d->out_stride[2];
    uint32_t boxO[4] = {64, 128, 1, 1};
    if (d->a_major == 0) { boxO[1] = (uint32_t)p.bw; boxO[2] = (uint32_t)p.bh; boxO[3] = (uint32_t)p.bnn; }
    rc = make_map(&p.tmO, ov, boxO, ones); if (rc) return rc;
    if (d->out2) { ov.ptr = d->out2; rc = make_map(&p.tmO2, ov, boxO, ones); if (rc) return rc; }
    if (d->resid) { ov.ptr = d->resid; rc = make_map(&p.tmR, ov, boxO, ones); if (rc) return rc; }
  }
  p.tiles_total = p.tiles_m * p.tiles_n * p.tiles_b * p.tiles_z;
  B3D_CHECK_ARG(p.tiles_total > 0, "b3d_gemm: empty problem");

  if (d->gn_sums != nullptr) {
    B3D_CHECK_ARG(p.epi_tma && gn_stats_ok(d), "b3d_gemm: gn_sums cannot be honoured for this descriptor (see b3d_gemm_stats_supported)");
    p.gn_sums = d->gn_sums; p.gn_cpg = d->gn_cpg; p.gn_rpi = d->gn_rows_per_image;
  }
  p.out = d->out; p.out2 = d->out2; p.bias = d->bias; p.rowvec = d->rowvec; p.rowvec_ld = d->rowvec_ld;
  p.resid = reinterpret_cast<const bf16*>(d->resid);
  p.os_x = d->out_stride[0]; p.os_y = d->out_stride[1]; p.os_n = d->out_stride[2]; p.os_tap = d->out_tap_stride;
  p.alpha = d->alpha; p.out_scale = d->out_scale; p.act = d->act; p.out_fp32 = d->out_fp32; p.out_atomic = d->out_atomic;
  B3D_CHECK_ARG(!d->out_atomic || d->out_fp32, "b3d_gemm: out_atomic needs fp32 output");
  B3D_CHECK_ARG(d->act != 2 || (d->resid != nullptr && !d->out_fp32 && d->out2 == nullptr), "b3d_gemm: act 2 (times silu'(resid)) needs resid, a bf16 output and no out2");

  const size_t smem = (p.halo ? (size_t)p.epi_off : (size_t)p.num_stages * p.stage_bytes) + epi_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    B3D_CUDA(cudaFuncSetAttribute(b3d_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
    B3D_CUDA(cudaFuncSetAttribute(b3d_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
    attr_set = true;
  }
  int grid = p.tiles_total < b3d_num_sms() ? p.tiles_total : b3d_num_sms();
  if (p.c2) {
    const int pairs = p.tiles_total < b3d_num_sms() / 2 ? p.tiles_total : b3d_num_sms() / 2;
    grid = 2 * pairs;
  }
  if (p.c2) B3D_CUDA(b3d_launch_cluster_pdl(b3d_gemm_kernel<true>, dim3(grid), dim3(kThreads), smem, stream, 2, p));
  else B3D_CUDA(b3d_launch_pdl(b3d_gemm_kernel<false>, dim3(grid), dim3(kThreads), smem, stream, p));
  B3D_LAUNCH_CHECK();
  b3d_count_launch(1);
  return B3D_OK;
}
