// Shared helpers for the b3d sm_100a kernels (host error plumbing + small device utilities).
#pragma once
#include <cstring>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define B3D_OK 0
#define B3D_ERR_ARG -1
#define B3D_ERR_CUDA -2
#define B3D_ERR_UNSUPPORTED -3

void b3d_set_error(const char* fmt, ...);

#define B3D_CHECK_ARG(cond, ...)            \
  do {                                      \
    if (!(cond)) {                          \
      b3d_set_error(__VA_ARGS__);           \
      return B3D_ERR_ARG;                   \
    }                                       \
  } while (0)

#define B3D_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      b3d_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return B3D_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

#define B3D_LAUNCH_CHECK()                                                          \
  do {                                                                              \
    cudaError_t e__ = cudaGetLastError();                                           \
    if (e__ != cudaSuccess) {                                                       \
      b3d_set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return B3D_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

int b3d_num_sms();

typedef __nv_bfloat16 bf16;

// ---- programmatic dependent launch (PDL): a kernel launched through b3d_launch_pdl may start while the previous kernel
// in the stream is still draining; its blocks run their prologue and then block in b3d_pdl_sync() until that kernel has
// completed and flushed.  Hides the launch latency between the ~1300 short, dependent launches of a step (the fork /
// join structure survives stream capture as programmatic graph edges).  B3D_NO_PDL=1 in the environment turns it off.
// Rule for kernels launched this way: no global-memory access before b3d_pdl_sync().
__device__ __forceinline__ void b3d_pdl_sync() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
#ifdef B3D_PDL_EARLY_TRIGGER
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
bool b3d_pdl_enabled();
// cluster_x > 1: thread-block clusters of that many CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
inline cudaError_t b3d_launch_cluster_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x,
                                          Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (b3d_pdl_enabled()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster_x > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = (unsigned)cluster_x; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = at; cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t b3d_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  return b3d_launch_cluster_pdl(kernel, grid, block, smem, st, 1, static_cast<Args&&>(args)...);
}

__device__ __forceinline__ float b3d_silu(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// d/dx silu(x) = s + x*s*(1-s), s = sigmoid(x)
// One MUFU op (tanh.approx, |rel err| <= 2^-11 on tanh => |abs err| <= 2.5e-4 on the sigmoid): the derivative multiplies
// a bf16 gradient (2^-9 relative), so this is below the storage rounding; the forward activation keeps ex2 + rcp because
// its small negative tail would see that absolute error as a large relative one.
__device__ __forceinline__ float b3d_silu_grad(float x) {
#ifdef B3D_SILU_GRAD_EXACT
  float s = __fdividef(1.0f, 1.0f + __expf(-x));
#else
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  const float s = fmaf(0.5f, t, 0.5f);
#endif
  return s * (1.0f + x * (1.0f - s));
}

__device__ __forceinline__ float b3d_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Counter-based dropout bits, recomputable in backward: one 32-bit hash per PAIR of consecutive elements (16 bits
// each; element 2k takes the low half).  key = seed * golden ratio + (element index >> 1).  An element is dropped when
// its 16-bit field is below b3d_drop_thr16(p).  (tests/emu_ops.py mirrors both functions bit for bit.)
// stride of the device-side step counter inside a dropout seed (2^40 + 435): host seeds (block index, rank) stay far below it
#define B3D_SEED_STEP_STRIDE 0x100000001B3ULL
__host__ __device__ __forceinline__ uint32_t b3d_drop_pair(uint64_t key) {
  uint32_t x = (uint32_t)key ^ ((uint32_t)(key >> 32) * 0x9E3779B9u);
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t b3d_drop_thr16(float p) { return (uint32_t)(p * 65536.0f + 0.5f); }

// 8 bf16 values moved as ONE 128-bit word (a struct of four __nv_bfloat162 is copied member-wise and the compiler
// then emits four 32-bit loads/stores instead of one LDG.128/STG.128 -- measured 4x the L1 sector requests).
struct alignas(16) bf16x8 { uint4 u; };

__device__ __forceinline__ void b3d_unpack8(const bf16x8& p, float* f) {
  const uint32_t w[4] = {p.u.x, p.u.y, p.u.z, p.u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t b3d_pack2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ bf16x8 b3d_pack8(const float* f) {
  bf16x8 p;
  p.u.x = b3d_pack2(f[0], f[1]); p.u.y = b3d_pack2(f[2], f[3]); p.u.z = b3d_pack2(f[4], f[5]); p.u.w = b3d_pack2(f[6], f[7]);
  return p;
}
