// Shared helpers for the b3d sm_100a kernels (host error plumbing + small device utilities).
#pragma once
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

__device__ __forceinline__ float b3d_silu(float x) { return x / (1.0f + __expf(-x)); }
// d/dx silu(x) = s + x*s*(1-s), s = sigmoid(x)
__device__ __forceinline__ float b3d_silu_grad(float x) {
  float s = 1.0f / (1.0f + __expf(-x));
  return s * (1.0f + x * (1.0f - s));
}

__device__ __forceinline__ float b3d_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// counter-based uniform in [0,1): one 32-bit hash per (seed, index) -- recomputable in backward.
__device__ __forceinline__ uint32_t b3d_hash32(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return (uint32_t)x;
}
__device__ __forceinline__ float b3d_uniform(uint64_t seed, uint64_t idx) {
  return (float)(b3d_hash32(seed * 0x9E3779B97F4A7C15ULL + idx) >> 8) * (1.0f / 16777216.0f);
}

struct alignas(16) bf16x8 { __nv_bfloat162 v[4]; };

__device__ __forceinline__ void b3d_unpack8(const bf16x8& p, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(p.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ bf16x8 b3d_pack8(const float* f) {
  bf16x8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return p;
}
