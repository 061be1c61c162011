// common.cuh — shared helpers for libmdv2 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mdv2.h"

#define MDV2_VERSION 100

void mdv2_set_error(const char* fmt, ...);
void mdv2_note_kernel(const char* name);   // name of the kernel family the last C-ABI call launched

#define MDV2_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      mdv2_set_error(__VA_ARGS__);         \
      return MDV2_EINVAL;                  \
    }                                      \
  } while (0)

#define MDV2_CHECK_LAUNCH(name)                                              \
  do {                                                                       \
    cudaError_t e_ = cudaGetLastError();                                     \
    if (e_ != cudaSuccess) {                                                 \
      mdv2_set_error("%s: launch failed: %s", name, cudaGetErrorString(e_)); \
      return MDV2_ECUDA;                                                     \
    }                                                                        \
    mdv2_note_kernel(name);                                                  \
  } while (0)

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// 16-byte vector of T: 4 floats or 8 bf16
template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  float4 raw;
  __device__ __forceinline__ void unpack(float* f) const { f[0] = raw.x; f[1] = raw.y; f[2] = raw.z; f[3] = raw.w; }
  __device__ __forceinline__ void pack(const float* f) { raw = make_float4(f[0], f[1], f[2], f[3]); }
};
template <> struct Vec16<bf16> {
  static constexpr int N = 8;
  uint4 raw;
  __device__ __forceinline__ void unpack(float* f) const {
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __bfloat1622float2(p[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  __device__ __forceinline__ void pack(const float* f) {
    __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float gelu_tanh(float x) {
  // nn.GELU(approximate="tanh"): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}

__device__ __forceinline__ float gelu_tanh_fast(float x) {   // MUFU.TANH (rel. error ~2^-11): bf16 outputs only
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x), t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.f + t);
}

// row r (β,t,s) -> row of the [b·T, K, C] modulation table
__device__ __forceinline__ int64_t mod_row(int64_t r, int T, int S, int NC) {
  int64_t g = r / S;
  return (g / T / NC) * T + g % T;
}

static inline int mdv2_num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// cudaFuncSetAttribute (dynamic shared memory opt-in) is per DEVICE: one flag per (call site, device), so a process that
// drives several GPUs configures each kernel on each of them
struct PerDeviceFlag {
  bool done[64] = {};
  bool& here() {
    int d = 0;
    cudaGetDevice(&d);
    return done[d & 63];
  }
};

// output scatter of the tcgen05 attention kernel under sequence parallelism (see AParams in attn_sm100.cu)
struct AttnScatter {
  void* const* dst;
  int n_dst, rows_per_rank, dst_col;
  int64_t dst_ld;
};
