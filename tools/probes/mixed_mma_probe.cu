// mixed_mma_probe.cu — (1) may a kind::f16 tcgen05.mma take A as fp16 (from TMEM) and B as bf16 (shared memory)?
// (2) issue cost of ex2.approx.ftz.f16x2 / bf16x2 and cvt.rn.f16x2.f32 next to the f32 forms.
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"
using namespace sm100;

__global__ void __launch_bounds__(128, 1) mixed(float* out, int a_fmt, int b_fmt, uint32_t a_bits, uint16_t b_bits) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 16384 / 2; i += 128) reinterpret_cast<uint16_t*>(smem)[i] = b_bits;
  fence_async_smem();
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
  uint32_t av[8];
  for (int i = 0; i < 8; ++i) av[i] = a_bits;             // 16 K-elements (pairs) of A for this row
  tmem_st_x8(lane_addr + 128, av);
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) {
    // idesc: c_format f32 (1<<4), a_format at bit 7, b_format at bit 10, N>>3 at 17, M>>4 at 24
    const uint32_t idesc = (1u << 4) | ((uint32_t)a_fmt << 7) | ((uint32_t)b_fmt << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t db = make_desc_k_sw128(smem_u32(smem));
    if (elect_one()) {
      umma_bf16_ts(tmem, tmem + 128, db, idesc, 0);
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t r[8];
  tmem_ld_x8(lane_addr, r);
  tmem_ld_wait();
  if (threadIdx.x == 5) { out[0] = __uint_as_float(r[0]); out[1] = __uint_as_float(r[7]); }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

template <int KIND>
__global__ void __launch_bounds__(256, 1) rate(float* out, int iters, long long* clk) {
  uint32_t u[8]; float a[8];
  for (int i = 0; i < 8; ++i) { u[i] = 0x3c003c00u + threadIdx.x + i; a[i] = -(float)(threadIdx.x % 7) - i * 0.1f; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (KIND == 0) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u[i]));
        else if (KIND == 1) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(u[i]));
        else if (KIND == 2) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(a[i]), "f"(a[(i + 1) & 7]));
        else asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(u[i]);
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) *clk = t1 - t0;
}

int main() {
  float* out; long long* clk;
  cudaMallocManaged(&out, 4096); cudaMallocManaged(&clk, 8);
  cudaFuncSetAttribute(mixed, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * 1024);
  const char* names[] = {"ex2.approx.ftz.f16x2", "ex2.approx.ftz.bf16x2", "cvt.rn.f16x2.f32", "ex2.approx.ftz.f32"};
  const int iters = 2000;
  for (int kind = 0; kind < 4; ++kind) {
    for (int rep = 0; rep < 2; ++rep) {
      switch (kind) { case 0: rate<0><<<1, 256>>>(out, iters, clk); break; case 1: rate<1><<<1, 256>>>(out, iters, clk); break;
                      case 2: rate<2><<<1, 256>>>(out, iters, clk); break; default: rate<3><<<1, 256>>>(out, iters, clk); }
      cudaDeviceSynchronize();
    }
    printf("%-24s %6.2f clk per warp-instruction per sub-partition (2 warps each)\n", names[kind], (double)*clk / (64.0 * iters * 2));
  }
  struct { const char* name; int af, bf; uint32_t ab; uint16_t bb; float expect; } cases[] = {
      {"A bf16 x B bf16 (1 x 1)", 1, 1, 0x3f803f80u, 0x3f80, 16.f},
      {"A f16  x B f16  (1 x 1)", 0, 0, 0x3c003c00u, 0x3c00, 16.f},
      {"A f16  x B bf16 (1 x 1)", 0, 1, 0x3c003c00u, 0x3f80, 16.f},
      {"A f16  x B bf16 (0.5 x 3)", 0, 1, 0x38003800u, 0x4040, 24.f},
  };
  for (auto& c : cases) {
    out[0] = out[1] = -1.f;
    mixed<<<1, 128, 32 * 1024>>>(out, c.af, c.bf, c.ab, c.bb);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%-28s -> D = %g, %g (expect %g)  %s\n", c.name, out[0], out[1], c.expect, e == cudaSuccess ? "" : cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
