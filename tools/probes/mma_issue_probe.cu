// mma_issue_probe.cu — how fast can tcgen05.mma instructions be ISSUED?  One CTA per SM, W issuing warps, each
// issues `n` MMAs (128 x N x 16, bf16, operands = whatever is in shared memory / TMEM) back to back and commits.
// Prints per-instruction issue interval and total time to completion, for SS and TS (A from TMEM) forms.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../magicdrive_v2_b200/csrc mma_issue_probe.cu -o mma_issue_probe
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"
using namespace sm100;

template <int MODE>   // 0: loop + elect per MMA, 1: unrolled x16 + elect per MMA, 2: lane 0 branch, unrolled x16, 3: one elect around 16
__global__ void __launch_bounds__(128, 1) probe(int N, int n, int W, int ts, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar[4];   // one per issuing warp
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp < W) {
    const uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
    const uint64_t da = make_desc_k_sw128(smem_u32(smem));
    const uint64_t db = make_desc_k_sw128(smem_u32(smem + 32768));
    const uint32_t d = tmem + warp * 128;   // N <= 128 with 4 warps
    __syncwarp();
    long long t0 = 0, t1 = 0, t2 = 0;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0)::"memory");
    auto one = [&](int i) {
      if (ts) umma_bf16_ts(d, tmem + 480, db + (uint64_t)((i & 3) * 2), idesc, 1);
      else umma_bf16_ss(d, da + (uint64_t)((i & 3) * 2), db + (uint64_t)((i & 3) * 2), idesc, 1);
    };
    if constexpr (MODE == 0) {
      for (int i = 0; i < n; ++i)
        if (elect_one()) one(i);
    } else if constexpr (MODE == 1) {
      for (int b = 0; b < n; b += 16) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (elect_one()) one(i);
      }
    } else if constexpr (MODE == 2) {
      if (lane == 0) {
        for (int b = 0; b < n; b += 16) {
#pragma unroll
          for (int i = 0; i < 16; ++i) one(i);
        }
      }
      __syncwarp();
    } else {
      for (int b = 0; b < n; b += 16) {
        if (elect_one()) {
#pragma unroll
          for (int i = 0; i < 16; ++i) one(i);
        }
        __syncwarp();
      }
    }
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1)::"memory");
    if (elect_one()) umma_commit(&bar[warp]);
    __syncwarp();
    mbar_wait(&bar[warp], 0);
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t2)::"memory");
    if (lane == 0 && blockIdx.x == 0) {
      out[warp * 2] = t1 - t0;
      out[warp * 2 + 1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 64);
  const int n = 64;
  for (int mode = 0; mode < 4; ++mode)
    for (int ts = 0; ts < 2; ++ts)
      for (int W : {1, 2, 4})
        for (int N : {16, 64, 80, 128}) {
          for (int rep = 0; rep < 2; ++rep) {
            switch (mode) {
              case 0: cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); probe<0><<<1, 128, 100 * 1024>>>(N, n, W, ts, out); break;
              case 1: cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); probe<1><<<1, 128, 100 * 1024>>>(N, n, W, ts, out); break;
              case 2: cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); probe<2><<<1, 128, 100 * 1024>>>(N, n, W, ts, out); break;
              default: cudaFuncSetAttribute(probe<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); probe<3><<<1, 128, 100 * 1024>>>(N, n, W, ts, out); break;
            }
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          }
          printf("mode %d %s W=%d N=%3d: issue %6.1f clk/MMA, complete %6.1f (math %5.1f)", mode, ts ? "TS" : "SS", W, N,
                 (double)out[0] / n, (double)out[1] / n, 128.0 * N / 256);
          if (W >= 2) printf(" | warp1: issue %6.1f, complete %6.1f", (double)out[2] / n, (double)out[3] / n);
          printf("\n");
        }
  return 0;
}
