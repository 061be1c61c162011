// pipe_probe.cu — issue cost (clk per warp instruction per SM sub-partition) of the instructions in the attention
// softmax loop on sm_100a: one CTA of W warps per sub-partition x 4, each warp runs a long unrolled chain of
// independent instructions of one kind.  clk/inst = elapsed * (4 sub-partitions) / (total warp instructions) ... printed
// per sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

template <int KIND>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, long long* clk) {
  float a[8], b = threadIdx.x * 1e-3f + 1.f, c = 0.5f;
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i * 0.25f;
  uint32_t u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = threadIdx.x * 3 + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        if (KIND == 0) {          // FFMA x2 (scalar)
          a[i] = fmaf(a[i], b, c); a[i + 1] = fmaf(a[i + 1], b, c);
        } else if (KIND == 1) {   // FFMA2
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*reinterpret_cast<uint64_t*>(&a[i])) : "l"(*reinterpret_cast<uint64_t*>(&a[(i + 2) & 7])), "l"(*reinterpret_cast<uint64_t*>(&a[(i + 4) & 7])));
        } else if (KIND == 2) {   // MUFU.EX2 x2
          asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
          asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i + 1]));
        } else if (KIND == 3) {   // FMNMX3
          asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(a[i + 1]), "f"(b));
          asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i + 1]) : "f"(a[i]), "f"(c));
        } else if (KIND == 4) {   // F2FP pack
          asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(a[i]), "f"(a[i + 1]));
          asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u[i + 1]) : "f"(a[i + 1]), "f"(a[i]));
          a[i] = __uint_as_float(u[i]); a[i + 1] = __uint_as_float(u[i + 1]);
        } else if (KIND == 5) {   // FADD2
          asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(*reinterpret_cast<uint64_t*>(&a[i])) : "l"(*reinterpret_cast<uint64_t*>(&a[(i + 2) & 7])));
        } else if (KIND == 6) {   // IMAD (shift-add) x2
          u[i] = u[i] * 0x800000u + u[i + 1]; u[i + 1] = u[i + 1] * 0x800000u + u[i];
        } else if (KIND == 7) {   // FADD scalar x2
          a[i] += b; a[i + 1] += c;
        }
      }
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 1 << 20); cudaMallocManaged(&clk, 8);
  const char* names[] = {"FFMA (scalar) x2", "FFMA2", "MUFU.EX2 x2", "FMNMX3 x2", "F2FP.BF16 pack x2", "FADD2", "IMAD x2", "FADD (scalar) x2"};
  const int per_iter[] = {64, 32, 64, 64, 64, 32, 64, 64};   // warp instructions of the kind per iteration
  const int iters = 2000;
  for (int W : {4, 8, 16})
    for (int kind = 0; kind < 8; ++kind) {
      for (int rep = 0; rep < 2; ++rep) {
        switch (kind) {
          case 0: k<0><<<1, W * 32>>>(out, iters, clk); break; case 1: k<1><<<1, W * 32>>>(out, iters, clk); break;
          case 2: k<2><<<1, W * 32>>>(out, iters, clk); break; case 3: k<3><<<1, W * 32>>>(out, iters, clk); break;
          case 4: k<4><<<1, W * 32>>>(out, iters, clk); break; case 5: k<5><<<1, W * 32>>>(out, iters, clk); break;
          case 6: k<6><<<1, W * 32>>>(out, iters, clk); break; default: k<7><<<1, W * 32>>>(out, iters, clk); break;
        }
        cudaDeviceSynchronize();
      }
      const double inst_per_smsp = (double)per_iter[kind] * iters * (W / 4.0);
      printf("W=%2d warps/SM (%d per sub-partition)  %-20s %6.2f clk per warp-instruction per sub-partition\n", W, W / 4,
             names[kind], (double)*clk / inst_per_smsp);
    }
  return 0;
}
