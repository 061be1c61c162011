// elementwise.cu — HBM-bound glue kernels of the STDiT3 step (vectorised 16-byte accesses,
// warp-shuffle reductions, one warp per token row).  See include/mdv2.h for the contracts.
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[512] = "";
void mdv2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* mdv2_last_error(void) { return g_err; }
static thread_local const char* g_kernel = "";
void mdv2_note_kernel(const char* name) { g_kernel = name; }
extern "C" const char* mdv2_last_kernel(void) { return g_kernel; }
extern "C" int mdv2_version(void) { return MDV2_VERSION; }

// ------------------------------------------------------------------------------------------------
// LayerNorm -> modulate.  One warp per row; the row lives in registers (VPL 16-byte vectors/lane).
// ------------------------------------------------------------------------------------------------
template <typename T, int VPL>
__global__ void __launch_bounds__(256) ln_modulate_kernel(const T* __restrict__ x, T* __restrict__ out,
                                                          const float* __restrict__ mod, int64_t n_rows,
                                                          int C, int K, int k_shift, int k_scale, int Tn,
                                                          int S, int NC, float eps) {
  constexpr int VN = Vec16<T>::N;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  const int nvec = C / VN;
  const Vec16<T>* xr = reinterpret_cast<const Vec16<T>*>(x + row * C);
  float v[VPL][VN];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + 32 * i;
    if (vi < nvec) {
      Vec16<T> t = xr[vi];
      t.unpack(v[i]);
#pragma unroll
      for (int j = 0; j < VN; ++j) sum += v[i][j];
    }
  }
  const float mean = warp_sum(sum) / C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    if (lane + 32 * i < nvec) {
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        float dlt = v[i][j] - mean;
        sq += dlt * dlt;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / C + eps);
  const float* mrow = mod + mod_row(row, Tn, S, NC) * (int64_t)K * C;
  const float* sh = mrow + (int64_t)k_shift * C;
  const float* sc = mrow + (int64_t)k_scale * C;
  Vec16<T>* orow = reinterpret_cast<Vec16<T>*>(out + row * C);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + 32 * i;
    if (vi < nvec) {
      float o[VN];
#pragma unroll
      for (int j = 0; j < VN; j += 4) {
        float4 s4 = *reinterpret_cast<const float4*>(sc + vi * VN + j);
        float4 h4 = *reinterpret_cast<const float4*>(sh + vi * VN + j);
        o[j + 0] = (v[i][j + 0] - mean) * rstd * (1.f + s4.x) + h4.x;
        o[j + 1] = (v[i][j + 1] - mean) * rstd * (1.f + s4.y) + h4.y;
        o[j + 2] = (v[i][j + 2] - mean) * rstd * (1.f + s4.z) + h4.z;
        o[j + 3] = (v[i][j + 3] - mean) * rstd * (1.f + s4.w) + h4.w;
      }
      Vec16<T> t;
      t.pack(o);
      orow[vi] = t;
    }
  }
}

template <typename T>
static int launch_ln_modulate(const void* x, void* out, const float* mod, int64_t n_rows, int C, int K,
                              int ks, int kc, int T_, int S, int NC, float eps, cudaStream_t st) {
  constexpr int VN = Vec16<T>::N;
  MDV2_REQUIRE(C % VN == 0, "ln_modulate: C=%d must be a multiple of %d", C, VN);
  const int nvec = C / VN;
  const int vpl = (nvec + 31) / 32;
  const int wpb = 8;
  dim3 grid((unsigned)((n_rows + wpb - 1) / wpb)), block(wpb * 32);
#define LN_CASE(V)                                                                                    \
  ln_modulate_kernel<T, V><<<grid, block, 0, st>>>((const T*)x, (T*)out, mod, n_rows, C, K, ks, kc, T_, \
                                                   S, NC, eps)
  if (vpl <= 2) LN_CASE(2);
  else if (vpl <= 5) LN_CASE(5);
  else if (vpl <= 10) LN_CASE(10);
  else MDV2_REQUIRE(false, "ln_modulate: C=%d too large", C);
#undef LN_CASE
  MDV2_CHECK_LAUNCH("ln_modulate");
  return MDV2_OK;
}

extern "C" int mdv2_ln_modulate(const void* x, void* out, const float* mod, int64_t n_rows, int C, int K,
                                int k_shift, int k_scale, int T, int S, int NC, float eps, int dtype,
                                void* stream) {
  if (n_rows == 0) return MDV2_OK;
  MDV2_REQUIRE(k_shift >= 0 && k_shift < K && k_scale >= 0 && k_scale < K, "ln_modulate: bad table index");
  MDV2_REQUIRE(T > 0 && S > 0 && NC > 0, "ln_modulate: bad geometry");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) return launch_ln_modulate<float>(x, out, mod, n_rows, C, K, k_shift, k_scale, T, S, NC, eps, st);
  if (dtype == 1) return launch_ln_modulate<bf16>(x, out, mod, n_rows, C, K, k_shift, k_scale, T, S, NC, eps, st);
  MDV2_REQUIRE(false, "ln_modulate: bad dtype %d", dtype);
}

// ------------------------------------------------------------------------------------------------
// per-head RMSNorm of the q and k column blocks, in place (SIMT-GEMM path; the tcgen05 GEMM fuses it)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void qk_rmsnorm_kernel(T* __restrict__ qkv, int64_t ld, int64_t M, int D, int n_q, int n_k,
                                  const T* __restrict__ wq, const T* __restrict__ wk, float eps) {
  const int lane = threadIdx.x & 31;
  const int heads = (n_q + n_k) / D;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (wid >= M * heads) return;
  const int64_t row = wid / heads;
  const int h = (int)(wid % heads);
  T* p = qkv + row * ld + (int64_t)h * D;
  const T* w = (h * D < n_q) ? wq : wk;
  float ss = 0.f;
  for (int d = lane; d < D; d += 32) {
    float f = to_f(p[d]);
    ss += f * f;
  }
  const float r = rsqrtf(warp_sum(ss) / D + eps);
  for (int d = lane; d < D; d += 32) {
    // LlamaRMSNorm: weight * (x * rstd).to(input_dtype)
    T normed = from_f<T>(to_f(p[d]) * r);
    p[d] = from_f<T>(to_f(w[d]) * to_f(normed));
  }
}

int mdv2_qk_rmsnorm(void* qkv, int64_t ld, int64_t M, int D, int n_q, int n_k, const void* wq,
                    const void* wk, float eps, int dtype, cudaStream_t st) {
  if (M == 0) return MDV2_OK;
  MDV2_REQUIRE(n_q % D == 0 && n_k % D == 0, "qk_rmsnorm: column blocks must be whole heads");
  const int64_t warps = M * ((n_q + n_k) / D);
  dim3 grid((unsigned)((warps + 7) / 8)), block(256);
  if (dtype == 0)
    qk_rmsnorm_kernel<float><<<grid, block, 0, st>>>((float*)qkv, ld, M, D, n_q, n_k, (const float*)wq, (const float*)wk, eps);
  else
    qk_rmsnorm_kernel<bf16><<<grid, block, 0, st>>>((bf16*)qkv, ld, M, D, n_q, n_k, (const bf16*)wq, (const bf16*)wk, eps);
  MDV2_CHECK_LAUNCH("qk_rmsnorm");
  return MDV2_OK;
}
