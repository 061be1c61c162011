// elementwise.cu — HBM-bound glue kernels of the STDiT3 step (vectorised 16-byte accesses,
// warp-shuffle reductions, one warp per token row).  See include/mdv2.h for the contracts.
#include <stdarg.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

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
// LayerNorm -> modulate.  One warp per row, the row lives in registers (VPL 16-byte vectors/lane).
// A CTA owns LN_ROWS consecutive rows (warp w: rows w, w+8, ...).  The shift / (1+scale) rows of the
// modulation table that its first row maps to are staged in shared memory once (every row of a (b,t)
// frame shares them; fetching them per row from L2 cost 4x the activation bytes), rows of another
// frame (a CTA straddling a frame boundary) read the table from global memory.  The next row's
// vectors are requested before the current row is reduced, so every warp keeps two rows in flight.
// ------------------------------------------------------------------------------------------------
constexpr int LN_WARPS = 8, LN_ROWS = 64;

template <typename T, int VPL>
__global__ void __launch_bounds__(LN_WARPS * 32, VPL <= 5 ? 3 : 1) ln_modulate_kernel(const T* __restrict__ x, T* __restrict__ out,
                                                                   const float* __restrict__ mod, uint32_t n_rows,
                                                                   int C, int K, int k_shift, int k_scale,
                                                                   uint32_t Tn, uint32_t S, uint32_t NC, float eps) {
  constexpr int VN = Vec16<T>::N;
  extern __shared__ float s_tab[];                       // [0,C): shift, [C,2C): 1 + scale
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t row0 = blockIdx.x * LN_ROWS;
  const uint32_t row_end = min(row0 + (uint32_t)LN_ROWS, n_rows);
  const int nvec = C / VN;
  auto table_row = [&](uint32_t r) {
    const uint32_t g = r / S;
    return (g / Tn / NC) * Tn + g % Tn;
  };
  const uint32_t tab0 = table_row(row0);
  {
    const float* mrow = mod + (int64_t)tab0 * K * C;
    const float4* sh4 = reinterpret_cast<const float4*>(mrow + (int64_t)k_shift * C);
    const float4* sc4 = reinterpret_cast<const float4*>(mrow + (int64_t)k_scale * C);
    for (int i = threadIdx.x; i < C / 4; i += LN_WARPS * 32) {
      reinterpret_cast<float4*>(s_tab)[i] = sh4[i];
      float4 sc = sc4[i];
      sc.x += 1.f; sc.y += 1.f; sc.z += 1.f; sc.w += 1.f;
      reinterpret_cast<float4*>(s_tab + C)[i] = sc;
    }
  }
  Vec16<T> cur[VPL], nxt[VPL];
  auto load_row = [&](uint32_t r, Vec16<T>* dst) {
    const Vec16<T>* xr = reinterpret_cast<const Vec16<T>*>(x + (int64_t)r * C);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) dst[i] = xr[vi];
    }
  };
  uint32_t row = row0 + warp;
  if (row < row_end) load_row(row, nxt);
  __syncthreads();
  for (; row < row_end; row += LN_WARPS) {
#pragma unroll
    for (int i = 0; i < VPL; ++i) cur[i] = nxt[i];
    if (row + LN_WARPS < row_end) load_row(row + LN_WARPS, nxt);
    // the row stays PACKED in registers (cur) and is unpacked once per pass: three cheap unpacks instead of
    // 8 live fp32 values per vector keep the kernel at 4 CTAs/SM
    // packed fp32 pairs (FADD2 / FFMA2 / FMUL2: two results per issue slot): inside the power-capped step the SM clock
    // drops to ~1.45 GHz and the ~10 scalar instructions per element of the three passes, not HBM, set this kernel's pace
    using sm100::fadd2;
    using sm100::ffma2;
    using sm100::fmul2;
    float2 sum2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      if (lane + 32 * i < nvec) {
        float v[VN];
        cur[i].unpack(v);
#pragma unroll
        for (int j = 0; j < VN; j += 2) sum2 = fadd2(sum2, make_float2(v[j], v[j + 1]));
      }
    }
    const float mean = warp_sum(sum2.x + sum2.y) / C;
    const float2 nmean2 = make_float2(-mean, -mean);
    float2 sq2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      if (lane + 32 * i < nvec) {
        float v[VN];
        cur[i].unpack(v);
#pragma unroll
        for (int j = 0; j < VN; j += 2) {
          const float2 dlt = fadd2(make_float2(v[j], v[j + 1]), nmean2);
          sq2 = ffma2(dlt, dlt, sq2);
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(sq2.x + sq2.y) / C + eps);
    const float2 rstd2 = make_float2(rstd, rstd);
    const uint32_t tab = table_row(row);
    const bool staged = tab == tab0;
    const float* mrow = mod + (int64_t)tab * K * C;
    const float* gsh = mrow + (int64_t)k_shift * C;
    const float* gsc = mrow + (int64_t)k_scale * C;
    Vec16<T>* orow = reinterpret_cast<Vec16<T>*>(out + (int64_t)row * C);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float o[VN], v[VN];
        cur[i].unpack(v);
#pragma unroll
        for (int j = 0; j < VN; j += 4) {
          float4 s4, h4;
          if (staged) {
            h4 = *reinterpret_cast<const float4*>(s_tab + vi * VN + j);
            s4 = *reinterpret_cast<const float4*>(s_tab + C + vi * VN + j);
          } else {
            h4 = *reinterpret_cast<const float4*>(gsh + vi * VN + j);
            s4 = *reinterpret_cast<const float4*>(gsc + vi * VN + j);
            s4.x += 1.f; s4.y += 1.f; s4.z += 1.f; s4.w += 1.f;
          }
          const float2 o01 = ffma2(fmul2(fadd2(make_float2(v[j + 0], v[j + 1]), nmean2), rstd2),
                                   make_float2(s4.x, s4.y), make_float2(h4.x, h4.y));
          const float2 o23 = ffma2(fmul2(fadd2(make_float2(v[j + 2], v[j + 3]), nmean2), rstd2),
                                   make_float2(s4.z, s4.w), make_float2(h4.z, h4.w));
          o[j + 0] = o01.x;
          o[j + 1] = o01.y;
          o[j + 2] = o23.x;
          o[j + 3] = o23.y;
        }
        Vec16<T> t;
        t.pack(o);
        orow[vi] = t;
      }
    }
  }
}

template <typename T>
static int launch_ln_modulate(const void* x, void* out, const float* mod, int64_t n_rows, int C, int K, int ks,
                              int kc, int T_, int S, int NC, float eps, cudaStream_t st) {
  constexpr int VN = Vec16<T>::N;
  MDV2_REQUIRE(C % VN == 0, "ln_modulate: C=%d must be a multiple of %d", C, VN);
  MDV2_REQUIRE(n_rows < (1ll << 31), "ln_modulate: too many rows");
  MDV2_REQUIRE(((uintptr_t)x % 16 == 0) && ((uintptr_t)out % 16 == 0) && ((uintptr_t)mod % 16 == 0),
               "ln_modulate: operands must be 16-byte aligned");
  const int nvec = C / VN;
  const int vpl = (nvec + 31) / 32;
  const size_t smem = 2 * (size_t)C * sizeof(float);
  MDV2_REQUIRE(smem <= 48 * 1024, "ln_modulate: C=%d too large", C);
  dim3 grid((unsigned)((n_rows + LN_ROWS - 1) / LN_ROWS)), block(LN_WARPS * 32);
#define LN_CASE(V)                                                                                             \
  ln_modulate_kernel<T, V><<<grid, block, smem, st>>>((const T*)x, (T*)out, mod, (uint32_t)n_rows, C, K, ks, kc, \
                                                      (uint32_t)T_, (uint32_t)S, (uint32_t)NC, eps)
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


// ------------------------------------------------------------------------------------------------
// adaLN modulation tables of ALL blocks in one pass:
//   out[n][bt][k][c] (fp32) = round_T(tables[n][k][c] + tsel[bt][k][c or 0])
// tables: the stacked scale_shift_table parameters [n, K, C]; tsel: the (x_mask-selected) timestep vector per (b, t)
// frame, [BT, K, Ct] with Ct = C (6-row tables) or 1 (the scalar modulation of the cross-view branch).  The sum is
// rounded to the model dtype like the reference's own add (bf16 + bf16 -> bf16) and widened for the kernels that read it.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void adaln_tables_kernel(const T* __restrict__ tables, const T* __restrict__ tsel, float* __restrict__ out,
                                    int64_t total4, int BT, int KC4, int C4, int Ct) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int kc = (int)(i % KC4);                        // (k, c/4)
    const int64_t r = i / KC4;
    const int bt = (int)(r % BT);
    const int64_t n = r / BT;
    const T* tp = tables + (n * KC4 + kc) * 4;
    float tv[4];
    if (Ct == 1) {
      const float v = to_f(tsel[(int64_t)bt * (KC4 / C4) + kc / C4]);
      tv[0] = tv[1] = tv[2] = tv[3] = v;
    } else {
      const T* sp = tsel + ((int64_t)bt * KC4 + kc) * 4;
#pragma unroll
      for (int e = 0; e < 4; ++e) tv[e] = to_f(sp[e]);
    }
    float4 o;
    o.x = to_f(from_f<T>(to_f(tp[0]) + tv[0]));
    o.y = to_f(from_f<T>(to_f(tp[1]) + tv[1]));
    o.z = to_f(from_f<T>(to_f(tp[2]) + tv[2]));
    o.w = to_f(from_f<T>(to_f(tp[3]) + tv[3]));
    reinterpret_cast<float4*>(out)[i] = o;
  }
}

extern "C" int mdv2_adaln_tables(const void* tables, const void* tsel, float* out, int n, int BT, int K, int C, int Ct,
                                 int dtype, void* stream) {
  if ((int64_t)n * BT * K * C == 0) return MDV2_OK;
  MDV2_REQUIRE(n > 0 && BT > 0 && K > 0 && C > 0 && C % 4 == 0, "adaln_tables: bad shape n=%d BT=%d K=%d C=%d", n, BT, K, C);
  MDV2_REQUIRE(Ct == C || Ct == 1, "adaln_tables: the timestep vector is [BT, K, C] or [BT, K, 1]");
  MDV2_REQUIRE(dtype == 0 || dtype == 1, "adaln_tables: bad dtype %d", dtype);
  MDV2_REQUIRE((uintptr_t)out % 16 == 0, "adaln_tables: output must be 16-byte aligned");
  const int64_t total4 = (int64_t)n * BT * K * C / 4;
  const int threads = 256;
  int64_t blocks = (total4 + threads - 1) / threads;
  const int64_t cap = (int64_t)mdv2_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0)
    adaln_tables_kernel<float><<<(unsigned)blocks, threads, 0, st>>>((const float*)tables, (const float*)tsel, out, total4, BT,
                                                                    K * C / 4, C / 4, Ct);
  else
    adaln_tables_kernel<bf16><<<(unsigned)blocks, threads, 0, st>>>((const bf16*)tables, (const bf16*)tsel, out, total4, BT,
                                                                   K * C / 4, C / 4, Ct);
  MDV2_CHECK_LAUNCH("adaln_tables");
  return MDV2_OK;
}
