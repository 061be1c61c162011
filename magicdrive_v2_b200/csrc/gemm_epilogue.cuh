// gemm_epilogue.cuh — the fused epilogue shared by the SIMT and the tcgen05 GEMM:
//   v = acc + bias_scale·bias[c] ; [GELU-tanh] ; [· gate(row,c)] ; [+ residual(row', c)]
// (per-head RMSNorm of q/k is applied row-wise by the caller before apply(): see gemm_sm100.cu;
// the SIMT path runs it as a separate in-place pass.)
#pragma once
#include "common.cuh"

struct RowCtx {
  const float* gate;  // row of the fp32 modulation table, or nullptr
  int64_t res_row;
};

template <typename T>
struct Epilogue {
  const T* bias;        // [N] or nullptr
  float bias_scale;
  int act;              // 0 none, 1 gelu-tanh
  const float* gate_mod;  // [b·T, mod_K, N] fp32 or nullptr
  int mod_K, k_gate;
  const T* res;         // residual or nullptr (may alias the output)
  int64_t ldr, res_mod, res_div;
  int T_, S, NC, N;

  __device__ __forceinline__ RowCtx row_ctx(int64_t r) const {
    RowCtx rc;
    rc.gate = gate_mod ? gate_mod + (mod_row(r, T_, S, NC) * mod_K + k_gate) * (int64_t)N : nullptr;
    if (res_mod > 0)
      rc.res_row = (res_div > 0 ? (r / (res_mod * res_div)) * res_mod : 0) + r % res_mod;
    else
      rc.res_row = r;
    return rc;
  }
  __device__ __forceinline__ float apply(float acc, int64_t r, int c, const RowCtx& rc) const {
    float v = acc;
    if (bias) v += bias_scale * to_f(bias[c]);
    if (act == 1) v = gelu_tanh(v);
    if (rc.gate) v *= rc.gate[c];
    if (res) v += to_f(res[rc.res_row * ldr + c]);
    return v;
  }
};

template <typename T>
int mdv2_gemm_simt(const void* a, int64_t lda, const void* w, int64_t ldw, void* out, int64_t ldo, int64_t M,
                   int N, int K, const Epilogue<T>& ep, cudaStream_t st);
int mdv2_qk_rmsnorm(void* qkv, int64_t ld, int64_t M, int D, int n_q, int n_k, const void* wq, const void* wk,
                    float eps, int dtype, cudaStream_t st);
// tcgen05 path (gemm_sm100.cu); returns MDV2_EINVAL without launching when the shape is not tiled
struct RmsArgs {
  const void* wq;
  const void* wk;
  int head_dim, n_q, n_k;
  float eps;
};
bool mdv2_gemm_sm100_supported(int64_t M, int N, int K, int64_t lda, int64_t ldw, int64_t ldo, const RmsArgs* rms);
// scattered qkv output (sequence parallelism): destinations = the peers' [seq][n_dst·rows_per_seq][3][W/n_dst] buffers
struct QkvScatter {
  void* const* dst;
  int n_dst, rows_per_seq, rank;
};
int mdv2_gemm_sm100(const void* a, int64_t lda, const void* w, int64_t ldw, void* out, int64_t ldo, int64_t M, int N,
                    int K, const Epilogue<bf16>& ep, const RmsArgs* rms, cudaStream_t st,
                    const QkvScatter* scatter = nullptr);
