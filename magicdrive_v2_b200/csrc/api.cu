// api.cu — C-ABI dispatch for the GEMM and attention entry points (see include/mdv2.h).
#include "gemm_epilogue.cuh"

int mdv2_attention_simt(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                        void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                        const int32_t* kv_map, int n_sets, int dtype, cudaStream_t st);
bool mdv2_attention_sm100_supported(int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int Lq, int Lk, int H, int D,
                                    int n_sets);
int mdv2_attention_sm100(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                         void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                         const int32_t* kv_map, int n_sets, cudaStream_t st, const AttnScatter* scatter = nullptr);

static bool device_is_sm100() {
  static int cached = -1;
  if (cached < 0) {
    int dev = 0, major = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cached = (major == 10) ? 1 : 0;
  }
  return cached == 1;
}

template <typename T>
static Epilogue<T> make_epilogue(const void* bias, float bias_scale, int act, const float* gate_mod, int mod_K,
                                 int k_gate, const void* res, int64_t ldr, int64_t res_mod, int64_t res_div, int T_,
                                 int S, int NC, int N) {
  Epilogue<T> ep;
  ep.bias = (const T*)bias;
  ep.bias_scale = bias_scale;
  ep.act = act;
  ep.gate_mod = gate_mod;
  ep.mod_K = mod_K;
  ep.k_gate = k_gate;
  ep.res = (const T*)res;
  ep.ldr = ldr;
  ep.res_mod = res_mod;
  ep.res_div = res_div;
  ep.T_ = T_ > 0 ? T_ : 1;
  ep.S = S > 0 ? S : 1;
  ep.NC = NC > 0 ? NC : 1;
  ep.N = N;
  return ep;
}

extern "C" int mdv2_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias, void* out,
                         int64_t ldo, int64_t M, int N, int K, int act, float bias_scale, const float* gate_mod,
                         int mod_K, int k_gate, const void* res, int64_t ldr, int64_t res_mod, int64_t res_div, int T,
                         int S, int NC, const void* rms_q, const void* rms_k, int head_dim, int n_q, int n_k, float eps,
                         int dtype, int backend, void* stream) {
  if (M == 0 || N == 0) return MDV2_OK;
  MDV2_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%lld N=%d K=%d", (long long)M, N, K);
  MDV2_REQUIRE(dtype == 0 || dtype == 1, "gemm: bad dtype %d", dtype);
  MDV2_REQUIRE(act == 0 || act == 1, "gemm: bad activation %d", act);
  MDV2_REQUIRE(!gate_mod || (T > 0 && S > 0 && NC > 0 && k_gate >= 0 && k_gate < mod_K), "gemm: gate needs geometry");
  const bool has_rms = rms_q != nullptr || rms_k != nullptr;
  if (has_rms) {
    MDV2_REQUIRE(rms_q && rms_k && head_dim > 0 && n_q % head_dim == 0 && n_k % head_dim == 0 && n_q + n_k <= N,
                 "gemm: bad RMSNorm column blocks");
    MDV2_REQUIRE(!gate_mod && !res && act == 0, "gemm: RMSNorm epilogue excludes gate/residual/activation");
  }
  cudaStream_t st = (cudaStream_t)stream;
  RmsArgs rms{rms_q, rms_k, head_dim, n_q, n_k, eps};
  if (dtype == 1 && backend != MDV2_BACKEND_SIMT) {
    const bool ok = device_is_sm100() && mdv2_gemm_sm100_supported(M, N, K, lda, ldw, ldo, has_rms ? &rms : nullptr);
    if (ok) {
      Epilogue<bf16> ep = make_epilogue<bf16>(bias, bias_scale, act, gate_mod, mod_K, k_gate, res, ldr, res_mod,
                                              res_div, T, S, NC, N);
      return mdv2_gemm_sm100(a, lda, w, ldw, out, ldo, M, N, K, ep, has_rms ? &rms : nullptr, st);
    }
    MDV2_REQUIRE(backend != MDV2_BACKEND_TCGEN05, "gemm: shape M=%lld N=%d K=%d not tiled by the tcgen05 kernel",
                 (long long)M, N, K);
  }
  MDV2_REQUIRE(!(dtype == 0 && backend == MDV2_BACKEND_TCGEN05), "gemm: the tcgen05 kernel is bf16 only");
  int rc;
  if (dtype == 0) {
    Epilogue<float> ep = make_epilogue<float>(bias, bias_scale, act, gate_mod, mod_K, k_gate, res, ldr, res_mod,
                                              res_div, T, S, NC, N);
    rc = mdv2_gemm_simt<float>(a, lda, w, ldw, out, ldo, M, N, K, ep, st);
  } else {
    Epilogue<bf16> ep = make_epilogue<bf16>(bias, bias_scale, act, gate_mod, mod_K, k_gate, res, ldr, res_mod,
                                            res_div, T, S, NC, N);
    rc = mdv2_gemm_simt<bf16>(a, lda, w, ldw, out, ldo, M, N, K, ep, st);
  }
  if (rc) return rc;
  if (has_rms) return mdv2_qk_rmsnorm(out, ldo, M, head_dim, n_q, n_k, rms_q, rms_k, eps, dtype, st);
  return MDV2_OK;
}

extern "C" int mdv2_gemm_qkv_scatter(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias,
                                     int64_t M, int N, int K, const void* rms_q, const void* rms_k, int head_dim,
                                     float eps, void* const* dst_ptrs, int n_dst, int rows_per_seq, int rank,
                                     void* stream) {
  if (M == 0 || N == 0) return MDV2_OK;
  MDV2_REQUIRE(M > 0 && N > 0 && K > 0 && N % 3 == 0, "gemm_qkv_scatter: bad shape M=%lld N=%d K=%d", (long long)M, N, K);
  MDV2_REQUIRE(rms_q && rms_k && head_dim > 0 && dst_ptrs && n_dst >= 1 && n_dst <= 8, "gemm_qkv_scatter: bad arguments");
  // there is no SIMT flavour of the fused scatter: the caller falls back to mdv2_gemm + mdv2_scatter_runs
  MDV2_REQUIRE(device_is_sm100(), "gemm_qkv_scatter: needs sm_100a");
  RmsArgs rms{rms_q, rms_k, head_dim, N / 3, N / 3, eps};
  MDV2_REQUIRE(mdv2_gemm_sm100_supported(M, N, K, lda, ldw, 8, &rms), "gemm_qkv_scatter: shape not tiled by the tcgen05 kernel");
  Epilogue<bf16> ep = make_epilogue<bf16>(bias, 1.f, 0, nullptr, 0, 0, nullptr, 0, 0, 1, 1, 1, 1, N);
  QkvScatter sc{dst_ptrs, n_dst, rows_per_seq, rank};
  return mdv2_gemm_sm100(a, lda, w, ldw, nullptr, 8, M, N, K, ep, &rms, (cudaStream_t)stream, &sc);
}

extern "C" int mdv2_attention_scatter(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                      void* partial, int64_t ldp, int n_seq, int Lq, int Lk, int H, int D, float scale,
                                      const int32_t* kv_map, int n_sets, void* const* dst_ptrs, int n_dst,
                                      int rows_per_rank, int64_t dst_ld, int dst_col, void* stream) {
  if (n_seq == 0 || Lq == 0) return MDV2_OK;
  MDV2_REQUIRE(n_seq > 0 && Lq > 0 && Lk > 0 && H > 0 && D > 0, "attention_scatter: bad shape");
  MDV2_REQUIRE(n_sets >= 1 && (kv_map != nullptr || n_sets == 1), "attention_scatter: n_sets > 1 needs kv_map");
  MDV2_REQUIRE(dst_ptrs != nullptr && (n_sets == 1 || partial != nullptr), "attention_scatter: missing buffers");
  // bf16 / tcgen05 only: other shapes go through mdv2_attention + mdv2_scatter_runs
  MDV2_REQUIRE(device_is_sm100() && mdv2_attention_sm100_supported(ldq, ldk, ldv, ldp, Lq, Lk, H, D, n_sets),
               "attention_scatter: shape not taken by the tcgen05 kernel");
  AttnScatter sc{dst_ptrs, n_dst, rows_per_rank, dst_col, dst_ld};
  return mdv2_attention_sm100(q, ldq, k, ldk, v, ldv, partial, ldp, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets,
                              (cudaStream_t)stream, &sc);
}

extern "C" int mdv2_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                              void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                              const int32_t* kv_map, int n_sets, int dtype, int backend, void* stream) {
  if (n_seq == 0 || Lq == 0) return MDV2_OK;
  MDV2_REQUIRE(n_seq > 0 && Lq > 0 && Lk > 0 && H > 0 && D > 0, "attention: bad shape");
  MDV2_REQUIRE(dtype == 0 || dtype == 1, "attention: bad dtype %d", dtype);
  MDV2_REQUIRE(n_sets >= 1 && (kv_map != nullptr || n_sets == 1), "attention: n_sets > 1 needs kv_map");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 1 && backend != MDV2_BACKEND_SIMT) {
    if (device_is_sm100() && mdv2_attention_sm100_supported(ldq, ldk, ldv, ldo, Lq, Lk, H, D, n_sets))
      return mdv2_attention_sm100(q, ldq, k, ldk, v, ldv, out, ldo, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets, st);
    MDV2_REQUIRE(backend != MDV2_BACKEND_TCGEN05, "attention: shape not taken by the tcgen05 kernel");
  }
  MDV2_REQUIRE(!(dtype == 0 && backend == MDV2_BACKEND_TCGEN05), "attention: the tcgen05 kernel is bf16 only");
  return mdv2_attention_simt(q, ldq, k, ldk, v, ldv, out, ldo, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets, dtype, st);
}
