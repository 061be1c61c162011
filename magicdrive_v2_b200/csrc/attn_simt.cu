// attn_simt.cu — CUDA-core softmax attention (online softmax, fp32 math).  The fp32 validation path
// and the fallback for shapes the tcgen05 kernel does not take.  One thread per query row, K/V tiles
// broadcast from shared memory.
#include "common.cuh"

namespace {
constexpr int QB = 128;  // queries (threads) per block
constexpr int KT = 16;   // keys per shared-memory tile

template <typename T, int DMAX>
__global__ void __launch_bounds__(QB) attn_simt_kernel(const T* __restrict__ q, int64_t ldq,
                                                       const T* __restrict__ k, int64_t ldk,
                                                       const T* __restrict__ v, int64_t ldv, T* __restrict__ out,
                                                       int64_t ldo, int Lq, int Lk, int H, int D, float scale,
                                                       const int32_t* __restrict__ kv_map, int n_sets) {
  __shared__ float Ks[KT][DMAX];
  __shared__ float Vs[KT][DMAX];
  const int seq = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * QB + threadIdx.x;
  const bool active = qi < Lq;
  float qr[DMAX], o[DMAX];
  if (active) {
    const T* qp = q + ((int64_t)seq * Lq + qi) * ldq + (int64_t)h * D;
#pragma unroll
    for (int d = 0; d < DMAX; ++d) qr[d] = d < D ? to_f(qp[d]) * scale : 0.f;
  }
  T* op = out + ((int64_t)seq * Lq + (active ? qi : 0)) * ldo + (int64_t)h * D;
  for (int set = 0; set < n_sets; ++set) {
    const int kseq = kv_map ? kv_map[seq * n_sets + set] : seq;
    const T* kb = k + (int64_t)kseq * Lk * ldk + (int64_t)h * D;
    const T* vb = v + (int64_t)kseq * Lk * ldv + (int64_t)h * D;
    float m = -INFINITY, l = 0.f;
#pragma unroll
    for (int d = 0; d < DMAX; ++d) o[d] = 0.f;
    for (int k0 = 0; k0 < Lk; k0 += KT) {
      __syncthreads();
      for (int idx = threadIdx.x; idx < KT * D; idx += QB) {
        const int kr = idx / D, d = idx % D;
        const bool ok = k0 + kr < Lk;
        Ks[kr][d] = ok ? to_f(kb[(int64_t)(k0 + kr) * ldk + d]) : 0.f;
        Vs[kr][d] = ok ? to_f(vb[(int64_t)(k0 + kr) * ldv + d]) : 0.f;
      }
      __syncthreads();
      if (!active) continue;
      float s[KT];
      float tm = -INFINITY;
#pragma unroll
      for (int j = 0; j < KT; ++j) {
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < DMAX; ++d)
          if (d < D) a = fmaf(qr[d], Ks[j][d], a);
        s[j] = (k0 + j < Lk) ? a : -INFINITY;
        tm = fmaxf(tm, s[j]);
      }
      const float mn = fmaxf(m, tm);
      const float corr = __expf(m - mn);
      l *= corr;
#pragma unroll
      for (int d = 0; d < DMAX; ++d) o[d] *= corr;
#pragma unroll
      for (int j = 0; j < KT; ++j) {
        const float p = __expf(s[j] - mn);
        l += p;
#pragma unroll
        for (int d = 0; d < DMAX; ++d)
          if (d < D) o[d] = fmaf(p, Vs[j][d], o[d]);
      }
      m = mn;
    }
    if (active) {
      const float inv = 1.f / l;
#pragma unroll
      for (int d = 0; d < DMAX; ++d) {
        if (d < D) {
          float r = o[d] * inv;
          if (set > 0) r += to_f(op[d]);
          op[d] = from_f<T>(r);
        }
      }
    }
  }
}

template <typename T>
int launch(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, void* out,
           int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale, const int32_t* kv_map, int n_sets,
           cudaStream_t st) {
  dim3 grid((Lq + QB - 1) / QB, H, n_seq), block(QB);
  MDV2_REQUIRE(n_seq <= 65535 && H <= 65535, "attention: too many sequences for the SIMT grid");
  if (D <= 64)
    attn_simt_kernel<T, 64><<<grid, block, 0, st>>>((const T*)q, ldq, (const T*)k, ldk, (const T*)v, ldv, (T*)out,
                                                    ldo, Lq, Lk, H, D, scale, kv_map, n_sets);
  else
    attn_simt_kernel<T, 80><<<grid, block, 0, st>>>((const T*)q, ldq, (const T*)k, ldk, (const T*)v, ldv, (T*)out,
                                                    ldo, Lq, Lk, H, D, scale, kv_map, n_sets);
  MDV2_CHECK_LAUNCH("attention_simt");
  return MDV2_OK;
}
}  // namespace

int mdv2_attention_simt(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                        void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                        const int32_t* kv_map, int n_sets, int dtype, cudaStream_t st) {
  MDV2_REQUIRE(D <= 80, "attention(simt): head_dim=%d > 80 unsupported", D);
  if (dtype == 0) return launch<float>(q, ldq, k, ldk, v, ldv, out, ldo, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets, st);
  return launch<bf16>(q, ldq, k, ldk, v, ldv, out, ldo, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets, st);
}
