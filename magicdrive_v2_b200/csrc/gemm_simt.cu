// gemm_simt.cu — CUDA-core GEMM with the full STDiT3 epilogue.  This is the fp32 validation path
// (parity <= 1e-4 needs true fp32 FMAs, not TF32) and the path for shapes the tcgen05 kernel does
// not tile (tiny K / N).  fp32 accumulate for both dtypes.
#include "gemm_epilogue.cuh"

namespace {
constexpr int BM = 64, BN = 64, BK = 16;

template <typename T>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const T* __restrict__ A, int64_t lda,
                                                        const T* __restrict__ W, int64_t ldw, T* __restrict__ O,
                                                        int64_t ldo, int64_t M, int N, int K, Epilogue<T> ep) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Ws[BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.y * BM;
  const int n0 = blockIdx.x * BN;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 4x4 outputs each
  float acc[4][4] = {};
  // loader mapping: 64 rows x 16 k -> 1024 elements, 4 per thread
  const int lr = tid >> 2;         // 0..63
  const int lk = (tid & 3) * 4;    // 0,4,8,12
  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + lk + j;
      const int64_t ar = m0 + lr;
      As[lk + j][lr] = (ar < M && k < K) ? to_f(A[ar * lda + k]) : 0.f;
      const int wr = n0 + lr;
      Ws[lk + j][lr] = (wr < N && k < K) ? to_f(W[(int64_t)wr * ldw + k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Ws[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = m0 + ty * 4 + i;
    if (r >= M) continue;
    RowCtx rc = ep.row_ctx(r);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + tx * 4 + j;
      if (c < N) O[r * ldo + c] = from_f<T>(ep.apply(acc[i][j], r, c, rc));
    }
  }
}
}  // namespace

template <typename T>
int mdv2_gemm_simt(const void* a, int64_t lda, const void* w, int64_t ldw, void* out, int64_t ldo, int64_t M,
                   int N, int K, const Epilogue<T>& ep, cudaStream_t st) {
  dim3 grid((N + BN - 1) / BN, (unsigned)((M + BM - 1) / BM)), block(256);
  MDV2_REQUIRE(grid.y <= 65535u, "gemm: M too large");
  gemm_simt_kernel<T><<<grid, block, 0, st>>>((const T*)a, lda, (const T*)w, ldw, (T*)out, ldo, M, N, K, ep);
  MDV2_CHECK_LAUNCH("gemm_simt");
  return MDV2_OK;
}
template int mdv2_gemm_simt<float>(const void*, int64_t, const void*, int64_t, void*, int64_t, int64_t, int, int,
                                   const Epilogue<float>&, cudaStream_t);
template int mdv2_gemm_simt<bf16>(const void*, int64_t, const void*, int64_t, void*, int64_t, int64_t, int, int,
                                  const Epilogue<bf16>&, cudaStream_t);
