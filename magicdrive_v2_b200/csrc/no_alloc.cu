// no_alloc.cu — the one entry point of libmdv2 whose scratch memory is a file-scope device array (include/mdv2.h:
// mdv2_attention_scatter_static).  For callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops,
// harnesses that forbid it): the running sum of the earlier K/V sets, which mdv2_attention_scatter takes from the caller
// as `partial`, lives in g_static_scratch.  Nothing here (or below it) calls cudaMalloc / cudaFree or an allocating
// library routine.
#include "common.cuh"

bool mdv2_attention_sm100_supported(int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int Lq, int Lk, int H, int D,
                                    int n_sets);
int mdv2_attention_sm100(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                         void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                         const int32_t* kv_map, int n_sets, cudaStream_t st, const AttnScatter* scatter = nullptr);

// Every GPU that loads the library holds these MDV2_STATIC_SCRATCH_BYTES until the process exits.  All calls share
// the array and it keeps the last call's contents; the kernel OVERWRITES its rows with K/V set 0 before set 1 reads
// them back (attn_sm100.cu: `set > 0` adds, set 0 stores), so nothing has to start at zero and nothing is cleared.
static __device__ __align__(256) unsigned char g_static_scratch[MDV2_STATIC_SCRATCH_BYTES];

// In host code `g_static_scratch` is only a host-side stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_ptr(void** out) {
  static void* ptrs[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    mdv2_set_error("attention_scatter_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return MDV2_ECUDA;
  }
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      mdv2_set_error("attention_scatter_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return MDV2_OK;
}

extern "C" int64_t mdv2_static_scratch_bytes(void) { return MDV2_STATIC_SCRATCH_BYTES; }

extern "C" int mdv2_attention_scatter_static(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                             int64_t ldv, int n_seq, int Lq, int Lk, int H, int D, float scale,
                                             const int32_t* kv_map, int n_sets, void* const* dst_ptrs, int n_dst,
                                             int rows_per_rank, int64_t dst_ld, int dst_col, void* stream) {
  if (n_seq == 0 || Lq == 0) return MDV2_OK;
  MDV2_REQUIRE(n_seq > 0 && Lq > 0 && Lk > 0 && H > 0 && D > 0, "attention_scatter_static: bad shape");
  MDV2_REQUIRE(n_sets >= 1 && (kv_map != nullptr || n_sets == 1), "attention_scatter_static: n_sets > 1 needs kv_map");
  MDV2_REQUIRE(dst_ptrs != nullptr, "attention_scatter_static: missing destination buffers");
  const int64_t ldp = (int64_t)H * D;                                  // partial sums: [n_seq·Lq, H·D] bf16, packed
  const int64_t need = (int64_t)n_seq * Lq * ldp * (int64_t)sizeof(bf16);
  MDV2_REQUIRE(need <= (int64_t)MDV2_STATIC_SCRATCH_BYTES,
               "attention_scatter_static: n_seq·Lq·H·D·2 = %lld bytes of partial sums exceed the static scratch of %lld "
               "bytes (MDV2_STATIC_SCRATCH_BYTES); use mdv2_attention_scatter with a caller-owned buffer",
               (long long)need, (long long)MDV2_STATIC_SCRATCH_BYTES);
  int major = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  MDV2_REQUIRE(major == 10 && mdv2_attention_sm100_supported(ldq, ldk, ldv, ldp, Lq, Lk, H, D, n_sets),
               "attention_scatter_static: shape not taken by the tcgen05 kernel");
  void* scratch = nullptr;
  int rc = static_scratch_ptr(&scratch);
  if (rc) return rc;
  AttnScatter sc{dst_ptrs, n_dst, rows_per_rank, dst_col, dst_ld};
  return mdv2_attention_sm100(q, ldq, k, ldk, v, ldv, scratch, ldp, n_seq, Lq, Lk, H, D, scale, kv_map, n_sets,
                              (cudaStream_t)stream, &sc);
}
