// attn_sm100.cu — tcgen05 flash attention (placeholder until the kernel lands: reports unsupported)
#include "common.cuh"
bool mdv2_attention_sm100_supported(int64_t, int64_t, int64_t, int64_t, int, int, int, int, int) { return false; }
int mdv2_attention_sm100(const void*, int64_t, const void*, int64_t, const void*, int64_t, void*, int64_t, int, int, int,
                         int, int, float, const int32_t*, int, cudaStream_t) {
  mdv2_set_error("attention_sm100: not built");
  return MDV2_EINVAL;
}
