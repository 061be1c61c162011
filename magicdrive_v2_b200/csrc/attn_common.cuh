// attn_common.cuh — what the two tcgen05 attention kernels (attn_sm100.cu: exact online softmax, 64-key flavour;
// attn_fast_sm100.cu: the 128-key fast path) share: launch parameters, the debug timeline, 2^x helpers.
#pragma once
#include "common.cuh"
#include "sm100_ptx.cuh"

namespace attn {
constexpr int BMQ = 128;          // query rows per tile

struct AParams {
  bf16* out;
  int64_t ldo;
  int Lq, Lk, H, D, n_sets;
  float scale_log2;
  const int32_t* kv_map;
  // sequence parallelism: the normalised output rows of the LAST K/V set go straight to the rank that owns the row
  // (peer-mapped [n_seq·rows_per_rank, peer_ld] buffers, this rank's head group at column peer_col); `out` then only
  // carries the running sum of earlier sets.  n_peer == 0: everything goes to `out`.
  bf16* peer_out[8];
  int n_peer, rows_per_rank, peer_col;
  int64_t peer_ld;
  long long* trace;        // debug timeline (tools/attn_trace.py) or nullptr
  int dbg;                 // debug switches (MDV2_ATTN_DBG): bit 0 swaps the TMA / MMA warps, bit 1 disables the ping-pong
};

constexpr int TRACE_STEPS = 96, TRACE_EV = 12;
__device__ __forceinline__ long long clk64() {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
  return t;
}

__device__ __forceinline__ long long gtime_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");
  return t;
}
__device__ __forceinline__ uint32_t smid() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
  return r;
}
// per-CTA life line behind the per-warp timeline (TRACE builds): 8 int64 per CTA —
// 0 start, 1 SM id, 2 first S tile seen by softmax warp 0, 3 last P of the CTA stored (warp 0), 4 O complete, 5 end
constexpr int TRACE_CTA_OFF = 20 * TRACE_STEPS * TRACE_EV, TRACE_CTA_EV = 8;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x for a pair on the FMA pipe (no MUFU): x = n + f, n = round(x), f in [-1/2, 1/2]; 2^f by the degree-3 minimax
// polynomial (max rel. error 7.5e-5 = 2^-13.7, far below the bf16 rounding of P), 2^n added into the exponent field.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  const float MAGIC = 12582912.f;                        // 1.5 * 2^23: x + MAGIC rounds x to an integer in the mantissa
  x.x = fmaxf(x.x, -126.f);
  x.y = fmaxf(x.y, -126.f);
  const float2 t = sm100::fadd2(x, make_float2(MAGIC, MAGIC));
  const float2 nf = sm100::fadd2(t, make_float2(-MAGIC, -MAGIC));
  const float2 f = sm100::ffma2(nf, make_float2(-1.f, -1.f), x);
  float2 q = sm100::ffma2(f, make_float2(0.05517202615737915f, 0.05517202615737915f),
                   make_float2(0.2426111400127411f, 0.2426111400127411f));
  q = sm100::ffma2(q, f, make_float2(0.693260908126831f, 0.693260908126831f));
  q = sm100::ffma2(q, f, make_float2(0.9999280571937561f, 0.9999280571937561f));
  float2 r;
  r.x = __uint_as_float(__float_as_uint(q.x) + (__float_as_uint(t.x) << 23));
  r.y = __uint_as_float(__float_as_uint(q.y) + (__float_as_uint(t.y) << 23));
  return r;
}
// pair index (0..) -> evaluated by the polynomial?  PF of every 8 pairs, spread evenly.
__host__ __device__ constexpr bool poly_pair(int pair, int PF) { return ((pair % 8) * PF) / 8 != (((pair % 8) + 1) * PF) / 8; }
}  // namespace attn

int mdv2_make_tmap_3d_heads_sw32(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                                 uint32_t box_rows);
int mdv2_make_tmap_4d_store(CUtensorMap* out, void* ptr, uint64_t n_seq, uint64_t L, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows);
int mdv2_make_tmap_3d_heads(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows);
