// attn_sm100.cu — flash attention on tcgen05 / TMEM / TMA (bf16 in, fp32 softmax, bf16 out).
//
// One CTA = one (sequence, head) x TWO 128-query tiles (256 query rows), 10 warps:
//   warp 0      TMA producer: Q (both tiles, once), K ring, V ring (independent rings: K_j is released as
//               soon as both S_t(j) are done, V_j after both PV_t(j))
//   warp 1      tcgen05.mma issuer (warp-uniform control flow, one elected lane per instruction) + TMEM owner
//   warps 2-5   softmax of query tile 0   (thread = query row = TMEM lane)
//   warps 6-9   softmax of query tile 1
// Per K/V tile j and query tile t:
//   S_t = Q_t · K_j^T   A = Q_t (smem, K-major SW128), B = K_j (smem, K-major SW128)   -> TMEM S_t (fp32)
//   softmax thread: one tcgen05.ld of its whole S row, running max with LAZY rescaling (the reference
//     maximum only moves when the row max grows by > 2^8, so O is rescaled in TMEM a handful of times
//     per row instead of once per tile), p = 2^(s·c − m·c), row sum, bf16 pack, tcgen05.st of P over
//     the S columns it just consumed
//   O_t += P_t · V_j    A = P_t (TMEM, bf16 pairs), B = V_j (smem, MN-major SW128)      -> TMEM O_t (fp32)
// The two query tiles ping-pong: while the tensor core runs PV_0(j) and S_0(j+1), warps 6-9 run the
// softmax of tile 1, and vice versa.  The tensor pipe executes MMAs in issue order, which is what makes
// P aliasing S safe: S_t(j+1) is issued after PV_t(j).
// head_dim 72 is not padded in HBM: TMA boxes are 64 columns wide, the second box (d = 64..127) is
// zero-filled beyond 72 by the TMA unit; QK^T runs K = 80, PV runs N = 80.
// Several K/V "sets" (cross-view neighbours) are processed back to back; each set has its own softmax
// and the bf16 outputs are summed like the reference's gather+sum (stdit3.py:311-322).
#include <cstdlib>

#include "common.cuh"
#include "sm100_ptx.cuh"

using namespace sm100;

int mdv2_make_tmap_3d_heads(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows);

namespace {
constexpr int BMQ = 128;          // query rows per tile
constexpr int NTHREADS = 320;
constexpr int KV_STAGES = 2;
constexpr int TMEM_COLS = 512;
constexpr int O_COL0 = 256;       // O_t at columns 256 + 128·t ; S_t (and P_t over it) at 128·t
constexpr float RESCALE_LOG2 = 8.f;

template <int DP, int BNK>
struct ACfg {
  static constexpr int NCH = DP > 64 ? 2 : 1;                 // 64-wide d-chunks per row
  static constexpr int Q_BYTES = NCH * BMQ * 128;             // one query tile
  static constexpr int KV_BYTES = NCH * BNK * 128;            // one K (or V) tile
  static constexpr int SMEM = 2 * Q_BYTES + KV_STAGES * 2 * KV_BYTES + 256 + 1024;
};

struct AParams {
  bf16* out;
  int64_t ldo;
  int Lq, Lk, H, D, n_sets;
  float scale_log2;
  const int32_t* kv_map;
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DP, int BNK>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_sm100_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                  const __grid_constant__ CUtensorMap tv, const AParams p) {
  using C = ACfg<DP, BNK>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                   // tile t at sQ + t·Q_BYTES
  uint8_t* sK = sQ + 2 * C::Q_BYTES;                    // stage s at sK + s·KV_BYTES
  uint8_t* sV = sK + KV_STAGES * C::KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KV_STAGES * C::KV_BYTES);
  uint64_t* q_full = bars;                              // 1
  uint64_t* k_full = bars + 1;                          // KV_STAGES
  uint64_t* k_empty = k_full + KV_STAGES;
  uint64_t* v_full = k_empty + KV_STAGES;
  uint64_t* v_empty = v_full + KV_STAGES;
  uint64_t* s_full = v_empty + KV_STAGES;               // 2 (per query tile)
  uint64_t* p_full = s_full + 2;                        // 2
  uint64_t* o_full = p_full + 2;                        // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qp = blockIdx.x, h = blockIdx.y, seq = blockIdx.z;
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int n_tiles = n_kt * p.n_sets;
  const int q_row0 = qp * 2 * BMQ;                      // first query row of this CTA within the sequence
  const int n_qt = (q_row0 + BMQ < p.Lq) ? 2 : 1;       // the second tile may lie entirely past Lq

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      mbar_expect_tx(q_full, n_qt * C::Q_BYTES);
      for (int t = 0; t < n_qt; ++t) {
        const int row = seq * p.Lq + q_row0 + t * BMQ;
#pragma unroll
        for (int c = 0; c < C::NCH; ++c) tma_load_3d(sQ + t * C::Q_BYTES + c * BMQ * 128, &tq, q_full, c * 64, h, row);
      }
      int j = 0;
      for (int set = 0; set < p.n_sets; ++set) {
        const int kseq = p.kv_map ? p.kv_map[seq * p.n_sets + set] : seq;
        for (int kt = 0; kt < n_kt; ++kt, ++j) {
          const int st = j % KV_STAGES;
          const uint32_t ph = (uint32_t)((j / KV_STAGES) & 1);
          const int row = kseq * p.Lk + kt * BNK;
          mbar_wait(&k_empty[st], ph ^ 1);
          mbar_expect_tx(&k_full[st], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c)
            tma_load_3d(sK + st * C::KV_BYTES + c * BNK * 128, &tk, &k_full[st], c * 64, h, row);
          mbar_wait(&v_empty[st], ph ^ 1);
          mbar_expect_tx(&v_full[st], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c)
            tma_load_3d(sV + st * C::KV_BYTES + c * BNK * 128, &tv, &v_full[st], c * 64, h, row);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // The whole warp runs the warp-uniform control flow and one elected lane issues each tcgen05 instruction:
    // descriptors then live in uniform registers and the MMAs of a tile go out back to back (with a lane-0
    // branch the compiler wrapped every MMA in an R2UR + ELECT loop, ~50 clk per 40-64 clk MMA).  Every
    // shared-memory descriptor is base + constant (the 14-bit address field never carries).
    {
      constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, BNK, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);     // A = P from TMEM, B = V is MN-major
      const uint64_t dq0 = make_desc_k_sw128(smem_u32(sQ));
      const uint64_t dk0 = make_desc_k_sw128(smem_u32(sK));
      const uint64_t dv0 = make_desc_mn_sw128(smem_u32(sV), BNK * 128, 1024);
      auto issue_qk = [&](int t, int st) {
        const uint64_t da = dq0 + (uint64_t)(t * (C::Q_BYTES >> 4));
        const uint64_t db = dk0 + (uint64_t)(st * (C::KV_BYTES >> 4));
#pragma unroll
        for (int ks = 0; ks < DP / 16; ++ks) {
          const uint64_t oa = (uint64_t)((ks / 4) * ((BMQ * 128) >> 4) + (ks % 4) * 2);
          const uint64_t ob = (uint64_t)((ks / 4) * ((BNK * 128) >> 4) + (ks % 4) * 2);
          if (elect_one()) umma_bf16_ss(tmem_base + t * 128, da + oa, db + ob, idesc_qk, ks != 0);
        }
        if (elect_one()) umma_commit(&s_full[t]);
        __syncwarp();
      };
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      for (int t = 0; t < n_qt; ++t) issue_qk(t, 0);
      if (elect_one()) umma_commit(&k_empty[0]);
      __syncwarp();
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % KV_STAGES, st1 = (j + 1) % KV_STAGES;
        const bool more = j + 1 < n_tiles;
        const bool fresh = (j % n_kt) == 0;              // first tile of a K/V set: O starts over
        mbar_wait(&v_full[st], (uint32_t)((j / KV_STAGES) & 1));
        if (more) mbar_wait(&k_full[st1], (uint32_t)(((j + 1) / KV_STAGES) & 1));
        const uint64_t dv = dv0 + (uint64_t)(st * (C::KV_BYTES >> 4));
        for (int t = 0; t < n_qt; ++t) {
          mbar_wait(&p_full[t], (uint32_t)(j & 1));
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < BNK / 16; ++ks)
            if (elect_one())
              umma_bf16_ts(tmem_base + O_COL0 + t * 128, tmem_base + t * 128 + ks * 8,
                           dv + (uint64_t)(ks * ((16 * 128) >> 4)), idesc_pv, (ks != 0) || !fresh);
          if (elect_one()) umma_commit(&o_full[t]);
          __syncwarp();
          if (more) issue_qk(t, st1);
        }
        if (elect_one()) {
          umma_commit(&v_empty[st]);
          if (more) umma_commit(&k_empty[st1]);
        }
        __syncwarp();
      }
    }
  } else if (((warp - 2) >> 2) < n_qt) {
    // ================= softmax warps: thread = query row =================
    const int t = (warp - 2) >> 2;
    const int quarter = warp & 3;                      // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_addr = lane_addr + t * 128, o_addr = lane_addr + O_COL0 + t * 128;
    const int q_idx = q_row0 + t * BMQ + row;
    const float c = p.scale_log2;
    bf16* orow = p.out + ((int64_t)seq * p.Lq + q_idx) * p.ldo + (int64_t)h * p.D;
    int j = 0;
    for (int set = 0; set < p.n_sets; ++set) {
      float m_used = -INFINITY, l_run = 0.f;
      for (int kt = 0; kt < n_kt; ++kt, ++j) {
        mbar_wait(&s_full[t], (uint32_t)(j & 1));
        tc_fence_after();
        uint32_t sr[BNK];
#pragma unroll
        for (int c0 = 0; c0 < BNK; c0 += 32) tmem_ld_x32(s_addr + c0, sr + c0);
        tmem_ld_wait();
        if (kt == n_kt - 1) {                          // ragged last tile: keys past Lk do not exist
          const int kv_left = p.Lk - kt * BNK;
          if (kv_left < BNK) {
#pragma unroll
            for (int i = 0; i < BNK; ++i)
              if (i >= kv_left) sr[i] = 0xff800000u;   // -inf
          }
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < BNK; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(sr[i]));
          mx1 = fmaxf(mx1, __uint_as_float(sr[i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(sr[i + 2]));
          mx3 = fmaxf(mx3, __uint_as_float(sr[i + 3]));
        }
        const float m_new = fmaxf(m_used, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)));
        const bool grow = (m_new - m_used) * c > RESCALE_LOG2;          // true on the first tile (m_used = -inf)
        if (__any_sync(0xffffffffu, grow)) {
          const float alpha = grow ? ex2((m_used - m_new) * c) : 1.f;  // 0 on the first tile
          if (grow) m_used = m_new;
          l_run *= alpha;
          if (kt > 0) {
            // rescale the running output in TMEM: PV(j-1) must have completed; PV(j) is not issued
            // before this warp arrives on p_full below
            mbar_wait(&o_full[t], (uint32_t)((j - 1) & 1));
            tc_fence_after();
#pragma unroll
            for (int c0 = 0; c0 < DP; c0 += 16) {
              uint32_t r[16];
              tmem_ld_x16(o_addr + c0, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
              tmem_st_x16(o_addr + c0, r);
            }
          }
        }
        const float moff = m_used * c;
        float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
        uint32_t pk[BNK / 2];
#pragma unroll
        for (int i = 0; i < BNK; i += 4) {
          const float e0 = ex2(fmaf(__uint_as_float(sr[i]), c, -moff));
          const float e1 = ex2(fmaf(__uint_as_float(sr[i + 1]), c, -moff));
          const float e2 = ex2(fmaf(__uint_as_float(sr[i + 2]), c, -moff));
          const float e3 = ex2(fmaf(__uint_as_float(sr[i + 3]), c, -moff));
          l0 += e0; l1 += e1; l2 += e2; l3 += e3;
          __nv_bfloat162 b01 = __floats2bfloat162_rn(e0, e1), b23 = __floats2bfloat162_rn(e2, e3);
          pk[i / 2] = *reinterpret_cast<uint32_t*>(&b01);
          pk[i / 2 + 1] = *reinterpret_cast<uint32_t*>(&b23);
        }
        l_run += (l0 + l1) + (l2 + l3);
#pragma unroll
        for (int c0 = 0; c0 < BNK / 2; c0 += 32) tmem_st_x32(s_addr + c0, pk + c0);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        if (kt == n_kt - 1) {
          // end of a K/V set: normalise, round to bf16, add to what earlier sets left in `out`
          mbar_wait(&o_full[t], (uint32_t)(j & 1));
          tc_fence_after();
          const float inv = 1.f / l_run;
#pragma unroll
          for (int c0 = 0; c0 < DP; c0 += 16) {
            uint32_t r[16];
            tmem_ld_x16(o_addr + c0, r);
            tmem_ld_wait();
            if (q_idx < p.Lq) {
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int d = c0 + u * 8;
                if (d < p.D) {
                  float f[8];
#pragma unroll
                  for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(r[u * 8 + i]) * inv;
                  Vec16<bf16> o;
                  o.pack(f);
                  if (set > 0) {
                    Vec16<bf16> prev;
                    prev.raw = *reinterpret_cast<const uint4*>(orow + d);
                    float g[8], a[8];
                    prev.unpack(g);
                    o.unpack(a);
#pragma unroll
                    for (int i = 0; i < 8; ++i) a[i] += g[i];
                    o.pack(a);
                  }
                  *reinterpret_cast<uint4*>(orow + d) = o.raw;
                }
              }
            }
          }
          tc_fence_before();
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int DP, int BNK>
int launch_attn(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p, int n_seq,
                cudaStream_t st) {
  using C = ACfg<DP, BNK>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_sm100_kernel<DP, BNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("attention_sm100: cannot set %d B dynamic smem: %s", C::SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    configured = true;
  }
  dim3 grid((p.Lq + 2 * BMQ - 1) / (2 * BMQ), p.H, n_seq);
  attn_sm100_kernel<DP, BNK><<<grid, NTHREADS, C::SMEM, st>>>(tq, tk, tv, p);
  MDV2_CHECK_LAUNCH("attention_sm100");
  return MDV2_OK;
}
}  // namespace

bool mdv2_attention_sm100_supported(int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int Lq, int Lk, int H, int D,
                                    int n_sets) {
  if (D != 64 && D != 72) return false;
  if (ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8) return false;
  if (H > 65535 || n_sets < 1) return false;
  return Lq > 0 && Lk > 0;
}

int mdv2_attention_sm100(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                         void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                         const int32_t* kv_map, int n_sets, cudaStream_t st) {
  MDV2_REQUIRE(mdv2_attention_sm100_supported(ldq, ldk, ldv, ldo, Lq, Lk, H, D, n_sets), "attention_sm100: unsupported shape");
  MDV2_REQUIRE(n_seq <= 65535, "attention_sm100: too many sequences");
  MDV2_REQUIRE(((uintptr_t)q % 16 == 0) && ((uintptr_t)k % 16 == 0) && ((uintptr_t)v % 16 == 0) && ((uintptr_t)out % 16 == 0),
               "attention_sm100: operands must be 16-byte aligned");
  const int bnk = Lk <= 64 ? 64 : 128;
  CUtensorMap tq, tk, tv;
  int rc = mdv2_make_tmap_3d_heads(&tq, q, (uint64_t)n_seq * Lq, H, D, ldq, BMQ);
  if (rc) return rc;
  // K/V hold their own number of sequences (cross-attention) — only the row count bounds the TMA zero fill
  rc = mdv2_make_tmap_3d_heads(&tk, k, (uint64_t)n_seq * Lk, H, D, ldk, bnk);
  if (rc) return rc;
  rc = mdv2_make_tmap_3d_heads(&tv, v, (uint64_t)n_seq * Lk, H, D, ldv, bnk);
  if (rc) return rc;
  AParams p;
  p.out = (bf16*)out;
  p.ldo = ldo;
  p.Lq = Lq; p.Lk = Lk; p.H = H; p.D = D; p.n_sets = n_sets;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.kv_map = kv_map;
  if (D == 72) {
    if (bnk == 64) return launch_attn<80, 64>(tq, tk, tv, p, n_seq, st);
    return launch_attn<80, 128>(tq, tk, tv, p, n_seq, st);
  }
  if (bnk == 64) return launch_attn<64, 64>(tq, tk, tv, p, n_seq, st);
  return launch_attn<64, 128>(tq, tk, tv, p, n_seq, st);
}
