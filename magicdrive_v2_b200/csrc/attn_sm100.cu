// attn_sm100.cu — flash attention on tcgen05 / TMEM / TMA (bf16 in, fp32 softmax, bf16 out).
//
// One CTA = one (sequence, head, 128-query tile).  warp 0: TMA producer (Q once, K/V ring),
// warp 1: tcgen05.mma issuer, warps 2-5: softmax (thread = query row = TMEM lane).
//   S = Q·K^T   : A = Q (K-major, SW128), B = K tile (K-major, SW128)      -> TMEM S[2] (double buffer)
//   O_j = P·V   : A = P (bf16, written to smem in the SW128 K-major layout), B = V tile (MN-major, SW128)
// Online softmax state (m, l) and the running output live in the softmax threads' registers; the
// per-tile PV product is read back from TMEM and folded in with the rescale factor.
// head_dim 72 is handled without padding in HBM: the TMA box is 64 wide, the second box (d = 64..127)
// is zero-filled beyond 72 by the TMA unit; QK^T runs K = 80, PV runs N = 80.
// Several K/V "sets" (cross-view neighbours) are processed back to back and their normalised
// outputs summed.
#include "common.cuh"
#include "sm100_ptx.cuh"

using namespace sm100;

int mdv2_make_tmap_3d_heads(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows);

namespace {
constexpr int BMQ = 128;        // query rows per CTA
constexpr int NTHREADS = 192;
constexpr int CHUNK_COLS = 64;  // bf16 columns per 128-byte swizzle row

template <int DP, int BNK>
struct ACfg {
  static constexpr int NCH = DP > 64 ? 2 : 1;                 // 64-wide d-chunks per tile
  static constexpr int Q_BYTES = NCH * BMQ * 128;
  static constexpr int KV_BYTES = NCH * BNK * 128;            // one of K or V
  static constexpr int STAGES = 2;
  static constexpr int P_BYTES = (BNK / 64) * BMQ * 128;
  static constexpr int SMEM = Q_BYTES + STAGES * 2 * KV_BYTES + P_BYTES + 256 + 1024;
  static constexpr int S_COLS = BNK;                          // fp32 columns per S buffer
  static constexpr int O_COL = 2 * 128;                       // TMEM column of the PV tile
};

struct AParams {
  bf16* out;
  int64_t ldo;
  int Lq, Lk, H, D, n_sets;
  float scale_log2;
  const int32_t* kv_map;
};

template <int DP, int BNK>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_sm100_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                  const __grid_constant__ CUtensorMap tv, const AParams p) {
  using C = ACfg<DP, BNK>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + C::Q_BYTES;                 // stage s: K at sKV + s*2*KV_BYTES, V right after
  uint8_t* sP = sKV + C::STAGES * 2 * C::KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + C::P_BYTES);
  uint64_t* q_full = bars;                        // 1
  uint64_t* k_full = bars + 1;                    // STAGES
  uint64_t* v_full = k_full + C::STAGES;          // STAGES
  uint64_t* kv_empty = v_full + C::STAGES;        // STAGES
  uint64_t* s_full = kv_empty + C::STAGES;        // 2
  uint64_t* s_empty = s_full + 2;                 // 2
  uint64_t* p_full = s_empty + 2;                 // 1
  uint64_t* o_full = p_full + 1;                  // 1
  uint64_t* o_empty = o_full + 1;                 // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, seq = blockIdx.z;
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int n_tiles = n_kt * p.n_sets;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    mbar_init(q_full, 1);
    for (int i = 0; i < C::STAGES; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 4);
    }
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_init(o_empty, 4);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      mbar_expect_tx(q_full, C::Q_BYTES);
      const int q_row = seq * p.Lq + qt * BMQ;
#pragma unroll
      for (int c = 0; c < C::NCH; ++c) tma_load_3d(sQ + c * BMQ * 128, &tq, q_full, c * 64, h, q_row);
      int stage = 0;
      uint32_t phase = 0;
      for (int set = 0; set < p.n_sets; ++set) {
        const int kseq = p.kv_map ? p.kv_map[seq * p.n_sets + set] : seq;
        for (int kt = 0; kt < n_kt; ++kt) {
          mbar_wait(&kv_empty[stage], phase ^ 1);
          uint8_t* sk = sKV + stage * 2 * C::KV_BYTES;
          uint8_t* sv = sk + C::KV_BYTES;
          const int row = kseq * p.Lk + kt * BNK;
          mbar_expect_tx(&k_full[stage], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c) tma_load_3d(sk + c * BNK * 128, &tk, &k_full[stage], c * 64, h, row);
          mbar_expect_tx(&v_full[stage], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c) tma_load_3d(sv + c * BNK * 128, &tv, &v_full[stage], c * 64, h, row);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, BNK, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);     // B (= V) is MN-major
      constexpr int KSTEPS_QK = DP / 16;
      auto issue_qk = [&](int j) {
        const int st = j % C::STAGES, buf = j & 1;
        mbar_wait(&k_full[st], (uint32_t)((j / C::STAGES) & 1));
        mbar_wait(&s_empty[buf], (uint32_t)(((j >> 1) & 1) ^ 1));
        tc_fence_after();
        const uint32_t aq = smem_u32(sQ), bk = smem_u32(sKV + st * 2 * C::KV_BYTES);
#pragma unroll
        for (int ks = 0; ks < KSTEPS_QK; ++ks) {
          const int c = ks / 4, kk = ks % 4;
          const uint64_t da = make_desc_k_sw128(aq + c * BMQ * 128 + kk * 32);
          const uint64_t db = make_desc_k_sw128(bk + c * BNK * 128 + kk * 32);
          umma_bf16_ss(tmem_base + buf * 128, da, db, idesc_qk, ks != 0);
        }
        umma_commit(&s_full[buf]);
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);
        const int st = j % C::STAGES;
        mbar_wait(&v_full[st], (uint32_t)((j / C::STAGES) & 1));
        mbar_wait(p_full, (uint32_t)(j & 1));
        mbar_wait(o_empty, (uint32_t)((j & 1) ^ 1));
        tc_fence_after();
        const uint32_t ap = smem_u32(sP), bv = smem_u32(sKV + st * 2 * C::KV_BYTES + C::KV_BYTES);
#pragma unroll
        for (int ks = 0; ks < BNK / 16; ++ks) {
          const uint64_t da = make_desc_k_sw128(ap + (ks / 4) * BMQ * 128 + (ks % 4) * 32);
          const uint64_t db = make_desc_mn_sw128(bv + ks * 16 * 128, BNK * 128, 1024);
          umma_bf16_ss(tmem_base + C::O_COL, da, db, idesc_pv, ks != 0);
        }
        umma_commit(o_full);
        umma_commit(&kv_empty[st]);
      }
    }
  } else {
    // ================= softmax warps: thread = query row =================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;               // row within the q tile == TMEM lane
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const int q_idx = qt * BMQ + row;
    float acc_out[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) acc_out[d] = 0.f;
    int j = 0;
    bool pv_pending = false;                           // a PV tile is in flight / waiting to be folded
    for (int set = 0; set < p.n_sets; ++set) {
      float o_run[DP];
#pragma unroll
      for (int d = 0; d < DP; ++d) o_run[d] = 0.f;
      float m_run = -INFINITY, l_run = 0.f;
      for (int kt = 0; kt < n_kt; ++kt, ++j) {
        const int buf = j & 1;
        const int kv_left = p.Lk - kt * BNK;           // valid keys in this tile (may exceed BNK)
        mbar_wait(&s_full[buf], (uint32_t)((j >> 1) & 1));
        tc_fence_after();
        // pass 1: row max
        float tmax = -INFINITY;
#pragma unroll
        for (int c0 = 0; c0 < BNK; c0 += 32) {
          uint32_t r[32];
          tmem_ld_x32(lane_addr + buf * 128 + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < kv_left) tmax = fmaxf(tmax, __uint_as_float(r[i]));
        }
        const float m_new = fmaxf(m_run, tmax);
        const float alpha = exp2f((m_run - m_new) * p.scale_log2);   // 0 on the first tile (m_run = -inf)
        const float moff = m_new * p.scale_log2;
        // the previous PV must have finished reading P before it is overwritten
        if (pv_pending) {
          mbar_wait(o_full, (uint32_t)((j - 1) & 1));
          tc_fence_after();
          // fold the previous PV tile into the running output (rescale was applied when it was formed)
          uint32_t r[32];
#pragma unroll
          for (int c0 = 0; c0 < DP; c0 += 16) {
            tmem_ld_x16(lane_addr + C::O_COL + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o_run[c0 + i] += __uint_as_float(r[i]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(o_empty);
          pv_pending = false;
        }
#pragma unroll
        for (int d = 0; d < DP; ++d) o_run[d] *= alpha;
        l_run *= alpha;
        // pass 2: probabilities -> bf16 -> smem (SW128 K-major), row sum
        float lsum = 0.f;
#pragma unroll
        for (int c0 = 0; c0 < BNK; c0 += 32) {
          uint32_t r[32];
          tmem_ld_x32(lane_addr + buf * 128 + c0, r);
          tmem_ld_wait();
          float pr[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float e = exp2f(__uint_as_float(r[i]) * p.scale_log2 - moff);
            pr[i] = (c0 + i < kv_left) ? e : 0.f;
            lsum += pr[i];
          }
          uint8_t* prow = sP + (c0 / 64) * BMQ * 128 + row * 128;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            Vec16<bf16> pk;
            pk.pack(pr + u * 8);
            const int unit = ((c0 % 64) / 8 + u) ^ (row & 7);
            *reinterpret_cast<uint4*>(prow + unit * 16) = pk.raw;
          }
        }
        l_run += lsum;
        m_run = m_new;
        tc_fence_before();
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&s_empty[buf]);
          mbar_arrive(p_full);
        }
        pv_pending = true;
        // last tile of a set: wait for its PV, finish the set
        if (kt == n_kt - 1) {
          mbar_wait(o_full, (uint32_t)(j & 1));
          tc_fence_after();
          uint32_t r[32];
#pragma unroll
          for (int c0 = 0; c0 < DP; c0 += 16) {
            tmem_ld_x16(lane_addr + C::O_COL + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o_run[c0 + i] += __uint_as_float(r[i]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(o_empty);
          pv_pending = false;
          const float inv = 1.f / l_run;
#pragma unroll
          for (int d = 0; d < DP; ++d) acc_out[d] += o_run[d] * inv;
        }
      }
    }
    if (q_idx < p.Lq) {
      bf16* op = p.out + ((int64_t)seq * p.Lq + q_idx) * p.ldo + (int64_t)h * p.D;
#pragma unroll
      for (int d = 0; d < DP; d += 8) {
        if (d < p.D) {
          Vec16<bf16> pk;
          pk.pack(acc_out + d);
          *reinterpret_cast<uint4*>(op + d) = pk.raw;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}
}  // namespace

namespace {
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
  dim3 grid((p.Lq + BMQ - 1) / BMQ, p.H, n_seq);
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
  // number of K/V sequences addressable: kv_map may point anywhere inside [0, n_seq) for self/cross-view;
  // for cross-attention the K/V tensor has its own n_seq·Lk rows
  const int bnk = Lk <= 64 ? 64 : 128;
  CUtensorMap tq, tk, tv;
  int rc = mdv2_make_tmap_3d_heads(&tq, q, (uint64_t)n_seq * Lq, H, D, ldq, BMQ);
  if (rc) return rc;
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
