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

#include "attn_common.cuh"

using namespace sm100;
using namespace attn;

namespace {
// CTA shape.  128-key tiles (self / cross-view attention): two query tiles per CTA, 11 warps (TMA, MMA of tile 0,
// 8 softmax warps, MMA of tile 1), all 512 TMEM columns: S_t (and P_t over it) at 128·t, O_t at 256 + 128·t,
// Q_t (bf16 pairs, DP/2 <= 40 columns) at 336 + 128·t in the gaps behind O_t.
// 64-key tiles (cross-attention onto <= 64 condition tokens, a single K/V tile): the CTA is one latency chain
// (load, one Q·K^T, one softmax, one P·V, store), so it takes ONE query tile, 6 warps, 256 TMEM columns
// (S at 0, O at 64, Q at 144) and 96 KB of shared memory — two CTAs per SM overlap each other's chain.
template <int BNK>
struct AShape {
  static constexpr int QT = BNK == 64 ? 1 : 2;                       // query tiles per CTA
  static constexpr int NTHREADS = QT == 2 ? 352 : 192;
  static constexpr int TMEM_COLS = QT == 2 ? 512 : 256;
  static constexpr int CTAS_PER_SM = QT == 2 ? 1 : 2;
  __host__ __device__ static constexpr int s_col(int t) { return QT == 2 ? t * 128 : 0; }
  __host__ __device__ static constexpr int o_col(int t) { return QT == 2 ? 256 + t * 128 : 64; }
  __host__ __device__ static constexpr int q_col(int t) { return QT == 2 ? 336 + t * 128 : 144; }
};
constexpr float RESCALE_LOG2 = 8.f;
constexpr int ATTN_PF_DEFAULT = 2;   // polynomial 2^x pairs per 8 pairs (0, 2, 3, 4)

template <int DP, int BNK>
struct ACfg {
  static constexpr int NCH = DP > 64 ? 2 : 1;                 // 64-wide d-chunks per row
  static constexpr int Q_BYTES = NCH * BMQ * 128;             // one query tile
  static constexpr int KV_BYTES = NCH * BNK * 128;            // one K (or V) tile
  // Q is copied into TMEM before the first MMA, so with 128-key tiles (Q_BYTES == KV_BYTES) its two shared-memory
  // tiles are recycled as the third K / V stage; the 64-key flavour (single K/V tile) keeps Q apart.
  static constexpr bool ALIAS_Q = BNK == 128;
  static constexpr int KV_STAGES = ALIAS_Q ? 3 : 2;
  static constexpr int SMEM = KV_STAGES * 2 * KV_BYTES + (ALIAS_Q ? 0 : AShape<BNK>::QT * Q_BYTES) + 320 + 1024;
  static_assert(!ALIAS_Q || Q_BYTES == KV_BYTES, "Q tile must fit a K/V stage");
};

// Measured (tools/probes/pipe_probe.cu, clk per warp instruction and SM sub-partition): MUFU.EX2 8, FFMA2 3, FADD2 /
// FMNMX3 / F2FP 2, FFMA / FADD 1.  The packed forms cost pipe time but save issue slots, and the loop below is
// bound by the MUFU pipe (64 exponentials x 2 resident warps x 8 clk ~ the 1000 clk a half-step takes): a scalar
// rewrite of the loop was 3 % slower, and moving 1/8..1/2 of the exponentials to this polynomial is neutral to
// slightly negative at D = 72 (sweep in profiles/), because the issue port then becomes the limit.
// What bounds the kernel (profiles/attn_trace_r01.txt): a softmax warp's half-step is ONE serial chain — wait S 150,
// tcgen05.ld 100, ~14 clk of issue per score (the pipe costs above add up within a warp) x 64, tcgen05.st 100 — of
// ~1400 clk, and only two such chains fit an SM sub-partition (TMEM holds two 128-row tiles).  Tried and measured:
// ping-pong turns between the two tiles (kept, +4 %: one tile's MMAs run under the other's exponentials); splitting a
// row over two warps (16 softmax warps at 96 registers, row statistics exchanged through shared memory + a named
// barrier per half-step): each chain got 30 % shorter but paid 140 clk of exchange, 725 vs 747 TF/s, removed
// (profiles/attn_trace_rowsplit_r01.txt).
// pair index (0..BNK/2) -> evaluated by the polynomial?  PF of every 8 pairs, spread evenly.

template <int DP, int BNK, int PF = 0, bool TRACE = false>
__global__ void __launch_bounds__(AShape<BNK>::NTHREADS, AShape<BNK>::CTAS_PER_SM)
attn_sm100_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                  const __grid_constant__ CUtensorMap tv, const AParams p) {
  using C = ACfg<DP, BNK>;
  using SH = AShape<BNK>;
  constexpr int QT = SH::QT, TMEM_COLS = SH::TMEM_COLS;
  extern __shared__ uint8_t smem_raw[];
#ifdef MDV2_SMEM_GENERIC                               // A/B hook: the previous form (loses the address space)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
#else
  // pointer arithmetic on the __shared__ array keeps the address space (LDS / STS instead of generic LD.E / ST.E)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
#endif
  constexpr int KV_STAGES = C::KV_STAGES;
  uint8_t* sK = smem;                                   // stage s at sK + s·KV_BYTES
  uint8_t* sV = sK + KV_STAGES * C::KV_BYTES;
  uint8_t* s_end = sV + KV_STAGES * C::KV_BYTES;
  // Q tile t: over the last K (t = 0) / V (t = 1) stage when aliased, else behind the rings
  uint8_t* sQ0 = C::ALIAS_Q ? sK + (KV_STAGES - 1) * C::KV_BYTES : s_end;
  uint8_t* sQ1 = C::ALIAS_Q ? sV + (KV_STAGES - 1) * C::KV_BYTES : s_end + C::Q_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(C::ALIAS_Q ? s_end : s_end + QT * C::Q_BYTES);
  uint64_t* q_full = bars;                              // 1
  uint64_t* k_full = bars + 1;                          // KV_STAGES
  uint64_t* k_empty = k_full + KV_STAGES;
  uint64_t* v_full = k_empty + KV_STAGES;
  uint64_t* v_empty = v_full + KV_STAGES;
  uint64_t* s_full = v_empty + KV_STAGES;               // [query tile][key half]
  uint64_t* p_full = s_full + 4;                        // [query tile][key half]
  uint64_t* o_full = p_full + 4;                        // [query tile][key half]: P·V of that half-step has completed
  uint64_t* qt_full = o_full + 4;                       // [query tile]: Q_t copied from shared memory into TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(qt_full + 2);
  float* pp_zero = reinterpret_cast<float*>(tmem_slot + 1);   // ping-pong turn: a 0.0f read behind the barrier ...
  float* pp_sink = pp_zero + 1;                               // ... and one word per warp written ahead of the arrive
  constexpr int mma1_warp = 10;                              // MMA issuer of query tile 1
  constexpr int HB = 64;                                // keys per half-step (S columns handed over at a time)
  constexpr int NH = BNK / HB;                          // half-steps per K/V tile

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int tma_warp = 0, mma_warp = 1;             // warp 1 also owns the TMEM allocation
  const int qp = blockIdx.x, h = blockIdx.y, seq = blockIdx.z;
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int n_tiles = n_kt * p.n_sets;
  const int q_row0 = qp * QT * BMQ;                     // first query row of this CTA within the sequence
  const int n_qt = (QT == 2 && q_row0 + BMQ < p.Lq) ? 2 : 1;   // the second tile may lie entirely past Lq
  // debug timeline of one CTA of a middle wave: trace[warp][step][event] = clock64
  const bool tracing = TRACE && p.trace != nullptr && qp == 3 && h == 1 && seq == (int)gridDim.z / 2;
  auto mark = [&](int j, int ev) {
    if constexpr (TRACE) {
      if (tracing && lane == 0 && j < TRACE_STEPS) p.trace[(warp * TRACE_STEPS + j) * TRACE_EV + ev] = clk64();
    }
  };

  if (warp == tma_warp && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], n_qt);                     // one arrival per query tile (its softmax warp 0)
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], n_qt);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
    }
    for (int i = 0; i < 2; ++i) mbar_init(&qt_full[i], 4);
    fence_mbar_init();
  }
  if (warp == mma_warp) tmem_alloc(tmem_slot, TMEM_COLS);
  if (threadIdx.x == 64) *pp_zero = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == tma_warp) {
    // ================= TMA producer =================
    if (lane == 0) {
      mbar_expect_tx(q_full, n_qt * C::Q_BYTES);
      for (int t = 0; t < n_qt; ++t) {
        const int row = seq * p.Lq + q_row0 + t * BMQ;
#pragma unroll
        for (int c = 0; c < C::NCH; ++c) tma_load_3d((t ? sQ1 : sQ0) + c * BMQ * 128, &tq, q_full, c * 64, h, row);
      }
      int j = 0;
      for (int set = 0; set < p.n_sets; ++set) {
        const int kseq = p.kv_map ? p.kv_map[seq * p.n_sets + set] : seq;
        for (int kt = 0; kt < n_kt; ++kt, ++j) {
          const int st = j % KV_STAGES;
          const uint32_t ph = (uint32_t)((j / KV_STAGES) & 1);
          const int row = kseq * p.Lk + kt * BNK;
          if (C::ALIAS_Q && j == KV_STAGES - 1)          // first use of the stage that held Q: Q must be in TMEM
            for (int t = 0; t < n_qt; ++t) mbar_wait_parked(&qt_full[t], 0, 2000);
          mbar_wait_parked(&k_empty[st], ph ^ 1, 2000);
          mbar_expect_tx(&k_full[st], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c)
            tma_load_3d(sK + st * C::KV_BYTES + c * BNK * 128, &tk, &k_full[st], c * 64, h, row);
          mbar_wait_parked(&v_empty[st], ph ^ 1, 2000);
          mbar_expect_tx(&v_full[st], C::KV_BYTES);
#pragma unroll
          for (int c = 0; c < C::NCH; ++c)
            tma_load_3d(sV + st * C::KV_BYTES + c * BNK * 128, &tv, &v_full[st], c * 64, h, row);
        }
      }
    }
  } else if (warp == 1 || (QT == 2 && warp == mma1_warp)) {
    // ================= MMA issuers: warp 1 -> query tile 0, warp 10 -> query tile 1 =================
    // The whole warp runs the warp-uniform control flow and one elected lane issues each tcgen05 instruction
    // (descriptors live in uniform registers; every shared-memory descriptor is base + constant, the 14-bit
    // address field never carries).  One issuer per query tile: the chain  wait P -> P·V -> Q·K^T(next) -> commit
    // of one tile costs ~700 clk per half-step, of which only ~400 is tensor-pipe time (MMAs that accumulate into
    // the same TMEM tile issue in lock-step with their execution, tools/probes/mma_issue_probe.cu); with a single
    // issuer the four chains of a step ran back to back and the softmax warps idled half of the time.
    const int t = warp == 1 ? 0 : 1;
    if (t < n_qt) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, HB, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);     // A = P from TMEM, B = V is MN-major
      const uint64_t dk0 = make_desc_k_sw128(smem_u32(sK));
      const uint64_t dv0 = make_desc_mn_sw128(smem_u32(sV), BNK * 128, 1024);
      const uint32_t s_tm = tmem_base + SH::s_col(t), o_tm = tmem_base + SH::o_col(t), q_tm = tmem_base + SH::q_col(t);
      // S_t^hf(j) = Q_t · K_j[hf·64 .. hf·64+63]^T  ->  TMEM columns t·128 + hf·64 .. +63.  A = Q_t from TMEM: an SS
      // MMA would re-read the 4 KB Q slice from shared memory for every 64-key half and run at the 128 B/clk
      // shared-memory limit (58 clk per 128x64x16 MMA measured, 32 clk of math)
      // (called by ONE lane: the MMAs of a half-step are issued back to back from a single elected region — with an
      // election per instruction the issuer spent ~16 instructions per MMA, and on a sub-partition it shares with
      // busy softmax warps that, not the tensor pipe, set the pace of the chain)
      auto issue_qk = [&](int hf, int st) {
        const uint64_t db = dk0 + (uint64_t)(st * (C::KV_BYTES >> 4) + hf * ((HB * 128) >> 4));
#pragma unroll
        for (int ks = 0; ks < DP / 16; ++ks) {
          const uint64_t ob = (uint64_t)((ks / 4) * ((BNK * 128) >> 4) + (ks % 4) * 2);
          umma_bf16_ts(s_tm + hf * HB, q_tm + ks * 8, db + ob, idesc_qk, ks != 0);
        }
        umma_commit(&s_full[t * 2 + hf]);
      };
      mbar_wait(&qt_full[t], 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      if (elect_one()) {
        for (int hf = 0; hf < NH; ++hf) issue_qk(hf, 0);
      }
      __syncwarp();
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % KV_STAGES, st1 = (j + 1) % KV_STAGES;
        const bool more = j + 1 < n_tiles;
        const bool fresh = (j % n_kt) == 0;              // first tile of a K/V set: O starts over
        const bool last_of_set = (j % n_kt) == n_kt - 1;
        mbar_wait_parked(&v_full[st], (uint32_t)((j / KV_STAGES) & 1), 1000);
        if (more) mbar_wait_parked(&k_full[st1], (uint32_t)(((j + 1) / KV_STAGES) & 1), 1000);
        const uint64_t dv = dv0 + (uint64_t)(st * (C::KV_BYTES >> 4));
        mark(j, 0);
        for (int hf = 0; hf < NH; ++hf) {
          mbar_wait_parked(&p_full[t * 2 + hf], (uint32_t)(j & 1), 1000);
          tc_fence_after();
          mark(j, 1 + 2 * hf);
          // O_t += P_t^hf · V_j[hf·64 ..]: P^hf sits on the first 32 columns of S_t^hf
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < HB / 16; ++ks)
              umma_bf16_ts(o_tm, s_tm + hf * HB + ks * 8, dv + (uint64_t)((hf * HB + ks * 16) * 128 >> 4), idesc_pv,
                           (ks != 0) || (hf != 0) || !fresh);
            // O is only read back at the end of a K/V set; in between, "P·V of a half-step has completed" is implied
            // by the completion of the Q·K^T issued behind it (s_full), so the extra commit (~100 clk of this warp's
            // issue chain each) is only spent on the last tile of a set
            if (last_of_set) umma_commit(&o_full[t * 2 + hf]);
            if (more) issue_qk(hf, st1);                 // S_t^hf(j+1) over the P it just consumed (in-order pipe)
          }
          __syncwarp();
          mark(j, 2 + 2 * hf);
        }
        // K_{j+1} and V_j are handed back to the TMA producer by the softmax warps (see there): a tcgen05.commit
        // costs this warp's issue chain ~100 clk, an mbarrier.arrive from a warp that already knows costs nothing
      }
    }
  } else if (warp < 2 + 4 * QT && ((warp - 2) >> 2) < n_qt) {
    // ================= softmax warps: thread = query row =================
    const int t = (warp - 2) >> 2;
    const int quarter = warp & 3;                      // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_addr = lane_addr + SH::s_col(t), o_addr = lane_addr + SH::o_col(t);
    const int q_idx = q_row0 + t * BMQ + row;
    const float c = p.scale_log2;
    bf16* orow = p.out + ((int64_t)seq * p.Lq + q_idx) * p.ldo + (int64_t)h * p.D;
    bf16* frow = orow;                                 // where the final (last set) output row goes
    if (p.n_peer > 0 && q_idx < p.Lq) {
      const int owner = q_idx / p.rows_per_rank;
      frow = p.peer_out[owner] + ((int64_t)seq * p.rows_per_rank + (q_idx - owner * p.rows_per_rank)) * p.peer_ld +
             p.peer_col + (int64_t)h * p.D;
    }
    {
      // Q_t: this thread's row from the 128-byte-swizzled TMA tile (16-byte unit u of row r sits at u ^ (r % 8))
      // -> TMEM columns q_col(t) .. (+ DP/2), the A-operand layout (lane = row, bf16 pairs along K)
      mbar_wait(q_full, 0);
      uint32_t qw[DP / 2];
      const uint8_t* qrow = (t ? sQ1 : sQ0) + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
      for (int u = 0; u < DP / 8; ++u) {
        const uint4 v4 = *reinterpret_cast<const uint4*>(qrow + (u >> 3) * (BMQ * 128) + (((u & 7) ^ (row & 7)) << 4));
        qw[u * 4 + 0] = v4.x; qw[u * 4 + 1] = v4.y; qw[u * 4 + 2] = v4.z; qw[u * 4 + 3] = v4.w;
      }
      const uint32_t q_addr = lane_addr + SH::q_col(t);
      tmem_st_x32(q_addr, qw);
      if constexpr (DP / 2 > 32) tmem_st_x8(q_addr + 32, qw + 32);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&qt_full[t]);
    }
    // The K/V tile is consumed in NH half-steps of 64 keys: while this warp runs the softmax of one half, the
    // tensor core runs P·V of the previous half and Q·K^T of the same half of the NEXT tile, so S is (almost)
    // always ready when the warp comes back for it and the exponentials never wait for the tensor pipe.
    // Ping-pong of the exponential phases: the two softmax warps of an SM sub-partition (same `quarter`, one per
    // query tile) would otherwise lock in phase — both wait for S, then both queue on the MUFU unit (2 x 64 x 8 clk)
    // while it idles for the ~400 clk of the S wait / load / P store.  A turn token (two named barriers per pair)
    // makes them alternate: one warp's exponentials run under the other's wait + load + store and under the other
    // tile's MMAs.  Tile 0 goes first; tile 1 posts the first token.
    const bool pingpong = QT == 2 && n_qt == 2 && !(p.dbg & 2);
    const int bar_mine = 1 + 2 * quarter + t, bar_other = 1 + 2 * quarter + (1 - t);
    if (pingpong && t == 1) named_bar_arrive(bar_other, 64);
    int j = 0;
    for (int set = 0; set < p.n_sets; ++set) {
      float m_used = -INFINITY, l_run = 0.f;
      for (int kt = 0; kt < n_kt; ++kt, ++j) {
#pragma unroll
        for (int hf = 0; hf < NH; ++hf) {
          const uint32_t sh_addr = s_addr + hf * HB;
          mark(j * NH + hf, 0);
          mbar_wait(&s_full[t * 2 + hf], (uint32_t)(j & 1));
          tc_fence_after();
          mark(j * NH + hf, 1);
          if (hf == NH - 1 && quarter == 0 && lane == 0) {
            // S^{last half}(j) is complete, and the tensor pipe completes in issue order: every Q·K^T on K_j and
            // every P·V on V_{j-1} of this tile is done — hand both stages back to the TMA producer
            mbar_arrive(&k_empty[j % KV_STAGES]);
            if (j > 0) mbar_arrive(&v_empty[(j - 1) % KV_STAGES]);
          }
          uint32_t sr[HB];
          auto load_s = [&]() {
            tmem_ld_x32(sh_addr, sr);
            tmem_ld_x32(sh_addr + 32, sr + 32);
            tmem_ld_wait();
            if (kt == n_kt - 1) {                        // ragged last tile: keys past Lk do not exist
              const int kv_left = p.Lk - kt * BNK - hf * HB;
              if (kv_left < HB) {
#pragma unroll
                for (int i = 0; i < HB; ++i)
                  if (i >= kv_left) sr[i] = 0xff800000u;  // -inf
              }
            }
          };
          load_s();
          mark(j * NH + hf, 2);
          // One pass over the half row: p = 2^(s·c − m_used·c) against the STALE reference maximum while the
          // maximum of these 64 scores is reduced alongside (FMNMX3 in the shadow of the exponentials).  Only when
          // it outgrows the reference by more than 2^RESCALE_LOG2 (or on the first half-step of a set, where there
          // is no reference yet) is the reference moved, O rescaled in TMEM and the pass repeated.  PF of every 8
          // pairs take the polynomial 2^x on the FMA pipe instead of MUFU.EX2 (16/clk/SM: the tightest unit at D=72).
          uint32_t pk[HB / 2];
          float l_tile, m_tile;
          auto exp_pass = [&](float moff) {
            const float2 c2 = make_float2(c, c), nm2 = make_float2(-moff, -moff);
            float2 la = make_float2(0.f, 0.f), lb = make_float2(0.f, 0.f);
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int i = 0; i < HB; i += 4) {
              const float s0 = __uint_as_float(sr[i]), s1 = __uint_as_float(sr[i + 1]);
              const float s2 = __uint_as_float(sr[i + 2]), s3 = __uint_as_float(sr[i + 3]);
              mx0 = fmax3(mx0, s0, s1);
              mx1 = fmax3(mx1, s2, s3);
              const float2 x01 = ffma2(make_float2(s0, s1), c2, nm2);
              const float2 x23 = ffma2(make_float2(s2, s3), c2, nm2);
              float2 e01, e23;
              if (poly_pair(i / 2, PF)) e01 = ex2_poly2(x01);
              else e01 = make_float2(ex2(x01.x), ex2(x01.y));
              if (poly_pair(i / 2 + 1, PF)) e23 = ex2_poly2(x23);
              else e23 = make_float2(ex2(x23.x), ex2(x23.y));
              la = fadd2(la, e01);
              lb = fadd2(lb, e23);
              __nv_bfloat162 b01 = __floats2bfloat162_rn(e01.x, e01.y), b23 = __floats2bfloat162_rn(e23.x, e23.y);
              pk[i / 2] = *reinterpret_cast<uint32_t*>(&b01);
              pk[i / 2 + 1] = *reinterpret_cast<uint32_t*>(&b23);
            }
            l_tile = (la.x + la.y) + (lb.x + lb.y);
            m_tile = fmaxf(mx0, mx1);
          };
          const bool first = kt == 0 && hf == 0;
          if (first) {                                   // no reference yet: find the maximum first
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int i = 0; i < HB; i += 4) {
              mx0 = fmax3(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
              mx1 = fmax3(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
            }
            m_used = fmaxf(mx0, mx1);
            l_run = 0.f;
          }
          if constexpr (TRACE) asm volatile("" ::"f"(m_used) : "memory");
          mark(j * NH + hf, 3);
          // (ptxas moves arithmetic across bar instructions freely: the offset picks up a 0.0f that is read from
          // shared memory behind the barrier, and the row sum is stored ahead of the arrive, so the exponentials
          // stay inside the turn)
          float moff = m_used * c;
          if (pingpong) moff += named_bar_sync_ld(bar_mine, 64, pp_zero);
          exp_pass(moff);
          if (pingpong && !(t == 1 && set == p.n_sets - 1 && kt == n_kt - 1 && hf == NH - 1))
            named_bar_arrive_st(bar_other, 64, pp_sink + (warp - 2), l_tile);   // (tile 1's last token: no taker)
          if (!first) {
            const bool grow = (m_tile - m_used) * c > RESCALE_LOG2;
            if (__any_sync(0xffffffffu, grow)) {
              const float alpha = grow ? ex2((m_used - m_tile) * c) : 1.f;
              if (grow) m_used = m_tile;
              l_run *= alpha;
              // rescale the running output in TMEM: P·V of the previous half-step must have completed; the next
              // one is not issued before this warp arrives on p_full below.  The tensor pipe completes in issue
              // order, so the Q·K^T that was issued right behind that P·V (its s_full phase) stands in for it;
              // only the very last tile has no Q·K^T behind it and uses the end-of-set o_full commit.
              if (hf == 0) mbar_wait(&s_full[t * 2 + NH - 1], (uint32_t)(j & 1));          // behind P·V^{last}(j-1)
              else if (j + 1 < n_tiles) mbar_wait(&s_full[t * 2 + hf - 1], (uint32_t)((j + 1) & 1));  // behind P·V^{hf-1}(j)
              else mbar_wait(&o_full[t * 2 + hf - 1], (uint32_t)(set & 1));
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
              load_s();                                  // S is still in TMEM (P is stored below); the registers
              exp_pass(m_used * c);                      // were recycled.  Rows that did not grow reproduce P.
            }
          }
          l_run += l_tile;
          if constexpr (TRACE) asm volatile("" ::"f"(l_run), "r"(pk[HB / 2 - 1]) : "memory");
          mark(j * NH + hf, 4);
          tmem_st_x32(sh_addr, pk);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[t * 2 + hf]);
          mark(j * NH + hf, 5);
        }
        if (kt == n_kt - 1) {
          // end of a K/V set: normalise, round to bf16, add to what earlier sets left in `out`
          mbar_wait(&o_full[t * 2 + NH - 1], (uint32_t)(set & 1));
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
                    float gg[8], a[8];
                    prev.unpack(gg);
                    o.unpack(a);
#pragma unroll
                    for (int i = 0; i < 8; ++i) a[i] += gg[i];
                    o.pack(a);
                  }
                  *reinterpret_cast<uint4*>((set == p.n_sets - 1 ? frow : orow) + d) = o.raw;
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
  if (warp == mma_warp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

long long* g_attn_trace = nullptr;

template <int DP, int BNK, int PF = 0, bool TRACE = false>
int launch_attn(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p, int n_seq,
                cudaStream_t st) {
  using C = ACfg<DP, BNK>;
  constexpr int SMEM = C::SMEM;
  static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_sm100_kernel<DP, BNK, PF, TRACE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("attention_sm100: cannot set %d B dynamic smem: %s", SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    configured = true;
  }
  constexpr int QT = AShape<BNK>::QT;
  dim3 grid((p.Lq + QT * BMQ - 1) / (QT * BMQ), p.H, n_seq);
  attn_sm100_kernel<DP, BNK, PF, TRACE><<<grid, AShape<BNK>::NTHREADS, SMEM, st>>>(tq, tk, tv, p);
  MDV2_CHECK_LAUNCH("attention_sm100");
  return MDV2_OK;
}
}  // namespace

// debug hook (not part of include/mdv2.h): device buffer of 10 x TRACE_STEPS x TRACE_EV int64 clock samples
extern "C" void mdv2_debug_attn_trace(void* buf) { g_attn_trace = (long long*)buf; }

int mdv2_attention_fast_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                               int n_seq, int D, bool trace, cudaStream_t st);
int mdv2_attention_pers_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                               int n_seq, int D, bool trace, cudaStream_t st);
int mdv2_attention_cross_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                                int n_seq, int D, const void* q, int64_t ldq, const void* k, int64_t ldk, cudaStream_t st);

bool mdv2_attention_sm100_supported(int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int Lq, int Lk, int H, int D,
                                    int n_sets) {
  if (D != 64 && D != 72) return false;
  if (ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8) return false;
  if (H > 65535 || n_sets < 1) return false;
  return Lq > 0 && Lk > 0;
}

int mdv2_attention_sm100(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                         void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D, float scale,
                         const int32_t* kv_map, int n_sets, cudaStream_t st, const AttnScatter* scatter) {
  MDV2_REQUIRE(mdv2_attention_sm100_supported(ldq, ldk, ldv, ldo, Lq, Lk, H, D, n_sets), "attention_sm100: unsupported shape");
  MDV2_REQUIRE(n_seq <= 65535, "attention_sm100: too many sequences");
  MDV2_REQUIRE(((uintptr_t)q % 16 == 0) && ((uintptr_t)k % 16 == 0) && ((uintptr_t)v % 16 == 0) && ((uintptr_t)out % 16 == 0),
               "attention_sm100: operands must be 16-byte aligned");
  MDV2_REQUIRE(out != nullptr || (scatter != nullptr && n_sets == 1), "attention_sm100: missing output");
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
  p.n_peer = 0;
  if (scatter != nullptr) {
    MDV2_REQUIRE(scatter->n_dst >= 1 && scatter->n_dst <= 8 && scatter->rows_per_rank > 0 &&
                     (int64_t)scatter->n_dst * scatter->rows_per_rank == Lq && scatter->dst_ld % 8 == 0 &&
                     scatter->dst_col % 8 == 0,
                 "attention_sm100: bad output scatter geometry");
    p.n_peer = scatter->n_dst;
    p.rows_per_rank = scatter->rows_per_rank;
    p.peer_ld = scatter->dst_ld;
    p.peer_col = scatter->dst_col;
    for (int i = 0; i < 8; ++i) p.peer_out[i] = i < scatter->n_dst ? (bf16*)scatter->dst[i] : nullptr;
    for (int i = 0; i < scatter->n_dst; ++i)
      MDV2_REQUIRE(scatter->dst[i] != nullptr && (uintptr_t)scatter->dst[i] % 16 == 0, "attention_sm100: bad scatter destination");
  }
  p.trace = g_attn_trace;
  static const int dbg = [] {
    const char* e = getenv("MDV2_ATTN_DBG");
    return e ? atoi(e) : 0;
  }();
  p.dbg = dbg;
  static const int pf = [] {                             // tuning hook: polynomial pairs per 8 (tools/bench_attn.py)
    const char* e = getenv("MDV2_ATTN_PF");
    return e ? atoi(e) : ATTN_PF_DEFAULT;
  }();
  static const int fast = [] {                           // MDV2_ATTN_FAST=0: the exact online-softmax kernel below
    const char* e = getenv("MDV2_ATTN_FAST");
    return e ? atoi(e) : 1;
  }();
  static const int persist = [] {                        // MDV2_ATTN_PERSIST=1: the persistent kernel (attn_pers_sm100.cu), opt-in
    const char* e = getenv("MDV2_ATTN_PERSIST");
    return e ? atoi(e) : 0;
  }();
  static const int cross = [] {                          // MDV2_ATTN_CROSS=0: one query tile per CTA (the kernel below)
    const char* e = getenv("MDV2_ATTN_CROSS");
    return e ? atoi(e) : 1;
  }();
  // <= 64 keys, one K/V set, rows stay on this rank: the persistent single-tile kernel (attn_cross_sm100.cu)
  if (bnk == 64 && cross && n_sets == 1 && kv_map == nullptr && scatter == nullptr && g_attn_trace == nullptr)
    return mdv2_attention_cross_launch(tq, tk, tv, p, n_seq, D, q, ldq, k, ldk, st);
  if (bnk == 128 && fast && persist) return mdv2_attention_pers_launch(tq, tk, tv, p, n_seq, D, g_attn_trace != nullptr, st);
  if (bnk == 128 && fast) return mdv2_attention_fast_launch(tq, tk, tv, p, n_seq, D, g_attn_trace != nullptr, st);
  if (D == 72 && bnk == 128 && g_attn_trace) {
    if (pf == 0) return launch_attn<80, 128, 0, true>(tq, tk, tv, p, n_seq, st);
    return launch_attn<80, 128, ATTN_PF_DEFAULT ? ATTN_PF_DEFAULT : 3, true>(tq, tk, tv, p, n_seq, st);
  }
  if (D == 72) {
    if (bnk == 64) return launch_attn<80, 64>(tq, tk, tv, p, n_seq, st);
    switch (pf) {
      case 0: return launch_attn<80, 128, 0>(tq, tk, tv, p, n_seq, st);
      case 2: return launch_attn<80, 128, 2>(tq, tk, tv, p, n_seq, st);
      case 3: return launch_attn<80, 128, 3>(tq, tk, tv, p, n_seq, st);
      case 4: return launch_attn<80, 128, 4>(tq, tk, tv, p, n_seq, st);
      default: MDV2_REQUIRE(false, "attention_sm100: MDV2_ATTN_PF must be 0, 2, 3 or 4");
    }
  }
  if (bnk == 64) return launch_attn<64, 64>(tq, tk, tv, p, n_seq, st);
  return launch_attn<64, 128>(tq, tk, tv, p, n_seq, st);
}
