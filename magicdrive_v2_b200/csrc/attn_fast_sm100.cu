// attn_fast_sm100.cu — the 128-key flash-attention kernel of the spatial self / cross-view attention
// (tcgen05 / TMEM / TMA; bf16 in, fp32 softmax, bf16 out).  Replaces flash_attn_func / xformers / the eager softmax
// path and the cross-view gather + sum (reference blocks.py:456-504,607-613; magicdrive_stdit3.py:158-190,303-322).
//
// Same data path as attn_sm100.cu (one CTA = one (sequence, head) x two 128-query tiles; Q in TMEM, S = Q·K^T and
// O += P·V as TS MMAs, P written over the S columns it came from, K/V through a 3-stage TMA ring), but the softmax
// is organised for the unit that bounds this kernel at head_dim 72 — the exponentials (288 FLOP per exponential:
// MUFU.EX2 and the FMA-pipe work around it take ~7.5 clk per score per SM sub-partition, tools/probes/softmax_probe.cu,
// against 5 clk of tensor pipe):
//
//  * SIXTEEN softmax warps, four per SM sub-partition: warp (t, hf) owns the 64-key half `hf` of every K/V tile of
//    query tile t (thread = query row = TMEM lane).  Its loop  wait S^hf(j) -> ld -> exp -> st P -> arrive  contains
//    the tensor-core round trip P·V^hf(j), Q·K^hf(j+1) (600-900 clk), but the other three warps of the sub-partition
//    are in their exponentials meanwhile, the two halves of a tile naturally half a period apart.  (Round 1 ran two
//    warps per sub-partition that each did both halves: 1500 clk per 2 x 64 scores against ~1000 of pipe time, the
//    rest was the chain wait / ld / st / arrive with nobody to fill it.)
//  * No running maximum.  Two warps share a row, and an online-softmax reference that moves would have to be agreed
//    between them every step.  Instead the reference m of a row is FIXED per K/V set: the maximum of the first 32
//    scores each of the two warps sees, exchanged once through shared memory.  p = 2^(s·c − m·c) is exact for any m
//    (fp32 / bf16 are scale free) as long as nothing overflows, and a score would have to exceed the reference by 2^80
//    for that.  The row sum tells: if l >= 2^80 (or NaN) for any row of the CTA, the CTA REPEATS the set with the
//    references of those rows raised by 90 (log2 units) — a handful of CTA-uniform retries in the worst case, none for
//    attention logits (tests: test_attention_tcgen05_rescale_path drives it with logit spreads of hundreds).  With
//    the maximum gone the inner loop is FFMA2 / MUFU.EX2 / FADD2 / F2FP only, and the O accumulator is never rescaled.
//  * The last K/V tile's second half is skipped entirely when it holds no keys (Lk mod 128 in 1..64, e.g. 1600).
//
// Round-2 note: two epilogue changes were tried on top of this file and REVERTED — the earlier sets' partial sums loaded
// ahead of the wait for O, and the last set's rows leaving through shared memory + one bulk copy per row (the rows are
// written with 32 different lines per store instruction).  Both shorten "O complete -> CTA exit" in the life-line trace,
// but the extra live registers changed the allocation of the whole kernel (96 registers, no slack) and the same-box A/B
// against this version was 2.7 % slower at S = 5600, 6 % slower with two K/V sets (tools: MDV2_LIB_FILE A/B).
#include <cstdlib>

#include "attn_common.cuh"

using namespace sm100;
using namespace attn;

namespace {
constexpr int BNK = 128;          // keys per K/V tile
constexpr int HB = 64;            // keys per half-step = S columns owned by one softmax warp
constexpr int CH = 32;            // keys per tcgen05.ld / exponential chunk
constexpr int KV_STAGES = 3;
constexpr int QT = 2;             // query tiles per CTA
// warps 0..15 softmax, 16 TMA producer, 17 / 18 MMA issuer of query tile 0 / 1.  Measured (profiles/attn_trace_r02.txt):
// a control warp costs the four softmax warps of its sub-partition 15-30 % of their speed — the issuer's tcgen05.mma
// instructions issue in lock-step with the tensor pipe (~42 clk each, 18 per K/V tile) and hold the sub-partition's
// dispatch port meanwhile — whichever warp ids they carry and however they wait (spinning try_wait, try_wait with a
// suspend hint, nanosleep): the sub-partition WITHOUT a control warp finishes its 64 scores in ~1100 clk, the others
// in 1300-1500, and the slowest sets the period.
constexpr int N_SM_WARPS = 16;
constexpr int TMA_WARP = 16, MMA_WARP0 = 17;
constexpr int NTHREADS = (N_SM_WARPS + 3) * 32;
constexpr int TMEM_COLS = 512;    // S_t at 128·t, O_t at 256 + 128·t, Q_t (bf16 pairs) at 336 + 128·t
constexpr float RETRY_SHIFT = 90.f;             // log2 units added to an overflowed row's reference
constexpr float L_LIMIT = 1.2089258196146292e24f;   // 2^80
constexpr int MAX_RETRY = 6;
__host__ __device__ constexpr int s_col(int t) { return t * 128; }
__host__ __device__ constexpr int o_col(int t) { return 256 + t * 128; }
__host__ __device__ constexpr int q_col(int t) { return 336 + t * 128; }

template <int DP>
struct FCfg {
  static constexpr int NCH = DP > 64 ? 2 : 1;                 // 64-wide d-chunks per row
  static constexpr int Q_BYTES = NCH * BMQ * 128;
  static constexpr int KV_BYTES = NCH * BNK * 128;
  static constexpr int XCH_FLOATS = 3 * QT * 2 * BMQ;         // reference exchange + two row-sum exchanges
  static constexpr int SMEM = KV_STAGES * 2 * KV_BYTES + 512 + XCH_FLOATS * 4 + 1024;
  static_assert(Q_BYTES == KV_BYTES, "Q tile must fit a K/V stage");
};

template <int DP, int PF, bool TRACE>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_fast_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const AParams p) {
  using C = FCfg<DP>;
  extern __shared__ uint8_t smem_raw[];
#ifdef MDV2_SMEM_GENERIC                               // A/B hook: the previous form (loses the address space)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
#else
  // pointer arithmetic on the __shared__ array keeps the address space (LDS / STS instead of generic LD.E / ST.E)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
#endif
  uint8_t* sK = smem;                                   // stage s at sK + s·KV_BYTES
  uint8_t* sV = sK + KV_STAGES * C::KV_BYTES;
  uint8_t* s_end = sV + KV_STAGES * C::KV_BYTES;
  // Q is copied into TMEM before the first MMA: its two tiles sit on the last K (tile 0) / V (tile 1) stage
  uint8_t* sQ0 = sK + (KV_STAGES - 1) * C::KV_BYTES;
  uint8_t* sQ1 = sV + (KV_STAGES - 1) * C::KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_end);
  uint64_t* q_full = bars;                              // 1
  uint64_t* k_full = bars + 1;                          // KV_STAGES
  uint64_t* k_empty = k_full + KV_STAGES;
  uint64_t* v_full = k_empty + KV_STAGES;
  uint64_t* v_empty = v_full + KV_STAGES;
  uint64_t* s_full = v_empty + KV_STAGES;               // [query tile][key half]
  uint64_t* p_full = s_full + 4;                        // [query tile][key half]
  uint64_t* o_full = p_full + 4;                        // [query tile]: the last P·V of the pass has completed
  uint64_t* qt_full = o_full + 2;                       // [query tile]: Q_t copied from shared memory into TMEM
  uint64_t* set_done = qt_full + 2;                     // every softmax warp has posted its verdict on the pass
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(set_done + 1);
  uint32_t* retry_flag = tmem_slot + 1;                 // 1 + index of the last pass that overflowed
  float* xch_m = reinterpret_cast<float*>(s_end + 512); // [tile][half][row]: reference exchange
  float* xch_l = xch_m + QT * 2 * BMQ;                  // [pass parity][tile][half][row]: row-sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qp = blockIdx.x, h = blockIdx.y, seq = blockIdx.z;
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int nh_last = p.Lk - (n_kt - 1) * BNK > HB ? 2 : 1;      // half-steps of the last K/V tile
  const int q_row0 = qp * QT * BMQ;                     // first query row of this CTA within the sequence
  const int n_qt = (q_row0 + BMQ < p.Lq) ? 2 : 1;       // the second tile may lie entirely past Lq
  const bool tracing = TRACE && p.trace != nullptr && qp == 3 && h == 1 && seq == (int)gridDim.z / 2;
  long long* cta_line = nullptr;                        // TRACE: life line of EVERY CTA (global timer, ns)
  if constexpr (TRACE) {
    if (p.trace != nullptr)
      cta_line = p.trace + TRACE_CTA_OFF +
                 (int64_t)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) * TRACE_CTA_EV;
    if (cta_line && threadIdx.x == 0) {
      cta_line[0] = gtime_ns();
      cta_line[1] = smid();
    }
  }
  auto mark = [&](int j, int ev) {
    if constexpr (TRACE) {
      if (tracing && lane == 0 && j < TRACE_STEPS) p.trace[(warp * TRACE_STEPS + j) * TRACE_EV + ev] = clk64();
    }
  };

  if (warp == TMA_WARP && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], n_qt);                     // one arrival per query tile
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], n_qt);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&o_full[i], 1);
      mbar_init(&qt_full[i], 4);
    }
    mbar_init(set_done, 8 * n_qt);
    *retry_flag = 0;
    fence_mbar_init();
  }
  if (warp == MMA_WARP0) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Every role walks the same sequence of PASSES (a pass = one attempt at one K/V set) and learns at the end of each
  // whether it has to be repeated.
  auto verdict = [&](int pass, int attempt) -> bool {
    mbar_wait(set_done, (uint32_t)(pass & 1));
    const uint32_t f = *reinterpret_cast<volatile uint32_t*>(retry_flag);
    return f == (uint32_t)(pass + 1) && attempt < MAX_RETRY;
  };

  if (warp == TMA_WARP) {
    // ================= TMA producer =================
    if (lane == 0) {
      mbar_expect_tx(q_full, n_qt * C::Q_BYTES);
      for (int t = 0; t < n_qt; ++t) {
        const int row = seq * p.Lq + q_row0 + t * BMQ;
#pragma unroll
        for (int c = 0; c < C::NCH; ++c) tma_load_3d((t ? sQ1 : sQ0) + c * BMQ * 128, &tq, q_full, c * 64, h, row);
      }
      int jg = 0, pass = 0;
      for (int set = 0; set < p.n_sets; ++set) {
        const int kseq = p.kv_map ? p.kv_map[seq * p.n_sets + set] : seq;
        for (int attempt = 0;; ++attempt) {
          for (int kt = 0; kt < n_kt; ++kt, ++jg) {
            const int st = jg % KV_STAGES;
            const uint32_t ph = (uint32_t)((jg / KV_STAGES) & 1);
            const int row = kseq * p.Lk + kt * BNK;
            if (jg == KV_STAGES - 1)                     // first use of the stage that held Q: Q must be in TMEM
              for (int t = 0; t < n_qt; ++t) mbar_wait_parked(&qt_full[t], 0, 4000);
            mbar_wait_parked(&k_empty[st], ph ^ 1, 4000);
            mark(jg, 0);
            mbar_expect_tx(&k_full[st], C::KV_BYTES);
#pragma unroll
            for (int c = 0; c < C::NCH; ++c)
              tma_load_3d(sK + st * C::KV_BYTES + c * BNK * 128, &tk, &k_full[st], c * 64, h, row);
            mbar_wait_parked(&v_empty[st], ph ^ 1, 4000);
            mark(jg, 1);
            mbar_expect_tx(&v_full[st], C::KV_BYTES);
#pragma unroll
            for (int c = 0; c < C::NCH; ++c)
              tma_load_3d(sV + st * C::KV_BYTES + c * BNK * 128, &tv, &v_full[st], c * 64, h, row);
          }
          const bool again = verdict(pass, attempt);
          ++pass;
          if (!again) break;
        }
      }
    }
  } else if (warp > TMA_WARP) {
    // ================= MMA issuers: warp 17 -> query tile 0, warp 18 -> query tile 1 =================
    // The whole warp runs the warp-uniform control flow and one elected lane issues the tcgen05 instructions of a
    // half-step back to back (descriptors are base + constant, the 14-bit address field never carries).
    const int t = warp - MMA_WARP0;
    if (t < n_qt) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, HB, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);     // A = P from TMEM, B = V is MN-major
      const uint64_t dk0 = make_desc_k_sw128(smem_u32(sK));
      const uint64_t dv0 = make_desc_mn_sw128(smem_u32(sV), BNK * 128, 1024);
      const uint32_t s_tm = tmem_base + s_col(t), o_tm = tmem_base + o_col(t), q_tm = tmem_base + q_col(t);
      // S_t^hf = Q_t · K[hf·64 .. +63]^T -> TMEM columns s_col(t) + hf·64 .. +63, A = Q_t from TMEM (one lane calls this)
      auto issue_qk = [&](int hf, int st) {
        const uint64_t db = dk0 + (uint64_t)(st * (C::KV_BYTES >> 4) + hf * ((HB * 128) >> 4));
#pragma unroll
        for (int ks = 0; ks < DP / 16; ++ks) {
          const uint64_t ob = (uint64_t)((ks / 4) * ((BNK * 128) >> 4) + (ks % 4) * 2);
          umma_bf16_ts(s_tm + hf * HB, q_tm + ks * 8, db + ob, idesc_qk, ks != 0);
        }
        umma_commit(&s_full[t * 2 + hf]);
      };
      mbar_wait_parked(&qt_full[t], 0, 1000);
      int jg = 0, pass = 0;
      uint32_t cp[2] = {0, 0};                           // P^hf half-steps consumed so far (parity of p_full)
      for (int set = 0; set < p.n_sets; ++set) {
        for (int attempt = 0;; ++attempt) {
          // first tile of the pass: both halves of S (behind every MMA of the previous pass: in-order pipe)
          mbar_wait_parked(&k_full[jg % KV_STAGES], (uint32_t)((jg / KV_STAGES) & 1), 1000);
          tc_fence_after();
          if (elect_one()) {
            const int nh0 = n_kt == 1 ? nh_last : 2;
            for (int hf = 0; hf < nh0; ++hf) issue_qk(hf, jg % KV_STAGES);
          }
          __syncwarp();
          for (int kt = 0; kt < n_kt; ++kt, ++jg) {
            const int st = jg % KV_STAGES, st1 = (jg + 1) % KV_STAGES;
            const bool more = kt + 1 < n_kt;
            const int nh = kt == n_kt - 1 ? nh_last : 2;
            const int nh1 = kt + 1 == n_kt - 1 ? nh_last : 2;             // half-steps of the next tile
            mark(jg, 5);
            mbar_wait_parked(&v_full[st], (uint32_t)((jg / KV_STAGES) & 1), 1000);
            mark(jg, 6);
            if (more) mbar_wait_parked(&k_full[st1], (uint32_t)(((jg + 1) / KV_STAGES) & 1), 1000);
            const uint64_t dv = dv0 + (uint64_t)(st * (C::KV_BYTES >> 4));
            mark(jg, 0);
            for (int hf = 0; hf < nh; ++hf) {
              mbar_wait_parked(&p_full[t * 2 + hf], cp[hf] & 1, 1000);
              ++cp[hf];
              tc_fence_after();
              mark(jg, 1 + 2 * hf);
              if (elect_one()) {
                // O_t (+)= P_t^hf · V[hf·64 ..]: P^hf sits on the first 32 columns of S_t^hf
#pragma unroll
                for (int ks = 0; ks < HB / 16; ++ks)
                  umma_bf16_ts(o_tm, s_tm + hf * HB + ks * 8, dv + (uint64_t)((hf * HB + ks * 16) * 128 >> 4), idesc_pv,
                               (ks != 0) || (hf != 0) || (kt != 0));
                if (!more && hf == nh - 1) umma_commit(&o_full[t]);
                if (more && hf < nh1) issue_qk(hf, st1);   // S_t^hf of the next tile over the P it just consumed
              }
              __syncwarp();
              mark(jg, 2 + 2 * hf);
            }
          }
          const bool again = verdict(pass, attempt);
          ++pass;
          if (!again) break;
        }
      }
    }
  } else if (warp < N_SM_WARPS && (warp >> 3) < n_qt) {
    // ================= softmax warps: warp (t, hf), thread = query row =================
    const int quarter = warp & 3;                      // TMEM lane quarter this warp may touch (= its sub-partition)
    const int role = warp >> 2;                        // 0..3 among the four warps of a quarter
    const int t = role >> 1, hf = role & 1;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t sh_addr = lane_addr + s_col(t) + hf * HB, o_addr = lane_addr + o_col(t);
    const int q_idx = q_row0 + t * BMQ + row;
    const float c = p.scale_log2;
    const int pair_bar = 1 + t * 4 + quarter;          // named barrier of the two warps that share these rows
    bf16* orow = p.out + ((int64_t)seq * p.Lq + q_idx) * p.ldo + (int64_t)h * p.D;
    bf16* frow = orow;                                 // where the final (last set) output row goes
    if (p.n_peer > 0 && q_idx < p.Lq) {
      const int owner = q_idx / p.rows_per_rank;
      frow = p.peer_out[owner] + ((int64_t)seq * p.rows_per_rank + (q_idx - owner * p.rows_per_rank)) * p.peer_ld +
             p.peer_col + (int64_t)h * p.D;
    }
    if (hf == 0) {
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
      const uint32_t q_addr = lane_addr + q_col(t);
      tmem_st_x32(q_addr, qw);
      if constexpr (DP / 2 > 32) tmem_st_x8(q_addr + 32, qw + 32);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&qt_full[t]);
    }
    // p = 2^(s·c − moff) of one 32-key chunk, row sum alongside.  PF of every 8 pairs take the polynomial 2^x on the
    // FMA pipe instead of MUFU.EX2.
    auto exp_chunk = [&](const uint32_t (&s)[CH], float moff, uint32_t (&pk)[CH / 2], float& l_acc) {
      const float2 c2 = make_float2(c, c), nm2 = make_float2(-moff, -moff);
      float2 la = make_float2(0.f, 0.f), lb = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < CH; i += 4) {
        const float2 x01 = ffma2(make_float2(__uint_as_float(s[i]), __uint_as_float(s[i + 1])), c2, nm2);
        const float2 x23 = ffma2(make_float2(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])), c2, nm2);
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
      l_acc += (la.x + la.y) + (lb.x + lb.y);
    };
    auto mask_tail = [&](uint32_t (&s)[CH], int left) {   // ragged last tile: keys past Lk do not exist
      if (left < CH) {
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (i >= left) s[i] = 0xff800000u;  // -inf
      }
    };
    const int n_mine = hf == 0 ? n_kt : n_kt - (nh_last == 1 ? 1 : 0);   // half-steps of this warp per pass
    const int xi = (t * 2 + hf) * BMQ + row, xo = (t * 2 + (hf ^ 1)) * BMQ + row;   // own / partner exchange slot
    uint32_t cnt = 0;                                    // half-steps done so far (parity of s_full)
    int jg = 0, pass = 0;
    float moff = 0.f;                                    // the row's reference, log2 units
    for (int set = 0; set < p.n_sets; ++set) {
      for (int attempt = 0;; ++attempt) {
        float l_part = 0.f;
        for (int kt = 0; kt < n_mine; ++kt, ++cnt) {
          mbar_wait(&s_full[t * 2 + hf], cnt & 1);
          tc_fence_after();
          mark(cnt, 0);
          if constexpr (TRACE) {
            if (cta_line && warp == 0 && lane == 0 && cnt == 0) cta_line[2] = gtime_ns();
          }
          const int kv_left = kt == n_kt - 1 ? p.Lk - kt * BNK - hf * HB : HB;   // keys of this half that exist
          // two 32-key chunks; P chunk q goes to columns 16q..16q+15 of the half, over scores already in registers.
          // (Tried: four 16-key chunks double-buffered through registers so that the tcgen05.ld of the next chunk
          // flies under the exponentials — slower, 880 vs 930 TF/s: more TMEM instructions, longer MMA issue.)
          uint32_t sr[CH], pk[CH / 2];
          const bool release = quarter == 0 && hf == (kt == n_kt - 1 ? nh_last : 2) - 1;
#pragma unroll
          for (int q = 0; q < HB / CH; ++q) {
            tmem_ld_x32(sh_addr + q * CH, sr);
            if (q == 0 && release && lane == 0) {
              // the LAST half of S(tile jg + kt) is complete, and the tensor pipe completes in issue order: every
              // Q·K^T on this K tile and every P·V on the previous V tile is done — hand both stages back to the TMA
              // producer (under the latency of the load above)
              const int j = jg + kt;
              mbar_arrive(&k_empty[j % KV_STAGES]);
              if (j > 0) mbar_arrive(&v_empty[(j - 1) % KV_STAGES]);
            }
            tmem_ld_wait();
            mask_tail(sr, kv_left - q * CH);
            if (q == 0 && kt == 0 && attempt == 0) {
              // the set's reference: max over the first 32 scores of both warps of the row
              float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
              for (int i = 0; i < CH; i += 4) {
                mx0 = fmax3(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
                mx1 = fmax3(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
              }
              const float mine = fmaxf(mx0, mx1);
              xch_m[xi] = mine;
              named_bar_sync(pair_bar, 64);
              moff = fmaxf(mine, xch_m[xo]) * c;
            }
            exp_chunk(sr, moff, pk, l_part);
            tmem_st_x16(sh_addr + q * (CH / 2), pk);
            if (q == 0) mark(cnt, 1);
          }
          mark(cnt, 2);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[t * 2 + hf]);
          mark(cnt, 3);
          if constexpr (TRACE) {
            if (cta_line && warp == 0 && lane == 0) cta_line[3] = gtime_ns();
          }
        }
        jg += n_kt;
        // end of the pass: the row sum of both halves, and the CTA-wide verdict
        float* xl = xch_l + (pass & 1) * (QT * 2 * BMQ);
        xl[xi] = l_part;
        named_bar_sync(pair_bar, 64);
        const float l_row = l_part + xl[xo];
        const bool ok = l_row < L_LIMIT;                 // false for NaN, too
        if (!ok) atomicMax(retry_flag, (uint32_t)(pass + 1));
        __syncwarp();
        if (lane == 0) mbar_arrive(set_done);
        const bool again = verdict(pass, attempt);
        ++pass;
        if (again) {
          if (!ok) moff += RETRY_SHIFT;
          continue;
        }
        // normalise, round to bf16, add to what earlier sets left in `out`: this warp takes half of the columns
        mbar_wait(&o_full[t], (uint32_t)((pass - 1) & 1));
        tc_fence_after();
        if constexpr (TRACE) {
          if (cta_line && warp == 0 && lane == 0) cta_line[4] = gtime_ns();
        }
        const float inv = 1.f / l_row;
        constexpr int NU = DP / 16;                      // 8-column units per warp
        uint32_t r[NU * 8];
#pragma unroll
        for (int u = 0; u < NU; ++u) tmem_ld_x8(o_addr + hf * (DP / 2) + u * 8, r + u * 8);
        tmem_ld_wait();
        if (q_idx < p.Lq) {
#pragma unroll
          for (int u = 0; u < NU; ++u) {
            const int d = hf * (DP / 2) + u * 8;
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
        tc_fence_before();
        break;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
  if constexpr (TRACE) {
    if (cta_line && threadIdx.x == 0) cta_line[5] = gtime_ns();
  }
}

template <int DP, int PF, bool TRACE>
int launch_fast(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p, int n_seq,
                cudaStream_t st) {
  constexpr int SMEM = FCfg<DP>::SMEM;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_fast_kernel<DP, PF, TRACE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("attention_fast: cannot set %d B dynamic smem: %s", SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    configured = true;
  }
  dim3 grid((p.Lq + QT * BMQ - 1) / (QT * BMQ), p.H, n_seq);
  attn_fast_kernel<DP, PF, TRACE><<<grid, NTHREADS, SMEM, st>>>(tq, tk, tv, p);
  MDV2_CHECK_LAUNCH("attention_fast");
  return MDV2_OK;
}
}  // namespace

// called by mdv2_attention_sm100 (attn_sm100.cu) for K/V sequences longer than 64 keys
int mdv2_attention_fast_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                               int n_seq, int D, bool trace, cudaStream_t st) {
  // polynomial pairs per 8 (tuning hook, tools/bench_attn.py).  Alone at full clocks 3 is the fastest setting by ~1 %; inside
  // the power-capped step (1.40-1.55 GHz) 2 is: 1127 ms of attention per step against 1146-1151 (3), 1165 (0), 1212 (4) in
  // a same-box sweep (profiles/attn_pf_instep_r02.txt) — the polynomial's FMA-pipe instructions cost clock under the cap
  static const int pf = [] {
    const char* e = getenv("MDV2_ATTN_PF");
    return e ? atoi(e) : 2;
  }();
  if (D == 72) {
    if (trace) return launch_fast<80, 3, true>(tq, tk, tv, p, n_seq, st);
    switch (pf) {
      case 0: return launch_fast<80, 0, false>(tq, tk, tv, p, n_seq, st);
      case 2: return launch_fast<80, 2, false>(tq, tk, tv, p, n_seq, st);
      case 3: return launch_fast<80, 3, false>(tq, tk, tv, p, n_seq, st);
      case 4: return launch_fast<80, 4, false>(tq, tk, tv, p, n_seq, st);
      default: MDV2_REQUIRE(false, "attention_fast: MDV2_ATTN_PF must be 0, 2, 3 or 4");
    }
  }
  return launch_fast<64, 0, false>(tq, tk, tv, p, n_seq, st);
}
