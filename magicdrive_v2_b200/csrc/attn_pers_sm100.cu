// attn_pers_sm100.cu — PERSISTENT form of the 128-key flash-attention kernel (attn_fast_sm100.cu): the spatial self /
// cross-view attention of the STDiT3 blocks (reference blocks.py:456-504,607-613; magicdrive_stdit3.py:158-190,303-322).
//
// Why: the life line of every CTA of the one-shot kernel (tools/attn_cta_trace.py, profiles/attn_cta_lifeline_r02.txt)
// shows 2.3-2.6 us from CTA start to the first score tile, 2.6-3.3 us from the last probability tile to CTA exit (8-9 us
// with two K/V sets), and 1.0-1.3 us of empty SM between two CTAs — 5.9 us of a 22 us CTA at 1600 tokens per sequence
// (13 K/V tiles), 7.2 of 66 us at 5600.  Here ONE CTA per SM walks a list of work items (sequence, head, pair of
// 128-query tiles) and every role runs ahead of the item boundary:
//   * the TMA producer never stops: Q of the next item and its first K/V tiles ride in the same 3-stage ring (a Q pair
//     is one ring entry: tile 0 in the K slot, tile 1 in the V slot) while the current item is still in its last tiles;
//   * softmax warp (t, 0) copies Q_t of the next item into TMEM right after its last probability tile (all Q·K^T of
//     the item have completed by then), before it turns to the epilogue;
//   * the MMA issuer puts S = Q·K^T of the next pass's first tile behind the last P·V of this pass, so the scores are
//     waiting when the softmax warps return from normalising / storing O (an `o_empty` barrier keeps the first P·V of
//     the next pass off the accumulator until every warp has read it).
// The fixed-reference softmax needs a second attempt when a row overflows (attn_fast_sm100.cu); a pipeline that runs
// ahead cannot repeat a pass in place, so the FAST phase only records the item in a shared-memory bit mask, and after
// its last item the CTA runs the flagged items again in a CAREFUL phase — the one-shot kernel's protocol: nothing
// crosses a pass boundary, every role waits for the CTA-wide verdict, overflowed rows raise their reference by 90.
#include <cstdlib>

#include "attn_common.cuh"

using namespace sm100;
using namespace attn;

namespace {
constexpr int BNK = 128;          // keys per K/V tile
constexpr int HB = 64;            // keys per half-step = S columns owned by one softmax warp
constexpr int CH = 32;            // keys per tcgen05.ld / exponential chunk
constexpr int KV_STAGES = 3;
constexpr int QT = 2;             // query tiles per item
constexpr int N_SM_WARPS = 16;    // warps 0..15 softmax (t, hf, quarter), 16 TMA producer, 17 / 18 MMA issuer of tile 0 / 1
constexpr int TMA_WARP = 16, MMA_WARP0 = 17;
constexpr int NTHREADS = (N_SM_WARPS + 3) * 32;
constexpr int TMEM_COLS = 512;    // S_t at 128·t, O_t at 256 + 128·t, Q_t (bf16 pairs) at 336 + 128·t
constexpr float RETRY_SHIFT = 90.f;
constexpr float L_LIMIT = 1.2089258196146292e24f;   // 2^80
constexpr int MAX_RETRY = 6;
constexpr int MASK_WORDS = 64;    // redo mask: 2048 items per CTA
__host__ __device__ constexpr int s_col(int t) { return t * 128; }
__host__ __device__ constexpr int o_col(int t) { return 256 + t * 128; }
__host__ __device__ constexpr int q_col(int t) { return 336 + t * 128; }

template <int DP>
struct PCfg {
  static constexpr int NCH = DP > 64 ? 2 : 1;                 // 64-wide d-chunks per row
  static constexpr int Q_BYTES = NCH * BMQ * 128;
  static constexpr int KV_BYTES = NCH * BNK * 128;
  static constexpr int XCH_FLOATS = 3 * QT * 2 * BMQ;         // reference exchange + two row-sum exchanges
  static constexpr int STATE_BYTES = 16 * 64;                 // one schedule record per softmax warp
  static constexpr int SMEM = KV_STAGES * 2 * KV_BYTES + 512 + MASK_WORDS * 4 + XCH_FLOATS * 4 + STATE_BYTES + 1024;
  static_assert(Q_BYTES == KV_BYTES, "a Q tile must fit a K/V stage");
};

// p = 2^(s·c − moff) of one 32-key chunk, row sum alongside.  PF of every 8 pairs take the polynomial 2^x on the FMA pipe
// instead of MUFU.EX2.
template <int PF>
__device__ __forceinline__ void exp_chunk(const uint32_t (&s)[CH], float c, float moff, uint32_t (&pk)[CH / 2], float& l_acc) {
  const float2 c2 = make_float2(c, c), nm2 = make_float2(-moff, -moff);
  float2 la2 = make_float2(0.f, 0.f), lb2 = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < CH; i += 4) {
    const float2 x01 = ffma2(make_float2(__uint_as_float(s[i]), __uint_as_float(s[i + 1])), c2, nm2);
    const float2 x23 = ffma2(make_float2(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])), c2, nm2);
    float2 e01, e23;
    if (poly_pair(i / 2, PF)) e01 = ex2_poly2(x01);
    else e01 = make_float2(ex2(x01.x), ex2(x01.y));
    if (poly_pair(i / 2 + 1, PF)) e23 = ex2_poly2(x23);
    else e23 = make_float2(ex2(x23.x), ex2(x23.y));
    la2 = fadd2(la2, e01);
    lb2 = fadd2(lb2, e23);
    __nv_bfloat162 b01 = __floats2bfloat162_rn(e01.x, e01.y), b23 = __floats2bfloat162_rn(e23.x, e23.y);
    pk[i / 2] = *reinterpret_cast<uint32_t*>(&b01);
    pk[i / 2 + 1] = *reinterpret_cast<uint32_t*>(&b23);
  }
  l_acc += (la2.x + la2.y) + (lb2.x + lb2.y);
}
__device__ __forceinline__ void mask_tail(uint32_t (&s)[CH], int left) {   // ragged last tile: keys past Lk do not exist
  if (left < CH) {
#pragma unroll
    for (int i = 0; i < CH; ++i)
      if (i >= left) s[i] = 0xff800000u;  // -inf
  }
}

// The half-steps of ONE pass of a live softmax warp — its own function, never inlined: 96 registers per thread is the
// cap at 19 warps and the exponential loop wants ~75 of them, so the schedule state of the persistent loop around it
// must not be live in here (inlined, the allocator parked a dozen of those values in local memory and re-read them
// every half-step — and local memory shares the L1 / shared-memory port the MMAs read K and V through: 650 instead
// of 890 TF/s).  The caller's state is saved once per pass at the call; the arguments are a few scalars in registers
// (shared-memory ADDRESSES as 32-bit words, the small integers packed into `cfg`).
//   ctl: bit 0 DONE (no pass left) | 1 IDLE (the tile lies past Lq: hand-shakes only) | 16 rel_all: this lane hands slots back at every half-step (warp
//   (t, 1), quarter 0, lane 0) | 17 rel_last: ... at the last one only (warp (t, 0) when the last tile has one half) |
//   18 first pass of the item (the Q pair's slot goes back at the first tile) | 19 take a new reference | 20-21 ring slot
//   of the first K/V tile | 22-23 ring slot of the Q pair | 24-27 named barrier of the row pair | 28 key half | 29-30 2·tile + half
constexpr uint32_t KE_TO_VE = 2 * KV_STAGES * 8;         // byte distance k_empty[i] -> v_empty[i]
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
  // the operands are pinned in registers for the poll loop: left to itself the allocator re-read them from local memory
  // on EVERY poll, sixteen spinning warps hammering the L1 / shared-memory port the MMAs read K and V through
  asm volatile("" : "+r"(bar), "+r"(parity));
  uint32_t spins = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
    if (++spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// one half-step: S^hf(tile) -> P.  FIRST / LAST are compile-time so that the steady-state instance carries neither the
// reference exchange nor the tail mask (nor the registers their operands would occupy through the loop).
template <int DP, int PF, bool FIRST, bool LAST, bool TRACE>
__device__ __forceinline__ void half_step(uint32_t sh_addr, uint32_t bar_s, float c, float& moff, uint32_t& par_s,
                                          float& l_part, uint32_t cfg, int kt, long long* trace) {
  auto mark = [&](int ev) {
    if constexpr (TRACE) {
      if (trace != nullptr && kt < TRACE_STEPS) trace[kt * TRACE_EV + ev] = clk64();
    }
  };
  mbar_wait_a(bar_s, par_s);
  par_s ^= 1;
  tc_fence_after();
  mark(0);
  // two 32-key chunks; P chunk q goes to columns 16q..16q+15 of the half, over scores already in registers.
  // (Tried: four 16-key chunks double-buffered through registers so that the tcgen05.ld of the next chunk flies
  // under the exponentials — slower: more TMEM instructions, longer MMA issue.)
  uint32_t sr[CH], pk[CH / 2];
#pragma unroll
  for (int q = 0; q < HB / CH; ++q) {
    tmem_ld_x32(sh_addr + q * CH, sr);
    if (q == 0 && ((cfg & (1u << 16)) || (LAST && (cfg & (1u << 17))))) {
      // the LAST half of S(tile kt) is complete, and the tensor pipe completes in issue order: every Q·K^T on this
      // K tile is done, and so is every P·V on the previous V tile whenever a Q·K^T of this tile was issued behind
      // its last half (not so for a one-half last tile: the MMA issuer commits that V tile) — handed back to the
      // TMA producer under the latency of the load above.  (Everything this branch needs is recomputed in it: only
      // one lane of four warps takes it, and nothing of it may occupy a register through the exponential loop.)
      // k_empty[0] from s_full[t][hf]: the barrier array is k_full[3] k_empty[3] v_full[3] v_empty[3] s_full[4] ...
      const uint32_t bar_ke = bar_s - (4 * KV_STAGES + ((cfg >> 29) & 3u)) * 8 + KV_STAGES * 8;
      const uint32_t slot0 = (cfg >> 20) & 3u;
      const uint32_t slot = (slot0 + (uint32_t)kt) % KV_STAGES;                   // ring slot of tile kt
      const uint32_t slot_prev = FIRST ? (cfg >> 22) & 3u : (slot + KV_STAGES - 1) % KV_STAGES;   // ... of the entry before
      mbar_arrive_a(bar_ke + slot * 8);
      if (!FIRST && (cfg & (1u << 16))) mbar_arrive_a(bar_ke + KE_TO_VE + slot_prev * 8);
      if (FIRST && (cfg & (1u << 18))) {                 // Q is in TMEM (this S tile came from it): its ring entry
        mbar_arrive_a(bar_ke + slot_prev * 8);
        mbar_arrive_a(bar_ke + KE_TO_VE + slot_prev * 8);
      }
    }
    tmem_ld_wait();
    if constexpr (LAST) mask_tail(sr, (int)((cfg >> 8) & 0x7fu) - q * CH);       // keys of this half that exist
    if constexpr (FIRST) {
      if (q == 0 && (cfg & (1u << 19))) {
        // the set's reference: max over the first 32 scores of both warps of the row
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int i = 0; i < CH; i += 4) {
          mx0 = fmax3(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
          mx1 = fmax3(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
        }
        const float mine = fmaxf(mx0, mx1);
        float other;
        // own slot of the reference exchange [tile][half][row] (behind the barriers and the redo mask)
        const uint32_t th = (cfg >> 29) & 3u;
        const uint32_t xm_mine = bar_s - (4 * KV_STAGES + th) * 8 + 512 + MASK_WORDS * 4 +
                                 (th * BMQ + (((threadIdx.x >> 5) & 3u) * 32 + (threadIdx.x & 31u))) * 4;
        asm volatile("st.volatile.shared.f32 [%0], %1;" ::"r"(xm_mine), "f"(mine) : "memory");
        named_bar_sync((int)((cfg >> 24) & 15u), 64);
        // the partner warp's slot: key half hf (cfg bit 28) sits BMQ floats before / after
        const uint32_t xm_other = (cfg & (1u << 28)) ? xm_mine - BMQ * 4 : xm_mine + BMQ * 4;
        asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(other) : "r"(xm_other) : "memory");
        moff = fmaxf(mine, other) * c;
      }
    }
    exp_chunk<PF>(sr, c, moff, pk, l_part);
    tmem_st_x16(sh_addr + q * (CH / 2), pk);
    if (q == 0) mark(1);
  }
  mark(2);
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive_a(bar_s + 4 * 8);     // p_full[t][hf] sits four barriers behind s_full[t][hf]
  mark(3);
}

struct Item {
  int h, seq, q_row0, n_qt;
};

// shared-memory layout behind the K/V ring (all roles derive it from the aligned base)
template <int DP>
struct Lay {
  using C = PCfg<DP>;
  uint8_t *sK, *sV;
  uint64_t *k_full, *k_empty, *v_full, *v_empty, *s_full, *p_full, *o_full, *o_empty, *qt_full, *set_done;
  uint32_t *tmem_slot, *retry_flag, *redo_any, *mma_lock, *redo_mask;
  float *xch_m, *xch_l;
  uint8_t* state;
  __device__ __forceinline__ explicit Lay(uint8_t* smem) {
    sK = smem;                                          // ring slot s: K (or Q tile 0) at sK + s·KV_BYTES
    sV = sK + KV_STAGES * C::KV_BYTES;                  //              V (or Q tile 1) at sV + s·KV_BYTES
    uint8_t* s_end = sV + KV_STAGES * C::KV_BYTES;
    k_full = reinterpret_cast<uint64_t*>(s_end);        // KV_STAGES (a Q pair completes on the full barriers of its slot)
    k_empty = k_full + KV_STAGES;
    v_full = k_empty + KV_STAGES;
    v_empty = v_full + KV_STAGES;
    s_full = v_empty + KV_STAGES;                       // [query tile][key half]
    p_full = s_full + 4;                                // [query tile][key half]
    o_full = p_full + 4;                                // [query tile]: the last P·V of the pass has completed
    o_empty = o_full + 2;                               // [query tile]: every softmax warp has read (or dropped) O
    qt_full = o_empty + 2;                              // [query tile]: Q_t copied from shared memory into TMEM
    set_done = qt_full + 2;                             // careful phase: every softmax warp has posted its verdict
    tmem_slot = reinterpret_cast<uint32_t*>(set_done + 1);
    retry_flag = tmem_slot + 1;                         // 1 + index of the last careful pass that overflowed
    redo_any = tmem_slot + 2;
    mma_lock = tmem_slot + 3;                           // the two MMA issuers take turns (see the issuer role)
    redo_mask = reinterpret_cast<uint32_t*>(s_end + 512);            // bit k: this CTA's k-th item overflowed
    xch_m = reinterpret_cast<float*>(s_end + 512 + MASK_WORDS * 4);  // [tile][half][row]: reference exchange
    xch_l = xch_m + QT * 2 * BMQ;                       // [pass parity][tile][half][row]: row-sum exchange
    state = reinterpret_cast<uint8_t*>(xch_m + C::XCH_FLOATS);      // [softmax warp]: schedule record (SmState)
  }
};

__device__ __forceinline__ Item item_at(const AParams& p, int n_qp, int k) {
  const int it = (int)blockIdx.x + k * (int)gridDim.x;
  const int qp = it % n_qp, r = it / n_qp;
  Item x;
  x.h = r % p.H;
  x.seq = r / p.H;
  x.q_row0 = qp * QT * BMQ;
  x.n_qt = (x.q_row0 + BMQ < p.Lq || (p.dbg & 4)) ? 2 : 1;   // the second tile may lie entirely past Lq
  return x;
}
// next item of a phase after k (fast: k + 1; careful: the next flagged one), n_my when there is none
__device__ __forceinline__ int next_item(const uint32_t* redo_mask, bool fast, int k, int n_my) {
  if (fast) return k + 1;
  for (++k; k < n_my; ++k)
    if (redo_mask[k >> 5] >> (k & 31) & 1) break;
  return k;
}

// The schedule of a softmax warp lives in a record in SHARED memory (one per warp, every lane reads the same words),
// accessed through volatile pointers — not in registers of the exponential loop: 96 registers per thread is the cap at
// 19 warps and that loop wants ~75 of them.  (With the schedule in ordinary variables the allocator parked a dozen
// values in local memory and re-read them every half-step, two of them on every POLL of the S-tile barrier; with the
// schedule in a local-memory record behind a never-inlined call the steady state was fine but every pass boundary cost
// ~9000 clk: with 204 KB of shared memory the L1 has ~28 KB left, the 150 KB of stack frames thrash it, and each of the
// ~100 local loads / stores of the cold path is an L2 round trip — `tools/attn_pers_boundary.py`.)
struct SmState {
  int started, phase, k, set, attempt, cpass;
  int e6;                      // ring entries so far, mod 6: slot e6 % 3, full-barrier parity e6 / 3
  int slot_q;                  // ring slot of the current item's Q pair
  uint32_t par_h1;             // parity of key half 1's NEXT S tile (only warp (t, 0) looks: all Q·K^T of an item done?)
  uint32_t par_o;              // parity of the next pass (o_full[t], row-sum exchange buffer)
  int pre;                     // Q_t of the next item is in TMEM already (copied at the end of the previous item)
  int idle;                    // the current item's tile t lies past Lq
  int npass;                   // (TRACE) passes so far
};

// End of the pass a softmax warp has just run (row sums, verdict, O, epilogue) and the head of its next one (next set /
// item / phase, Q_t into TMEM).  Never inlined.  Returns the control word of the next pass.
template <int DP, bool TRACE>
__device__ __forceinline__ uint32_t sm_next(volatile SmState* st, const AParams* pp, uint8_t* smem, int n_qp, int n_items,
                                            float l_part, float* moff_io) {
  using C = PCfg<DP>;
  const AParams& p = *pp;
  const Lay<DP> L(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quarter = warp & 3;                        // TMEM lane quarter this warp may touch (= its sub-partition)
  const int t = warp >> 3, hf = (warp >> 2) & 1;
  const int row = quarter * 32 + lane;
  const uint32_t tmem_base = *L.tmem_slot;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
  const int pair_bar = 1 + t * 4 + quarter;            // named barrier of the two warps that share these rows
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int nh_last = p.Lk - (n_kt - 1) * BNK > HB ? 2 : 1;
  const int n_half1 = n_kt - (nh_last == 1 ? 1 : 0);   // half-steps of key half 1 per pass
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int xi = (t * 2 + hf) * BMQ + row, xo = (t * 2 + (hf ^ 1)) * BMQ + row;   // own / partner exchange slot
  const bool la_on = !(p.dbg & 1);
  auto add6 = [](int a, int n) { return (a + n) % (2 * KV_STAGES); };
  auto slot_of = [](int a) { return a >= KV_STAGES ? a - KV_STAGES : a; };
  // Q_t: this thread's row from the 128-byte-swizzled TMA tile (16-byte unit u of row r sits at u ^ (r % 8))
  // -> TMEM columns q_col(t) .. (+ DP/2), the A-operand layout (lane = row, bf16 pairs along K)
  auto copy_q = [&](int e6q, bool idle) {                 // e6q: ring position (mod 6) of the Q pair
    const int slot = slot_of(e6q);
    mbar_wait(t ? &L.v_full[slot] : &L.k_full[slot], (uint32_t)(e6q >= KV_STAGES));
    if (!idle) {
      uint32_t qw[DP / 2];
      const uint8_t* qrow = (t ? L.sV : L.sK) + slot * C::KV_BYTES + (row >> 3) * 1024 + (row & 7) * 128;
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
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&L.qt_full[t]);
  };
  auto verdict = [&](int cp, int attempt) -> bool {
    mbar_wait(L.set_done, (uint32_t)(cp & 1));
    const uint32_t f = *reinterpret_cast<volatile uint32_t*>(L.retry_flag);
    return f == (uint32_t)(cp + 1) && attempt < MAX_RETRY;
  };

  // TRACE builds: where the time between two passes goes (rows 80 + pass of this warp's timeline, lane 0 of CTA 5)
  long long* btr = nullptr;
  if constexpr (TRACE) {
    if (p.trace != nullptr && blockIdx.x == 5 && lane == 0 && st->npass < 12)
      btr = p.trace + ((int64_t)warp * TRACE_STEPS + 80 + st->npass) * TRACE_EV;
    st->npass = st->npass + 1;
  }
  auto bmark = [&](int ev) {
    if constexpr (TRACE) {
      if (btr != nullptr) btr[ev] = clk64();
    }
  };
  bmark(0);
  bool fast = st->phase == 0;
  int k = st->k;
  bool new_item = false;
  if (st->started) {
    // ---------------- end of the pass (k, set, attempt) ----------------
    const int set = st->set, attempt = st->attempt;
    const bool idle = st->idle != 0;
    st->par_h1 = st->par_h1 ^ (uint32_t)(n_half1 & 1);
    st->e6 = add6(st->e6, n_kt);
    const int kn = next_item(L.redo_mask, fast, k, n_my);
    const bool nxt_new = fast && la_on && set + 1 == p.n_sets && kn < n_my;
    if (nxt_new && hf == 0) {
      // Q_t of the NEXT item -> TMEM: every Q·K^T of this item has completed (key half 1's last S tile, too)
      if (n_half1 > 0) mbar_wait(&L.s_full[t * 2 + 1], st->par_h1 ^ 1);
      tc_fence_after();
      copy_q(st->e6, t >= item_at(p, n_qp, kn).n_qt);
    }
    bmark(1);
    // where this thread's output row lives
    const Item it = item_at(p, n_qp, k);
    const int q_idx = it.q_row0 + t * BMQ + row;
    const bool live_row = !idle && q_idx < p.Lq;
    bf16* orow = p.out + ((int64_t)it.seq * p.Lq + q_idx) * p.ldo + (int64_t)it.h * p.D;
    // earlier sets' partial sums of the row: requested now, needed after O (the loads fly under the waits)
    constexpr int NU = DP / 16;                           // 8-column units per warp
    uint4 prev[NU];
    if (set > 0 && live_row) {
#pragma unroll
      for (int u = 0; u < NU; ++u) {
        const int d = hf * (DP / 2) + u * 8;
        if (d < p.D) prev[u] = *reinterpret_cast<const uint4*>(orow + d);
      }
    }
    // the row sum of both halves (an idle tile's two warps skip the exchange together)
    float l_row = 1.f;
    if (!idle) {
      float* xl = L.xch_l + st->par_o * (QT * 2 * BMQ);
      xl[xi] = l_part;
      named_bar_sync(pair_bar, 64);
      l_row = l_part + xl[xo];
    }
    bmark(2);
    const bool ok = l_row < L_LIMIT;                       // false for NaN, too
    bool again = false;
    if (fast) {
      if (__any_sync(0xffffffffu, !ok) && lane == 0) {
        atomicOr(&L.redo_mask[k >> 5], 1u << (k & 31));
        *reinterpret_cast<volatile uint32_t*>(L.redo_any) = 1;
      }
    } else {
      if (!ok) atomicMax(L.retry_flag, (uint32_t)(st->cpass + 1));
      __syncwarp();
      if (lane == 0) mbar_arrive(L.set_done);
      again = verdict(st->cpass, attempt);
      st->cpass = st->cpass + 1;
    }
    mbar_wait(&L.o_full[t], st->par_o);                    // every MMA of this tile's pass has completed
    st->par_o = st->par_o ^ 1u;
    tc_fence_after();
    bmark(3);
    uint32_t r[NU * 8];
    if (!again && !idle) {
#pragma unroll
      for (int u = 0; u < NU; ++u) tmem_ld_x8(lane_addr + o_col(t) + hf * (DP / 2) + u * 8, r + u * 8);
      tmem_ld_wait();
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&L.o_empty[t]);             // the next pass may overwrite O_t
    bmark(4);
    st->pre = nxt_new;
    if (again) {
      if (!ok) *moff_io += RETRY_SHIFT;
      st->attempt = attempt + 1;
    } else {
      // normalise, round to bf16, add to what earlier sets left in `out`: this warp takes half of the columns
      if (live_row) {
        bf16* dst = orow;                                  // the final (last set) row may go to a peer
        if (set == p.n_sets - 1 && p.n_peer > 0) {
          const int owner = q_idx / p.rows_per_rank;
          dst = p.peer_out[owner] + ((int64_t)it.seq * p.rows_per_rank + (q_idx - owner * p.rows_per_rank)) * p.peer_ld +
                p.peer_col + (int64_t)it.h * p.D;
        }
        const float inv = 1.f / l_row;
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
              Vec16<bf16> pv;
              pv.raw = prev[u];
              float gg[8], a[8];
              pv.unpack(gg);
              o.unpack(a);
#pragma unroll
              for (int i = 0; i < 8; ++i) a[i] += gg[i];
              o.pack(a);
            }
            *reinterpret_cast<uint4*>(dst + d) = o.raw;
          }
        }
      }
      st->attempt = 0;
      if (set + 1 < p.n_sets) {
        st->set = set + 1;
      } else {
        st->set = 0;
        k = kn;
        new_item = true;
      }
    }
  } else {
    st->started = 1;
    k = next_item(L.redo_mask, true, -1, n_my);
    new_item = true;
  }
  bmark(5);
  // ---------------- head of the next pass ----------------
  if (new_item) {
    while (k >= n_my) {
      if (!fast) return 1u;                                // DONE
      // between the phases: the whole CTA (every role calls it at a warp-converged point)
      named_bar_sync(0, NTHREADS);
      if (*reinterpret_cast<volatile uint32_t*>(L.redo_any) == 0) return 1u;
      fast = false;
      st->phase = 1;
      st->pre = 0;
      k = next_item(L.redo_mask, false, -1, n_my);
    }
    st->k = k;
    const bool idle = fast && t >= item_at(p, n_qp, k).n_qt;   // this tile lies past Lq: walk the protocol, skip the math
    st->idle = idle;
    if (!st->pre && hf == 0) copy_q(st->e6, idle);
    st->slot_q = slot_of(st->e6);
    st->e6 = add6(st->e6, 1);
  }
  bmark(6);
  const bool releaser = quarter == 0 && lane == 0;         // lane that hands ring slots back to the producer
  const bool first = st->set == 0 && st->attempt == 0;
  // keys of this warp's half that exist in the pass's last tile (HB when it never gets there): ctl bits 8-14
  const int n_mine = hf == 0 ? n_kt : n_half1;
  const int tail_left = n_mine == n_kt ? p.Lk - (n_kt - 1) * BNK - hf * HB : HB;
  return (st->idle ? 2u : 0u) | (uint32_t)(tail_left < HB ? (tail_left < 0 ? 0 : tail_left) : HB) << 8 |
         (uint32_t)(releaser && hf == 1) << 16 | (uint32_t)(releaser && hf == 0 && nh_last == 1) << 17 |
         (uint32_t)first << 18 | (uint32_t)(st->attempt == 0) << 19 | (uint32_t)slot_of(st->e6) << 20 |
         (uint32_t)st->slot_q << 22 | (uint32_t)pair_bar << 24 | (uint32_t)hf << 28 | (uint32_t)(t * 2 + hf) << 29;
}

template <int DP, int PF, bool TRACE>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_pers_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const __grid_constant__ AParams p, const int n_qp,
                 const int n_items) {
  using C = PCfg<DP>;
  extern __shared__ uint8_t smem_raw[];
#ifdef MDV2_SMEM_GENERIC                               // A/B hook: the previous form (loses the address space)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
#else
  // pointer arithmetic on the __shared__ array keeps the address space (LDS / STS instead of generic LD.E / ST.E)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
#endif
  const Lay<DP> L(smem);
  uint8_t *const sK = L.sK, *const sV = L.sV;
  uint64_t *const bars = L.k_full, *const k_full = L.k_full, *const k_empty = L.k_empty, *const v_full = L.v_full,
                 *const v_empty = L.v_empty, *const s_full = L.s_full, *const p_full = L.p_full, *const o_full = L.o_full,
                 *const o_empty = L.o_empty, *const qt_full = L.qt_full, *const set_done = L.set_done;
  uint32_t *const tmem_slot = L.tmem_slot, *const retry_flag = L.retry_flag, *const redo_any = L.redo_any,
                 *const redo_mask = L.redo_mask;
  float* const xch_m = L.xch_m;

  // the warp index through a shuffle: the compiler then knows it is warp-uniform, so the role branches are uniform
  // branches and the schedule counters below live in uniform registers (96 registers per thread is the cap at 19 warps)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int n_kt = (p.Lk + BNK - 1) / BNK;
  const int nh_last = p.Lk - (n_kt - 1) * BNK > HB ? 2 : 1;      // half-steps of the last K/V tile
  const int nh_first = n_kt == 1 ? nh_last : 2;                  // ... of the first
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // items of this CTA
  const bool tracing = TRACE && p.trace != nullptr && blockIdx.x == 5;
  auto mark = [&](int j, int ev) {
    if constexpr (TRACE) {
      if (tracing && lane == 0 && j < TRACE_STEPS) p.trace[(warp * TRACE_STEPS + j) * TRACE_EV + ev] = clk64();
    }
  };
  auto item_of = [&](int k) { return item_at(p, n_qp, k); };

  if (warp == TMA_WARP && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 2);                        // one arrival per query tile (a lone tile arrives twice)
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 2);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&o_full[i], 1);
      mbar_init(&o_empty[i], 8);
      mbar_init(&qt_full[i], 4);
    }
    mbar_init(set_done, 16);
    *retry_flag = 0;
    *redo_any = 0;
    *L.mma_lock = 0;
    fence_mbar_init();
  }
  for (int i = threadIdx.x; i < MASK_WORDS; i += NTHREADS) redo_mask[i] = 0;
  if (warp == MMA_WARP0) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // TRACE builds: a wait that does not complete within ~2^18 polls is RECORDED (block, warp, source line, barrier index,
  // parity) behind the timeline and the thread carries on, so a protocol bug can be read back from a kernel that ends
  // (tools/attn_pers_debug.py) instead of a trapped context
  auto wait_dbg = [&](uint64_t* bar, uint32_t parity, int line) {
    if constexpr (TRACE) {
      uint32_t spins = 0;
      for (;;) {
        uint32_t ok;                                     // parked between polls like the product's waits (1 us per probe)
        asm volatile(
            "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
            : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(1000u) : "memory");
        if (ok) return;
        if (++spins > (1u << 14)) {
          if (p.trace != nullptr && lane == 0) {
            const unsigned long long slot = atomicAdd(reinterpret_cast<unsigned long long*>(p.trace + TRACE_CTA_OFF), 1ull);
            if (slot < 256) {
              long long* rec = p.trace + TRACE_CTA_OFF + 8 + slot * 8;
              rec[0] = blockIdx.x; rec[1] = warp; rec[2] = line; rec[3] = (long long)(bar - bars); rec[4] = parity;
            }
          }
          return;
        }
      }
    } else {
      mbar_wait(bar, parity);
    }
  };
  auto wait_dbg_parked = [&](uint64_t* bar, uint32_t parity, uint32_t hint, int line) {
    if constexpr (TRACE) wait_dbg(bar, parity, line);
    else mbar_wait_parked(bar, parity, hint);
  };
#define PW(bar, par) wait_dbg(bar, par, __LINE__)
#define PWP(bar, par, hint) wait_dbg_parked(bar, par, hint, __LINE__)

  // careful phase: the verdict on careful pass `cp` — does it have to be repeated?
  auto verdict = [&](int cp, int attempt) -> bool {
    PW(set_done, (uint32_t)(cp & 1));
    const uint32_t f = *reinterpret_cast<volatile uint32_t*>(retry_flag);
    return f == (uint32_t)(cp + 1) && attempt < MAX_RETRY;
  };
  auto next_k = [&](bool fast, int k) { return next_item(redo_mask, fast, k, n_my); };
  // between the phases: the whole CTA (every role calls it at a warp-converged point; 0 is the __syncthreads barrier)
  auto phase_sync = [&]() -> bool {
    named_bar_sync(0, NTHREADS);
    return *reinterpret_cast<volatile uint32_t*>(redo_any) != 0;
  };
  const bool la_on = !(p.dbg & 1);

  // Every role walks the same schedule with its OWN counters (shared ones would be live in every role's code and spill
  // wherever the softmax loop is short of registers): e = ring entries so far (a Q pair or a K/V tile; slot e % 3,
  // barrier parity (e / 3) & 1), cpass = careful passes so far (set_done parity).
  if (warp == TMA_WARP) {
    // ================= TMA producer =================
    int e = 0, cpass = 0;
    for (int phase = 0; phase < 2; ++phase) {
      if (phase == 1 && !phase_sync()) break;
      const bool fast = phase == 0;
      if (lane == 0) {
        for (int k = next_k(fast, -1); k < n_my; k = next_k(fast, k)) {
          const Item it = item_of(k);
          const int n_qt = fast ? it.n_qt : 2;
          {                                              // the Q pair: one ring entry
            const int st = e % KV_STAGES;
            const uint32_t ph = (uint32_t)((e / KV_STAGES) & 1);
            PWP(&k_empty[st], ph ^ 1, 4000);
            PWP(&v_empty[st], ph ^ 1, 4000);
            // (the slot's own full barriers, so that their parity stays (entry / 3) & 1 for every kind of entry)
            for (int t = 0; t < 2; ++t) {
              uint64_t* qf = t ? &v_full[st] : &k_full[st];
              if (t >= n_qt) {
                mbar_arrive(qf);                           // a lone tile 0: the V slot stays empty
                break;
              }
              mbar_expect_tx(qf, C::Q_BYTES);
              const int row = it.seq * p.Lq + it.q_row0 + t * BMQ;
#pragma unroll
              for (int c = 0; c < C::NCH; ++c)
                tma_load_3d((t ? sV : sK) + st * C::KV_BYTES + c * BMQ * 128, &tq, qf, c * 64, it.h, row);
            }
            ++e;
          }
          for (int set = 0; set < p.n_sets; ++set) {
            const int kseq = p.kv_map ? p.kv_map[it.seq * p.n_sets + set] : it.seq;
            for (int attempt = 0;; ++attempt) {
              for (int kt = 0; kt < n_kt; ++kt, ++e) {
                const int st = e % KV_STAGES;
                const uint32_t ph = (uint32_t)((e / KV_STAGES) & 1);
                const int row = kseq * p.Lk + kt * BNK;
                PWP(&k_empty[st], ph ^ 1, 4000);
                mark(e, 0);
                mbar_expect_tx(&k_full[st], C::KV_BYTES);
#pragma unroll
                for (int c = 0; c < C::NCH; ++c)
                  tma_load_3d(sK + st * C::KV_BYTES + c * BNK * 128, &tk, &k_full[st], c * 64, it.h, row);
                PWP(&v_empty[st], ph ^ 1, 4000);
                mark(e, 1);
                mbar_expect_tx(&v_full[st], C::KV_BYTES);
#pragma unroll
                for (int c = 0; c < C::NCH; ++c)
                  tma_load_3d(sV + st * C::KV_BYTES + c * BNK * 128, &tv, &v_full[st], c * 64, it.h, row);
              }
              if (fast) break;
              const bool again = verdict(cpass, attempt);
              ++cpass;
              if (!again) break;
            }
          }
        }
      }
      __syncwarp();
    }
    if (lane == 0 && n_my > 0) {
      // Drain: the V tiles of the last pass come back through tcgen05.commit — an ASYNCHRONOUS arrive on this CTA's
      // shared memory that nobody waits for.  If the CTA retired first, the arrive would land in the barriers of
      // whichever CTA (of whichever kernel) owns that shared memory next.  Every ring entry is handed back exactly
      // once, so the producer simply waits for the last KV_STAGES entries as if it were to reuse their slots.
      for (int i = 0; i < KV_STAGES; ++i, ++e) {
        const int st = e % KV_STAGES;
        const uint32_t ph = (uint32_t)((e / KV_STAGES) & 1);
        PWP(&k_empty[st], ph ^ 1, 4000);
        PWP(&v_empty[st], ph ^ 1, 4000);
      }
    }
    __syncwarp();
  } else if (warp > TMA_WARP) {
    // ================= MMA issuers: warp 17 -> query tile 0, warp 18 -> query tile 1 =================
    // The whole warp runs the warp-uniform control flow and one elected lane issues the tcgen05 instructions.
    const int t = warp - MMA_WARP0;
    constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, HB, 0, 0);
    constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);     // A = P from TMEM, B = V is MN-major
    const uint64_t dk0 = make_desc_k_sw128(smem_u32(sK));
    const uint64_t dv0 = make_desc_mn_sw128(smem_u32(sV), BNK * 128, 1024);
    const uint32_t s_tm = tmem_base + s_col(t), o_tm = tmem_base + o_col(t), q_tm = tmem_base + q_col(t);
    // S_t^hf = Q_t · K[hf·64 .. +63]^T -> TMEM columns s_col(t) + hf·64 .. +63, A = Q_t from TMEM (one lane calls this)
    // (idle: a query tile that lies entirely past Lq — it walks the whole barrier protocol, so that every role stays
    // within one phase of the ring, but issues no MMA: a commit with nothing before it arrives at once)
    auto issue_qk = [&](int hf, int st, bool idle) {
      if (!idle) {
        const uint64_t db = dk0 + (uint64_t)(st * (C::KV_BYTES >> 4) + hf * ((HB * 128) >> 4));
#pragma unroll
        for (int ks = 0; ks < DP / 16; ++ks) {
          const uint64_t ob = (uint64_t)((ks / 4) * ((BNK * 128) >> 4) + (ks % 4) * 2);
          umma_bf16_ts(s_tm + hf * HB, q_tm + ks * 8, db + ob, idesc_qk, ks != 0);
        }
      }
      umma_commit(&s_full[t * 2 + hf]);
    };
    // The two issuers take TURNS at the tensor pipe: a block (P·V of a half + Q·K^T of the next tile's half, ~9 MMAs,
    // ~360 clk) is issued under a shared-memory lock.  Left alone, the blocks of the two query tiles — which start every
    // pass together — interleave instruction by instruction, BOTH take twice as long, both score tiles arrive late
    // together, and the symmetric schedule never breaks (measured: 2700 instead of 2180 clk per K/V tile); one after the
    // other, the first tile's scores are ready 360 clk earlier and the tiles settle half a block apart.
    uint32_t* const lock = L.mma_lock;
    const bool use_lock = (p.dbg & 8) != 0;   // experiment switch (MDV2_ATTN_DBG=8), off: see the measurements in DESIGN.md
    auto lock_acquire = [&]() {
      if (use_lock)
        while (atomicCAS(lock, 0u, 1u) != 0u) {}
    };
    auto lock_release = [&]() {
      if (use_lock) atomicExch(lock, 0u);
    };
    int e = 0, cpass = 0;
    int nqt = 0;                 // items so far (qt_full[t] parity)
    int np = 0;                  // passes so far (o_full / o_empty parity)
    uint32_t hs0 = 0, hs1 = 0;   // half-steps of key half 0 / 1 so far (p_full parity)
    bool pre = false;            // this pass was started by the previous one (Q in TMEM, first S tile issued)
    for (int phase = 0; phase < 2; ++phase) {
      if (phase == 1) {
        if (!phase_sync()) break;
        pre = false;
      }
      const bool fast = phase == 0;
      for (int k = next_k(fast, -1); k < n_my; k = next_k(fast, k)) {
        const bool idle = fast && t >= item_of(k).n_qt;
        const int kn = next_k(fast, k);
        const bool has_next = fast && la_on && kn < n_my;
        const bool idle_next = has_next && t >= item_of(kn).n_qt;
        ++e;                                             // the Q pair
        for (int set = 0; set < p.n_sets; ++set) {
          for (int attempt = 0;; ++attempt) {
            const int e0 = e;
            const bool nxt_same = fast && la_on && set + 1 < p.n_sets;
            const bool nxt_new = set + 1 == p.n_sets && has_next;
            const bool la = nxt_same || nxt_new;         // start the next pass behind the last tile of this one
            const bool idle_n = nxt_same ? idle : idle_next;
            const int eN = nxt_same ? e0 + n_kt : e0 + n_kt + 1;       // ring entry of the next pass's first K/V tile
            if (!pre) {
              if (set == 0 && attempt == 0) PWP(&qt_full[t], (uint32_t)(nqt & 1), 1000);
              PWP(&k_full[e0 % KV_STAGES], (uint32_t)((e0 / KV_STAGES) & 1), 1000);
              tc_fence_after();
              if (elect_one()) {
                lock_acquire();
                for (int hf = 0; hf < nh_first; ++hf) issue_qk(hf, e0 % KV_STAGES, idle);
                lock_release();
              }
              __syncwarp();
            }
            if (np > 0) {                                // O_t of the previous pass has been read
              PWP(&o_empty[t], (uint32_t)((np - 1) & 1), 1000);
              tc_fence_after();
            }
            for (int kt = 0; kt < n_kt; ++kt) {
              const int j = e0 + kt;
              const int st = j % KV_STAGES, st1 = (j + 1) % KV_STAGES;
              const bool more = kt + 1 < n_kt;
              const int nh = kt == n_kt - 1 ? nh_last : 2;
              const int nh1 = kt + 1 == n_kt - 1 ? nh_last : 2;               // half-steps of the next tile
              mark(j, 5);
              PWP(&v_full[st], (uint32_t)((j / KV_STAGES) & 1), 1000);
              mark(j, 6);
              if (more) PWP(&k_full[st1], (uint32_t)(((j + 1) / KV_STAGES) & 1), 1000);
              const uint64_t dv = dv0 + (uint64_t)(st * (C::KV_BYTES >> 4));
              mark(j, 0);
              for (int hf = 0; hf < nh; ++hf) {
                // operands of the block, ready BEFORE the wait: behind it every instruction is on the critical chain
                // P ready -> S of the next tile ready
                const uint64_t dvh = dv + (uint64_t)((hf * HB) * 128 >> 4);
                const uint64_t dbh = dk0 + (uint64_t)(st1 * (C::KV_BYTES >> 4) + hf * ((HB * 128) >> 4));
                const uint32_t p_tm = s_tm + hf * HB;
                const bool steady = more && !idle;
                const bool qk_next = hf < nh1;
                const bool v_back = hf == 1 && nh1 == 1 && kt + 2 == n_kt;
                PWP(&p_full[t * 2 + hf], (hf ? hs1 : hs0) & 1, 1000);
                if (hf) ++hs1; else ++hs0;
                tc_fence_after();
                mark(j, 1 + 2 * hf);
                if (steady) {
                  // the common case, straight-line: P·V of this half, then Q·K^T of the same half of the next tile
                  if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < HB / 16; ++ks)
                      umma_bf16_ts(o_tm, p_tm + ks * 8, dvh + (uint64_t)((ks * 16) * 128 >> 4), idesc_pv, (ks | hf | kt) != 0);
                    if (qk_next) {
#pragma unroll
                      for (int ks = 0; ks < DP / 16; ++ks)
                        umma_bf16_ts(p_tm, q_tm + ks * 8,
                                     dbh + (uint64_t)((ks / 4) * ((BNK * 128) >> 4) + (ks % 4) * 2), idesc_qk, ks != 0);
                      umma_commit(&s_full[t * 2 + hf]);
                    }
                    if (v_back) umma_commit(&v_empty[st]);
                  }
                  __syncwarp();
                  mark(j, 2 + 2 * hf);
                  continue;
                }
                if (elect_one()) {
                  lock_acquire();
                  // O_t (+)= P_t^hf · V[hf·64 ..]: P^hf sits on the first 32 columns of S_t^hf
                  if (!idle) {
#pragma unroll
                    for (int ks = 0; ks < HB / 16; ++ks)
                      umma_bf16_ts(o_tm, s_tm + hf * HB + ks * 8, dv + (uint64_t)((hf * HB + ks * 16) * 128 >> 4), idesc_pv,
                                   (ks != 0) || (hf != 0) || (kt != 0));
                  }
                  if (hf == nh - 1) {
                    // V tiles no later Q·K^T of the pass follows are handed back by a commit behind their last P·V: the
                    // last tile, and the one before a one-half last tile
                    if (!more) umma_commit(&o_full[t]);
                    if (!more || (nh1 == 1 && kt + 2 == n_kt)) umma_commit(&v_empty[st]);
                  }
                  if (more && hf < nh1) issue_qk(hf, st1, idle);   // S_t^hf of the next tile over the P it just consumed
                  lock_release();
                }
                __syncwarp();
                if (!more && la) {
                  // the next pass's first score tile, over the P columns the MMA above has just consumed
                  if (hf == 0) {
                    if (nxt_new) PWP(&qt_full[t], (uint32_t)((nqt + 1) & 1), 1000);
                    PWP(&k_full[eN % KV_STAGES], (uint32_t)((eN / KV_STAGES) & 1), 1000);
                    tc_fence_after();
                  }
                  if (hf < nh_first && elect_one()) {
                    lock_acquire();
                    issue_qk(hf, eN % KV_STAGES, idle_n);
                    lock_release();
                  }
                  __syncwarp();
                }
                mark(j, 2 + 2 * hf);
              }
              if (!more && la && nh == 1 && nh_first == 2) {
                if (elect_one()) {
                  lock_acquire();
                  issue_qk(1, eN % KV_STAGES, idle_n);
                  lock_release();
                }
                __syncwarp();
              }
            }
            ++np;
            e = e0 + n_kt;
            pre = la;
            if (fast) break;
            const bool again = verdict(cpass, attempt);
            ++cpass;
            if (!again) break;
          }
        }
        ++nqt;
      }
    }
  } else if (warp < N_SM_WARPS) {
    // ================= softmax warps: warp (t, hf), thread = query row =================
    // The loop below only runs half-steps; everything between two passes (row sums, verdict, O, the epilogue, the next
    // set / item / phase, Q_t into TMEM) is sm_next() above, with the schedule in a local-memory record.
    const int quarter = warp & 3;                      // TMEM lane quarter this warp may touch (= its sub-partition)
    const int t = warp >> 3, hf = (warp >> 2) & 1;
    const uint32_t sh_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + s_col(t) + hf * HB;
    const uint32_t bar_s = smem_u32(&s_full[t * 2 + hf]), bar_ke = smem_u32(k_empty);
    const float c = p.scale_log2;
    const int n_mine = hf == 0 ? n_kt : n_kt - (nh_last == 1 ? 1 : 0);     // half-steps of this warp per pass
    long long* tr = nullptr;
    if constexpr (TRACE) {
      if (tracing && lane == 0) tr = p.trace + (int64_t)warp * TRACE_STEPS * TRACE_EV;
    }
    volatile SmState* const stp = reinterpret_cast<volatile SmState*>(L.state + warp * 64);
    static_assert(sizeof(SmState) <= 64, "one 64-byte record per warp");
    stp->started = 0; stp->phase = 0; stp->k = 0; stp->set = 0; stp->attempt = 0; stp->cpass = 0; stp->e6 = 0;
    stp->slot_q = 0; stp->par_h1 = 0; stp->par_o = 0; stp->pre = 0; stp->idle = 0; stp->npass = 0;
    __syncwarp();
    float moff = 0.f;                                    // the row's reference, log2 units
    uint32_t par_s = 0;                                  // parity of this warp's next S tile (s_full[t][hf])
    uint32_t ctl = sm_next<DP, TRACE>(stp, &p, smem, n_qp, n_items, 0.f, &moff);
    while (!(ctl & 1u)) {
      float l_part = 0.f;
      if (!(ctl & 2u)) {
        if (n_mine == 1) {
          half_step<DP, PF, true, true, TRACE>(sh_addr, bar_s, c, moff, par_s, l_part, ctl, 0, tr);
        } else if (n_mine > 1) {
          half_step<DP, PF, true, false, TRACE>(sh_addr, bar_s, c, moff, par_s, l_part, ctl, 0, tr);
#pragma unroll 1
          for (int kt = 1; kt < n_mine - 1; ++kt)
            half_step<DP, PF, false, false, TRACE>(sh_addr, bar_s, c, moff, par_s, l_part, ctl, kt, tr);
          half_step<DP, PF, false, true, TRACE>(sh_addr, bar_s, c, moff, par_s, l_part, ctl, n_mine - 1, tr);
        }
      } else {
        // a tile past Lq: only the hand-shakes of the pass, so that the ring and the MMA issuer see the same protocol
        uint32_t slot = (ctl >> 20) & 3u, slot_prev = (ctl >> 22) & 3u;    // slot of tile kt / of the ring entry before it
        for (int kt = 0; kt < n_mine; ++kt) {
          mbar_wait_a(bar_s, par_s);
          par_s ^= 1;
          if ((ctl & (1u << 16)) || (kt == n_mine - 1 && (ctl & (1u << 17)))) {
            mbar_arrive_a(bar_ke + slot * 8);
            if (kt > 0 && (ctl & (1u << 16))) mbar_arrive_a(bar_ke + KE_TO_VE + slot_prev * 8);
            if (kt == 0 && (ctl & (1u << 18))) {
              mbar_arrive_a(bar_ke + slot_prev * 8);
              mbar_arrive_a(bar_ke + KE_TO_VE + slot_prev * 8);
            }
          }
          slot_prev = slot;
          slot = slot == KV_STAGES - 1 ? 0 : slot + 1;
          __syncwarp();
          if (lane == 0) mbar_arrive_a(bar_s + 4 * 8);
        }
      }
      ctl = sm_next<DP, TRACE>(stp, &p, smem, n_qp, n_items, l_part, &moff);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
#undef PW
#undef PWP
}

template <int DP, int PF, bool TRACE>
int launch_pers(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p, int n_seq,
                cudaStream_t st) {
  constexpr int SMEM = PCfg<DP>::SMEM;
  static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_pers_kernel<DP, PF, TRACE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("attention_pers: cannot set %d B dynamic smem: %s", SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    configured = true;
  }
  const int n_qp = (p.Lq + QT * BMQ - 1) / (QT * BMQ);
  const int64_t n_items = (int64_t)n_qp * p.H * n_seq;
  MDV2_REQUIRE(n_items <= 0x7fffffff, "attention_pers: too many work items");
  static const int forced = [] {                         // tuning hook: CTAs of the persistent grid
    const char* e = getenv("MDV2_ATTN_CTAS");
    return e ? atoi(e) : 0;
  }();
  int64_t grid = forced > 0 ? forced : mdv2_num_sms();
  const int64_t need = (n_items + MASK_WORDS * 32 - 1) / (MASK_WORDS * 32);      // redo mask capacity per CTA
  if (grid < need) grid = need;
  if (grid > n_items) grid = n_items;
  attn_pers_kernel<DP, PF, TRACE><<<(unsigned)grid, NTHREADS, SMEM, st>>>(tq, tk, tv, p, n_qp, (int)n_items);
  MDV2_CHECK_LAUNCH("attention_fast");
  return MDV2_OK;
}
}  // namespace

// called by mdv2_attention_sm100 (attn_sm100.cu) for K/V sequences longer than 64 keys (MDV2_ATTN_PERSIST=0: the one-shot
// kernel of attn_fast_sm100.cu)
int mdv2_attention_pers_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                               int n_seq, int D, bool trace, cudaStream_t st) {
  static const int pf = [] {                             // tuning hook: polynomial pairs per 8 (tools/bench_attn.py)
    const char* e = getenv("MDV2_ATTN_PF");
    return e ? atoi(e) : 3;
  }();
  if (D == 72) {
    if (trace) return launch_pers<80, 3, true>(tq, tk, tv, p, n_seq, st);
    switch (pf) {
      case 0: return launch_pers<80, 0, false>(tq, tk, tv, p, n_seq, st);
      case 2: return launch_pers<80, 2, false>(tq, tk, tv, p, n_seq, st);
      case 3: return launch_pers<80, 3, false>(tq, tk, tv, p, n_seq, st);
      case 4: return launch_pers<80, 4, false>(tq, tk, tv, p, n_seq, st);
      default: MDV2_REQUIRE(false, "attention_pers: MDV2_ATTN_PF must be 0, 2, 3 or 4");
    }
  }
  return launch_pers<64, 0, false>(tq, tk, tv, p, n_seq, st);
}
