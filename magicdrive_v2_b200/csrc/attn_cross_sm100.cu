// attn_cross_sm100.cu — cross-attention onto <= 64 condition tokens (text / box / camera / frame tokens, L ~ 62):
// MultiHeadCrossAttention of every STDiT3 block (reference blocks.py:678-747, magicdrive_stdit3.py:267-280).
//
// The whole K/V of a (sequence, head) is ONE 64-key tile, so a query tile is a single latency chain — load Q, one Q·K^T,
// one softmax, one P·V, store — with ~62 FLOP per byte: an HBM-bound kernel.  The one-tile-per-CTA flavour of
// attn_sm100.cu (two CTAs per SM, grid over (query tile, head, sequence)) ran it at 2.1 TB/s, and ncu showed why: 1.94 GB
// read from DRAM for 0.79 GB of queries.  A head is a 144-byte slice of a 2304-byte row; with the 16 heads of a row
// block read at sixteen unrelated moments every slice drags its neighbours' sectors in (and every 144-byte output slice
// its partial sectors) and nothing is left in L2 when the neighbour is finally wanted.  Here ONE CTA per SM walks a
// contiguous range of items ordered (sequence, query tile, HEAD fastest): the sixteen heads of a 128-row block are read
// and written back to back by the same SM, so DRAM sees whole rows.
//   warp 0      TMA producer: per item Q (3-slot ring), K (2 slots) and V (3 slots) — K/V are L2-resident (17 MB)
//   warp 1      MMA issuer: S_b = Q·K^T (SS, both operands in shared memory), O_b = P_b·V (P from TMEM over S_b);
//               two TMEM buffers b = item parity, issue order  P·V(i), Q·K^T(i + 2)  so the scores of the next item of
//               a buffer are under way while its softmax warps still read O.  Two commits per item (S ready, O ready)
//   warps 2-5   softmax + epilogue of the even items, warps 6-9 of the odd items (thread = query row = TMEM lane):
//               exact single-tile softmax (row maximum, 2^(s·c − m·c), row sum), P as bf16 pairs over S, then O / l
//               rounded to bf16, parked in shared memory and written by ONE TMA store per warp (32 rows of one head;
//               the 4-D output map clips the rows past the end of the sequence).  A bulk copy per ROW (128 copies of
//               144 B per tile) was tried first: 2.8 us per tile, the copy engine's per-operation cost.
//               The warps that have seen "S ready" / "O ready" hand the Q + K / the V slot back to the producer with a
//               plain mbarrier arrive — nothing asynchronous is outstanding when the CTA retires.
#include <cstdlib>

#include "attn_common.cuh"

using namespace sm100;
using namespace attn;

namespace {
// suspend-time hints (ns) of the producer's / the issuer's barrier waits; tuning hooks.  0 = plain try_wait spin.
#ifndef MDV2_XATTN_MMA_HINT
#define MDV2_XATTN_MMA_HINT 0
#define MDV2_XATTN_TMA_HINT 200
#endif
constexpr uint32_t MMA_HINT = MDV2_XATTN_MMA_HINT, TMA_HINT = MDV2_XATTN_TMA_HINT;
__device__ __forceinline__ void mbar_wait_hint(uint64_t* bar, uint32_t parity, uint32_t hint_ns) {
  if (hint_ns == 0) mbar_wait(bar, parity);
  else mbar_wait_parked(bar, parity, hint_ns);
}
constexpr int BNK = 64;           // keys per K/V tile = all of them
// Rings: Q is the only operand that comes from DRAM (K / V of a sequence are 0.3 MB and L2-resident), and at ~1.5 us of
// loaded DRAM latency an SM needs ~65 KB of reads in flight to draw its share of the bandwidth: Q runs Q_AHEAD items
// ahead of K.  (With a 3-slot Q ring and one in-order producer stream the issuer spent 31 % of its time waiting for Q.)
#ifndef MDV2_XATTN_QS
#define MDV2_XATTN_QS 5
#define MDV2_XATTN_VS 3
#endif
constexpr int Q_SLOTS = MDV2_XATTN_QS, K_SLOTS = 3, V_SLOTS = MDV2_XATTN_VS, Q_AHEAD = Q_SLOTS - K_SLOTS;
#ifndef MDV2_XATTN_LSU_STORE
#define MDV2_XATTN_LSU_STORE 0
#endif
constexpr bool LSU_STORE = MDV2_XATTN_LSU_STORE != 0;   // output rows by coalesced st.global instead of TMA stores
constexpr int V_LAG = V_SLOTS == 3 ? 1 : 2;      // step i requests V(i - V_LAG): its slot comes back with P·V(i - 4)
constexpr int N_SM_WARPS = 8;     // two groups of four
constexpr int NTHREADS = (2 + N_SM_WARPS) * 32;
constexpr int TMEM_COLS = 512;    // S_b (P_b over it) at 128·b, O_b at 256 + 128·b
__host__ __device__ constexpr int s_col(int b) { return b * 128; }
__host__ __device__ constexpr int o_col(int b) { return 256 + b * 128; }

template <int DP>
struct XCfg {
  // a head of 72 = one 64-wide d-chunk (128-byte rows, SW128) + a 16-wide tail chunk.  The tail of Q and K (K-major MMA
  // operands, one k16 step) is a 32-byte-row SW32 slab — 4 KB instead of the 16 KB a zero-filled 64-wide box takes — so
  // that SIX query tiles fit the ring; V (MN-major, both chunks in one MMA) keeps two SW128 chunks
  static constexpr bool TAIL = DP > 64;
  static constexpr int Q_MAIN = BMQ * 128, Q_BYTES = Q_MAIN + (TAIL ? BMQ * 32 : 0);
  static constexpr int K_MAIN = BNK * 128, K_BYTES = K_MAIN + (TAIL ? BNK * 32 : 0);
  static constexpr int V_BYTES = (TAIL ? 2 : 1) * BNK * 128;
  static constexpr int ROW_BYTES = DP == 64 ? 128 : 144;      // staged output row = D bf16, dense (the TMA box)
  static constexpr int STAGE_BYTES = 2 * BMQ * ROW_BYTES;     // output rows of the two item parities
  static constexpr int SMEM = Q_SLOTS * Q_BYTES + K_SLOTS * K_BYTES + V_SLOTS * V_BYTES + STAGE_BYTES + 256 + 1024;
  static_assert(Q_BYTES % 1024 == 0 && K_BYTES % 1024 == 0 && V_BYTES % 1024 == 0, "slots keep the swizzle alignment");
};

template <int DP>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_cross_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                  const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap to,
                  const __grid_constant__ CUtensorMap tq1, const __grid_constant__ CUtensorMap tk1, const AParams p,
                  const int n_qt, const int n_items, const int teams) {
  using C = XCfg<DP>;
  extern __shared__ uint8_t smem_raw[];
  // (pointer arithmetic on the __shared__ array keeps the address space: the staging stores stay STS, not generic ST.E)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                                   // slot s at sQ + s·Q_BYTES
  uint8_t* sK = sQ + Q_SLOTS * C::Q_BYTES;              // slot s at sK + s·K_BYTES
  uint8_t* sV = sK + K_SLOTS * C::K_BYTES;
  uint8_t* sO = sV + V_SLOTS * C::V_BYTES;              // [item parity][row] staged output rows
  uint64_t* bars = reinterpret_cast<uint64_t*>(sO + C::STAGE_BYTES);
  uint64_t* q_full = bars;                              // Q_SLOTS
  uint64_t* q_empty = q_full + Q_SLOTS;
  uint64_t* k_full = q_empty + Q_SLOTS;                 // K_SLOTS
  uint64_t* k_empty = k_full + K_SLOTS;
  uint64_t* v_full = k_empty + K_SLOTS;                 // V_SLOTS
  uint64_t* v_empty = v_full + V_SLOTS;
  uint64_t* s_full = v_empty + V_SLOTS;                 // [item parity]
  uint64_t* p_full = s_full + 2;
  uint64_t* o_full = p_full + 2;
  uint64_t* o_free = o_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  // this CTA's contiguous range of items; item -> (sequence, query tile, head), head fastest
  // Work split.  teams > 0 (experiment, off: see the launcher): the grid is `teams` teams of H CTAs; a team walks the
  // 128-row query blocks (sequence, query tile) team, team + teams, ... and CTA j of the team owns head j of every block —
  // the H slices of a row block are wanted by H SMs at about the same time (DRAM sees whole rows, L2 serves the
  // neighbours' sectors) and a CTA's head never changes.  teams == 0 (default): a contiguous range of (sequence, query
  // tile, head) items per CTA, head fastest.
  const int64_t G = gridDim.x;
  const int team = teams > 0 ? (int)blockIdx.x / p.H : 0, my_head = teams > 0 ? (int)blockIdx.x % p.H : 0;
  const int n_rb = n_items / p.H;                       // row blocks
  const int i0 = teams > 0 ? 0 : (int)((int64_t)n_items * blockIdx.x / G);
  const int n_my = teams > 0 ? (n_rb - team + teams - 1) / teams
                             : (int)((int64_t)n_items * (blockIdx.x + 1) / G) - i0;
  // local item i -> (head, query tile, sequence)
  auto coords = [&](int i, int& h, int& qt, int& seq) {
    int r;
    if (teams > 0) {
      h = my_head;
      r = team + teams * i;
    } else {
      const int it = i0 + i;
      r = it / p.H;
      h = it - r * p.H;
    }
    seq = r / n_qt;
    qt = r - seq * n_qt;
  };

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tq);
    prefetch_tmap(&tk);
    prefetch_tmap(&tv);
    prefetch_tmap(&to);
    if (C::TAIL) {
      prefetch_tmap(&tq1);
      prefetch_tmap(&tk1);
    }
    for (int i = 0; i < Q_SLOTS; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
    }
    for (int i = 0; i < K_SLOTS; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    for (int i = 0; i < V_SLOTS; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], 4);
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
    // step i requests V(i - V_LAG) — its slot comes back with P·V(i - 4) — then K(i) and Q(i + Q_AHEAD), whose slots came
    // back earlier, with Q·K^T(i - K_SLOTS).  V goes first: the copy engine serves requests in order, and a V tile queued
    // behind 384 row requests of Q and K reached the issuer late (it waited for V 37 % of its time, ncu)
    if (lane == 0) {
      for (int i = -Q_AHEAD; i < n_my + 2; ++i) {
        int h, qt, seq;
        const int jv = i - V_LAG;                      // first: V is the load with the least slack
        if (jv >= 0 && jv < n_my) {
          coords(jv, h, qt, seq);
          const int vs = jv % V_SLOTS;
          mbar_wait_hint(&v_empty[vs], (uint32_t)(((jv / V_SLOTS) & 1) ^ 1), TMA_HINT);
          if ((p.dbg & 1) && jv >= V_SLOTS) {
            mbar_arrive(&v_full[vs]);
          } else {
            mbar_expect_tx(&v_full[vs], C::V_BYTES);
#pragma unroll
            for (int c = 0; c < (C::TAIL ? 2 : 1); ++c)
              tma_load_3d(sV + vs * C::V_BYTES + c * BNK * 128, &tv, &v_full[vs], c * 64, h, seq * p.Lk);
          }
        }
        const int jq = i + Q_AHEAD;
        if (jq < n_my) {
          coords(jq, h, qt, seq);
          const int qs = jq % Q_SLOTS;
          uint8_t* dst = sQ + qs * C::Q_BYTES;
          mbar_wait_hint(&q_empty[qs], (uint32_t)(((jq / Q_SLOTS) & 1) ^ 1), TMA_HINT);
          if ((p.dbg & 4) && jq >= Q_SLOTS) {            // timing experiment (bit 2): stale Q
            mbar_arrive(&q_full[qs]);
          } else {
            mbar_expect_tx(&q_full[qs], C::Q_BYTES);
            tma_load_3d(dst, &tq, &q_full[qs], 0, h, seq * p.Lq + qt * BMQ);
            if (C::TAIL) tma_load_3d(dst + C::Q_MAIN, &tq1, &q_full[qs], 64, h, seq * p.Lq + qt * BMQ);
          }
        }
        if (i >= 0 && i < n_my) {
          coords(i, h, qt, seq);
          const int ks = i % K_SLOTS;
          uint8_t* dst = sK + ks * C::K_BYTES;
          mbar_wait_hint(&k_empty[ks], (uint32_t)(((i / K_SLOTS) & 1) ^ 1), TMA_HINT);
          if ((p.dbg & 1) && i >= K_SLOTS) {             // timing experiment (MDV2_ATTN_DBG bit 0): stale K / V, wrong results
            mbar_arrive(&k_full[ks]);
          } else {
            mbar_expect_tx(&k_full[ks], C::K_BYTES);
            tma_load_3d(dst, &tk, &k_full[ks], 0, h, seq * p.Lk);
            if (C::TAIL) tma_load_3d(dst + C::K_MAIN, &tk1, &k_full[ks], 64, h, seq * p.Lk);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc_qk = make_idesc_bf16(BMQ, BNK, 0, 0);
    constexpr uint32_t idesc_pv = make_idesc_bf16(BMQ, DP, 0, 1);       // A = P from TMEM, B = V is MN-major
    const uint64_t dq0 = make_desc_k_sw128(smem_u32(sQ)), dq1 = make_desc_k_sw32(smem_u32(sQ + C::Q_MAIN));
    const uint64_t dk0 = make_desc_k_sw128(smem_u32(sK)), dk1 = make_desc_k_sw32(smem_u32(sK + C::K_MAIN));
    const uint64_t dv0 = make_desc_mn_sw128(smem_u32(sV), BNK * 128, 1024);
    // S_b(i) = Q(i) · K(i)^T
    auto issue_qk = [&](int i) {
      const int ks = i % K_SLOTS, qs = i % Q_SLOTS, b = i & 1;
      mbar_wait_hint(&k_full[ks], (uint32_t)((i / K_SLOTS) & 1), MMA_HINT);
      mbar_wait_hint(&q_full[qs], (uint32_t)((i / Q_SLOTS) & 1), MMA_HINT);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t oq = (uint64_t)(qs * (C::Q_BYTES >> 4)), ok = (uint64_t)(ks * (C::K_BYTES >> 4));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)                   // d = 0 .. 63: 32-byte steps inside the SW128 rows
          umma_bf16_ss(tmem_base + s_col(b), dq0 + oq + (uint64_t)(kk * 2), dk0 + ok + (uint64_t)(kk * 2), idesc_qk, kk != 0);
        if (C::TAIL) umma_bf16_ss(tmem_base + s_col(b), dq1 + oq, dk1 + ok, idesc_qk, 1);   // d = 64 .. 79 (72 .. 79 are zeros)
        umma_commit(&s_full[b]);                         // also tells the softmax warps that Q and K have been read
      }
      __syncwarp();
    };
    if (n_my > 0) issue_qk(0);
    if (n_my > 1) issue_qk(1);
    for (int i = 0; i < n_my; ++i) {
      const int b = i & 1, vs = i % V_SLOTS;
      mbar_wait_hint(&v_full[vs], (uint32_t)((i / V_SLOTS) & 1), MMA_HINT);
      if (i >= 2) mbar_wait_hint(&o_free[b], (uint32_t)(((i >> 1) - 1) & 1), MMA_HINT);   // O_b of item i - 2 has been read
      mbar_wait_hint(&p_full[b], (uint32_t)((i >> 1) & 1), MMA_HINT);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t dv = dv0 + (uint64_t)(vs * (C::V_BYTES >> 4));
#pragma unroll
        for (int kk = 0; kk < BNK / 16; ++kk)
          umma_bf16_ts(tmem_base + o_col(b), tmem_base + s_col(b) + kk * 8, dv + (uint64_t)((kk * 16) * 128 >> 4), idesc_pv,
                       kk != 0);
        umma_commit(&o_full[b]);                         // also: V has been read
      }
      __syncwarp();
      if (i + 2 < n_my) issue_qk(i + 2);                 // over the P this P·V has just consumed (in-order pipe)
    }
  } else {
    // ================= softmax + epilogue: group b = (warp - 2) / 4 takes the items of parity b =================
    const int b = (warp - 2) >> 2;
    const int quarter = warp & 3;                        // TMEM lane quarter this warp may touch
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_addr = lane_addr + s_col(b), o_addr = lane_addr + o_col(b);
    uint8_t* swarp = sO + (b * BMQ + quarter * 32) * C::ROW_BYTES;   // this warp's 32 staged rows: one TMA box
    uint8_t* srow = swarp + lane * C::ROW_BYTES;
    const float c = p.scale_log2;
    // (sequence, query tile, head) of the warp's current item, stepped by two items per turn: four runtime divisions
    // per item were 7 % of the kernel's instructions, on the chain between two tiles
    int h, qt, seq;
    coords(b, h, qt, seq);
    for (int i = b; i < n_my; i += 2) {
      const uint32_t ph = (uint32_t)((i >> 1) & 1);
      mbar_wait(&s_full[b], ph);
      tc_fence_after();
      if (quarter == 0 && lane == 0) {                   // Q·K^T has completed: its operands go back to the producer
        mbar_arrive(&q_empty[i % Q_SLOTS]);
        mbar_arrive(&k_empty[i % K_SLOTS]);
      }
      uint32_t sr[BNK];
      tmem_ld_x32(s_addr, sr);
      tmem_ld_x32(s_addr + 32, sr + 32);
      tmem_ld_wait();
      // keys past Lk do not exist
      if (p.Lk < BNK) {                                  // (a compare + select per score was 14 % of the instructions)
        if (p.Lk >= BNK - 8) {
#pragma unroll
          for (int j = BNK - 8; j < BNK; ++j)
            if (j >= p.Lk) sr[j] = 0xff800000u;
        } else {
#pragma unroll
          for (int j = 0; j < BNK; ++j)
            if (j >= p.Lk) sr[j] = 0xff800000u;
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int j = 0; j < BNK; j += 4) {
        mx0 = fmax3(mx0, __uint_as_float(sr[j]), __uint_as_float(sr[j + 1]));
        mx1 = fmax3(mx1, __uint_as_float(sr[j + 2]), __uint_as_float(sr[j + 3]));
      }
      const float moff = fmaxf(mx0, mx1) * c;
      const float2 c2 = make_float2(c, c), nm2 = make_float2(-moff, -moff);
      float2 la = make_float2(0.f, 0.f), lb = make_float2(0.f, 0.f);
      uint32_t pk[BNK / 2];
#pragma unroll
      for (int j = 0; j < BNK; j += 4) {
        const float2 x01 = ffma2(make_float2(__uint_as_float(sr[j]), __uint_as_float(sr[j + 1])), c2, nm2);
        const float2 x23 = ffma2(make_float2(__uint_as_float(sr[j + 2]), __uint_as_float(sr[j + 3])), c2, nm2);
        const float2 e01 = make_float2(ex2(x01.x), ex2(x01.y)), e23 = make_float2(ex2(x23.x), ex2(x23.y));
        la = fadd2(la, e01);
        lb = fadd2(lb, e23);
        __nv_bfloat162 b01 = __floats2bfloat162_rn(e01.x, e01.y), b23 = __floats2bfloat162_rn(e23.x, e23.y);
        pk[j / 2] = *reinterpret_cast<uint32_t*>(&b01);
        pk[j / 2 + 1] = *reinterpret_cast<uint32_t*>(&b23);
      }
      const float l_row = (la.x + la.y) + (lb.x + lb.y);
      tmem_st_x32(s_addr, pk);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[b]);
      // ---- O = P·V back, normalised, through shared memory, one TMA store per warp ----
      if constexpr (!LSU_STORE) {
        if (lane == 0) bulk_wait_read0();                // the warp's previous store (two items ago) has left shared memory
        __syncwarp();
      }
      mbar_wait(&o_full[b], ph);
      tc_fence_after();
      if (quarter == 0 && lane == 0) mbar_arrive(&v_empty[i % V_SLOTS]);   // P·V has completed: V goes back
      uint32_t r[DP];
      tmem_ld_x32(o_addr, r);
      tmem_ld_x32(o_addr + 32, r + 32);
      if constexpr (DP > 64) tmem_ld_x16(o_addr + 64, r + 64);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_free[b]);
      const float inv = 1.f / l_row;
#pragma unroll
      for (int u = 0; u < C::ROW_BYTES / 16; ++u) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[u * 8 + e]) * inv;
        Vec16<bf16> o;
        o.pack(f);
        // D = 64: 128-byte rows, SW128 like the store map; D = 72: 144-byte rows already rotate over the bank groups
        const int uu = DP == 64 ? (u ^ (lane & 7)) : u;
        *reinterpret_cast<uint4*>(srow + uu * 16) = o.raw;
      }
      if constexpr (LSU_STORE) {
        // rows back out of the staging buffer with coalesced 16-byte stores (whole 144-byte row pieces per warp turn)
        __syncwarp();
        constexpr int UPR = C::ROW_BYTES / 16;
        const int r0 = qt * BMQ + quarter * 32;
        bf16* obase = p.out + ((int64_t)seq * p.Lq + r0) * p.ldo + (int64_t)h * p.D;
        if (!(p.dbg & 2)) {
#pragma unroll
          for (int t = 0; t < UPR; ++t) {
            const int idx = lane + 32 * t, rl = idx / UPR, u = idx - rl * UPR;
            const int uu = DP == 64 ? (u ^ (rl & 7)) : u;
            if (r0 + rl < p.Lq)
              *reinterpret_cast<uint4*>(obase + (int64_t)rl * p.ldo + u * 8) =
                  *reinterpret_cast<const uint4*>(swarp + rl * C::ROW_BYTES + uu * 16);
          }
        }
        __syncwarp();
      } else {
        fence_async_smem();                              // generic-proxy writes -> the copy engine
        __syncwarp();
        if (lane == 0) {
          const int r0 = qt * BMQ + quarter * 32;
          if (r0 < p.Lq && !(p.dbg & 2)) tma_store_4d(&to, swarp, 0, h, r0, seq);
          bulk_commit();
        }
      }
      if (teams > 0) {
        qt += 2 * teams;
        while (qt >= n_qt) {
          qt -= n_qt;
          ++seq;
        }
      } else {
        h += 2;
        while (h >= p.H) {
          h -= p.H;
          if (++qt == n_qt) {
            qt = 0;
            ++seq;
          }
        }
      }
    }
    if (!LSU_STORE && lane == 0) bulk_wait_read0();      // shared memory must outlive the engine's reads
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int DP>
int launch_cross(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p, int n_seq,
                 const void* q, int64_t ldq, const void* k, int64_t ldk, cudaStream_t st) {
  CUtensorMap to, tq1 = tq, tk1 = tk;
  if (XCfg<DP>::TAIL) {
    int rc1 = mdv2_make_tmap_3d_heads_sw32(&tq1, q, (uint64_t)n_seq * p.Lq, (uint32_t)p.H, (uint32_t)p.D, (uint64_t)ldq, BMQ);
    if (rc1) return rc1;
    rc1 = mdv2_make_tmap_3d_heads_sw32(&tk1, k, (uint64_t)n_seq * p.Lk, (uint32_t)p.H, (uint32_t)p.D, (uint64_t)ldk, BNK);
    if (rc1) return rc1;
  }
  int rc = mdv2_make_tmap_4d_store(&to, p.out, (uint64_t)n_seq, (uint64_t)p.Lq, (uint32_t)p.H, (uint32_t)p.D, (uint64_t)p.ldo, 32);
  if (rc) return rc;
  constexpr int SMEM = XCfg<DP>::SMEM;
  static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_cross_kernel<DP>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("attention_cross: cannot set %d B dynamic smem: %s", SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    configured = true;
  }
  const int n_qt = (p.Lq + BMQ - 1) / BMQ;
  const int64_t n_items = (int64_t)n_qt * p.H * n_seq;
  MDV2_REQUIRE(n_items <= 0x7fffffff, "attention_cross: too many work items");
  int64_t grid = mdv2_num_sms();
  if (grid > n_items) grid = n_items;
  // MDV2_ATTN_CROSS_TEAMS=1: teams of H CTAs in lock-step on one row block (see the kernel).  Measured 0.52 ms against
  // 0.36 ms for contiguous item ranges at 60 x 5600 x 62: sixteen SMs asking for the same 128 rows at the same moment
  // queue on the same L2 slices / DRAM pages instead of sharing them.  Off.
  static const int team_mode = [] {
    const char* e = getenv("MDV2_ATTN_CROSS_TEAMS");
    return e ? atoi(e) : 0;
  }();
  int teams = 0;
  if (team_mode && p.H <= mdv2_num_sms()) {
    teams = mdv2_num_sms() / p.H;
    const int64_t n_rb = (int64_t)n_qt * n_seq;
    if (teams > n_rb) teams = (int)n_rb;
    grid = (int64_t)teams * p.H;
  }
  attn_cross_kernel<DP><<<(unsigned)grid, NTHREADS, SMEM, st>>>(tq, tk, tv, to, tq1, tk1, p, n_qt, (int)n_items, teams);
  MDV2_CHECK_LAUNCH("attention_cross");
  return MDV2_OK;
}
}  // namespace

// called by mdv2_attention_sm100 (attn_sm100.cu) for Lk <= 64, one K/V set, no output scatter (MDV2_ATTN_CROSS=0: the
// one-tile-per-CTA flavour of attn_sm100.cu)
int mdv2_attention_cross_launch(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AParams& p,
                                int n_seq, int D, const void* q, int64_t ldq, const void* k, int64_t ldk, cudaStream_t st) {
  if (D == 72) return launch_cross<80>(tq, tk, tv, p, n_seq, q, ldq, k, ldk, st);
  return launch_cross<64>(tq, tk, tv, p, n_seq, q, ldq, k, ldk, st);
}
