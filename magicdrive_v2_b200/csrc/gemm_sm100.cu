// gemm_sm100.cu — bf16 GEMM  out = epilogue(A[M,K] · W[N,K]^T)  on the 5th-gen tensor cores.
//
// Persistent, warp-specialised, CTA PAIRS (cluster of 2, tcgen05 cta_group::2): one pair computes a
// 256 x BN output tile.  Each CTA loads its 128 rows of A and its BN/2 rows of W per 64-wide k-block
// (TMA, 128-byte swizzle; the bytes of both CTAs are counted on the leader's mbarrier), the leader's
// warp 1 issues UMMA 256 x BN x 16 for the pair, and each CTA's TMEM receives the fp32 accumulators of
// its own 128 rows.  Sharing W across the pair halves the shared-memory traffic per SM — the single-CTA
// 128 x 256 tile measured 68 % tensor-pipe activity, exactly the 128 B/clk smem limit (ncu, profiles/).
// warp 0 = TMA producer, warp 1 = MMA issuer (leader) + TMEM owner, warps 2-9 = epilogue (TMEM ->
// registers -> fused bias / per-head RMSNorm / GELU -> shared staging -> gate / residual -> coalesced
// stores).  Two TMEM accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
#include <cstdlib>
#include <mutex>
#include <string>
#include <unordered_map>

#include "gemm_epilogue.cuh"
#include "sm100_ptx.cuh"

using namespace sm100;

namespace {
constexpr int BM = 128, BK = 64;                     // rows per CTA (256 per pair), k-block
constexpr int A_BYTES = BM * BK * 2;
constexpr int NUM_EPI_WARPS = 8;                      // 2 per TMEM lane quarter, interleaved over column chunks
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
constexpr int TMEM_COLS = 512, ACC_STRIDE = 256;

template <int BN, int HD>
struct Cfg {
  static constexpr int CW = HD ? HD : 32;                // epilogue chunk width (columns)
  // BN = 216 = three heads of 72: 216 is not an MMA width (N % 16 == 0), so the MMA runs 224 wide — each CTA of the pair
  // fetches 112 W rows, the last 8 of the pair belong to the next tile (or are TMA zero fill) and their accumulator
  // columns are never read.  3.7 % of wasted tensor work for 20 % less operand traffic per FLOP than two heads per tile.
  static constexpr int BNM = BN == 216 ? 224 : BN;       // MMA width = W rows fetched per pair and k-block
  static constexpr int NCH = BN / CW;                    // chunks per tile
  static constexpr int CPW = (NCH + 1) / 2;              // chunks per epilogue warp and tile (column halves interleave)
  static constexpr int B_BYTES = (BNM / 2) * BK * 2;     // this CTA's half of the W tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int PITCH = ((CW * 2 / 16) % 2 == 1) ? CW * 2 : CW * 2 + 16;   // odd multiple of 16 B
  static constexpr int EPW = 32 * PITCH + 512 + CPW * CW * 4;   // per warp: staging + row info + bias (fp32)
  static constexpr int STAGING = NUM_EPI_WARPS * EPW;
  static constexpr int RMSW = HD ? 2 * HD * 4 : 0;       // q / k RMSNorm weights (fp32), staged once
  static constexpr int STAGES_RAW = (232448 - 1024 - 256 - STAGING - RMSW) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int SMEM = STAGES * STAGE_BYTES + STAGING + RMSW + 256 + 1024;  // + barriers + align slack
  static_assert(BN % CW == 0 && BNM <= ACC_STRIDE, "whole chunks, one accumulator stage per tile");
  static_assert(EPW % 16 == 0, "per-warp staging must stay 16-byte aligned");
  static_assert((PITCH / 16) % 2 == 1, "staging pitch must be an odd multiple of 16 B");
  static_assert(B_BYTES % 1024 == 0, "B stage must keep 1024-B alignment");
};

// Optional scattered output of the qkv projection under sequence parallelism (HD != 0 flavour only): the N = 3·W
// output columns are q | k | v of width W; head group g = (col % W) / grp of every part goes to dst[g] (the peers of a
// sequence-parallel group, mapped through NVLink) in the layout the attention kernel reads,
//   dst[g][seq][row_base + r][part][grp]   with seq = row / rows_per_seq, r = row % rows_per_seq,
// i.e. the head-scatter / sequence-gather all-to-all is done by the epilogue's own stores, under the MMAs.
struct ScatterOut {
  bf16* dst[8];
  int n_dst;                 // 0: plain output
  int grp, part_w, rows_per_seq;
  int64_t dst_ld, seq_rows, row_base;
};

struct Params {
  ScatterOut sc;
  int64_t M;
  int N, K;
  int64_t ldo;
  bf16* out;
  Epilogue<bf16> ep;
  const bf16* rms_q;
  const bf16* rms_k;
  int n_q, n_k;
  float eps;
};

// MC = CTA pairs per cluster.  MC = 2 (cluster of four): the two pairs compute N-ADJACENT tiles of the same 256 rows and
// share the A operand through TMA multicast — each of the four CTAs fetches 64 of the 128 A rows it needs and
// multicasts them to the CTA of the other pair that needs the same rows.  Why: with one pair per cluster every CTA pulls
// (128 + BN/2) x 128 B per k-block through the L2 -> SM fabric, 64 / 75 / 89 B per clock and SM at full tensor rate for
// BN = 256 / 192 / 144; sharing A brings that to 47 / 52 / 59 B/clk/SM.  (Experiment, off by default: see launch().)
template <int BN, int HD, int MC>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                  const __grid_constant__ CUtensorMap tmap_wr, const Params p) {
  using C = Cfg<BN, HD>;
  extern __shared__ uint8_t smem_raw[];
  // aligned up to 1024 B by POINTER arithmetic on the __shared__ array: a round trip through uintptr_t loses the address
  // space, and every staging access of the epilogue became a generic LD.E / ST.E (long-scoreboard, ~5x the latency of
  // LDS: the dependent staging load of the column phase was 15 % of the gate + residual GEMMs' stall samples)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* staging = smem + C::STAGES * C::STAGE_BYTES;
  uint32_t* s_rms = reinterpret_cast<uint32_t*>(staging + C::STAGING);   // [2][HD / 2]: q, k RMSNorm weights as bf16 pairs
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + C::STAGING + C::RMSW);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::STAGES;
  uint64_t* tfull = bars + 2 * C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t rank = crank & 1;                      // within the pair: 0 = leader (issues the pair's MMAs)
  const uint32_t q = crank >> 1;                        // pair within the cluster: N tile n_super·MC + q
  const int64_t pair = blockIdx.x / (2 * MC), n_pairs = gridDim.x / (2 * MC);      // cluster index / clusters
  // Column tiles: BN wide, plus — when N is not a multiple of BN — ONE narrower last tile of R = N % BN columns (R a
  // multiple of 64; MC == 1 only).  N = 1152 runs as 4 x 256 + 128 instead of 6 x 192: the L2 -> SM fabric, not the tensor
  // pipe, bounds these GEMMs (a CTA pulls (128 + BN/2) x 128 B per k-block: 64 / 75 / 89 B per clock at full tensor rate
  // for BN = 256 / 192 / 144), so the widest tile wins and the remainder is cheaper as one narrow tile than as 192s.
  const int R = MC == 1 ? p.N % BN : 0;
  const int n_tiles_n = p.N / BN / MC + (R ? 1 : 0);    // cluster tiles along N (MC adjacent BN-wide tiles each)
  const int64_t n_tiles_m = (p.M + 2 * BM - 1) / (2 * BM);
  const int64_t n_tiles = n_tiles_m * n_tiles_n;
  const int n_kb = p.K / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_w);
    for (int i = 0; i < C::STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], MC);                         // one commit per pair that reads (or multicasts into) the stage
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 2 * NUM_EPI_WARPS);         // the epilogue warps of BOTH CTAs (leader's copy is used)
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, TMEM_COLS);
  if constexpr (HD != 0) {
    for (int i = threadIdx.x; i < HD; i += NUM_THREADS) {                 // pair i % (HD/2) of the q (i < HD/2) or k weights
      const bf16* w = i < HD / 2 ? p.rms_q : p.rms_k;
      const int j = (i % (HD / 2)) * 2;
      const __nv_bfloat162 one = __floats2bfloat162_rn(1.f, 1.f);
      __nv_bfloat162 v2 = w ? __halves2bfloat162(w[j], w[j + 1]) : one;
      s_rms[i] = *reinterpret_cast<uint32_t*>(&v2);
    }
  }
  tc_fence_before();
  cluster_sync();                                        // barriers + TMEM of both CTAs are ready (and s_rms)
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t tile = pair; tile < n_tiles; tile += n_pairs) {
        const int m0 = (int)(tile / n_tiles_n) * 2 * BM + (int)rank * BM;
        const int tn = (int)(tile % n_tiles_n);
        const bool narrow = R != 0 && tn == n_tiles_n - 1;
        const int n0 = (tn * MC + (int)q) * BN + (int)rank * ((narrow ? R : C::BNM) / 2);
        const uint32_t tx = narrow ? 2 * (A_BYTES + (R / 2) * BK * 2) : 2 * C::STAGE_BYTES;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          if (rank == 0) mbar_expect_tx(&full[stage], tx);
          if constexpr (MC == 1) {
            tma_load_2d_pair(sa, &tmap_a, &full[stage], kb * BK, m0);
          } else {
            // rows [q·64, q·64 + 64) of this CTA's 128 A rows, to itself and to the same-rank CTA of the other pair
            tma_load_2d_pair_mc(sa + q * (A_BYTES / MC), &tmap_a, &full[stage], kb * BK, m0 + (int)q * (BM / MC),
                                (uint16_t)(0x5u << rank));
          }
          tma_load_2d_pair(sa + A_BYTES, narrow ? &tmap_wr : &tmap_w, &full[stage], kb * BK, n0);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    // The whole warp walks the loop and ONE ELECTED lane issues the four MMAs + the commit of a k-block.  Under a plain
    // `lane == 0` branch the compiler cannot keep the descriptors in uniform registers: every UTCHMMA came wrapped in an
    // election loop with seven register -> uniform-register moves, ~19 instructions per MMA on one lane — longer than a
    // 256 x 144 x 16 MMA takes to execute, so the two-head qkv tiles were issue-bound (the same finding as in the
    // attention kernel, attn_sm100.cu).
    if (rank == 0) {
      constexpr uint32_t idesc_full = make_idesc_bf16(2 * BM, C::BNM);
      const uint32_t idesc_narrow = make_idesc_bf16(2 * BM, R ? R : BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int64_t tile = pair; tile < n_tiles; tile += n_pairs) {
        const uint32_t idesc = (R != 0 && (int)(tile % n_tiles_n) == n_tiles_n - 1) ? idesc_narrow : idesc_full;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * ACC_STRIDE;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_u32(smem + stage * C::STAGE_BYTES);
            const uint64_t da = make_desc_k_sw128(sa), db = make_desc_k_sw128(sa + A_BYTES);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_ss_pair(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
            umma_commit_mc(&empty[stage], (uint16_t)((1u << (2 * MC)) - 1));     // every CTA of the cluster
          }
          __syncwarp();
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit_mc(&tfull[acc], (uint16_t)(3u << (2 * q)));  // the two CTAs of this pair
        __syncwarp();
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue warps: TMEM lane quarter = warp % 4, column chunks ch ≡ half (mod 2) =====
    // Everything the epilogue reads from global memory is requested BEFORE it is needed: the bias of all of the
    // warp's chunks goes to shared memory ahead of the accumulator wait, and the residual / gate vectors of
    // chunk i+1 are loaded into registers before chunk i is processed (a dependent load per 16-byte store had
    // the K = 1152 GEMMs with gate + residual running at 24 % tensor-pipe activity, ncu profiles/).
    constexpr int CW = C::CW, CPW = C::CPW;
    constexpr int UPR = CW / 8;                       // 16-byte units per staged row
    constexpr int NIT = UPR;                          // column-phase iterations: 32 rows x UPR units / 32 lanes
    constexpr bool FIXED_U = (32 % UPR) == 0;         // lane -> constant column unit (gate vector reused over rows)
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    uint8_t* stg = staging + (warp - 2) * C::EPW;
    int64_t* row_res = reinterpret_cast<int64_t*>(stg + 32 * C::PITCH);
    int64_t* row_gate = row_res + 32;
    float* sbias = reinterpret_cast<float*>(row_gate + 32);   // [CPW][CW]
    int acc = 0;
    uint32_t acc_phase = 0;
    const Epilogue<bf16>& ep = p.ep;
    // the RMSNorm flavour (qkv projections) never carries a gate or a residual (api.cu rejects the combination)
    const bool has_res = HD == 0 && ep.res != nullptr, has_gate = HD == 0 && ep.gate_mod != nullptr;
    for (int64_t tile = pair; tile < n_tiles; tile += n_pairs) {
      const int64_t m0 = (tile / n_tiles_n) * 2 * BM + (int64_t)rank * BM;
      const int tn = (int)(tile % n_tiles_n);
      const int n0 = (tn * MC + (int)q) * BN;
      // chunks of this warp in the tile: all CPW of them, fewer in the narrow last tile (>= 1: R is a multiple of 2·CW)
      const int n_ci = ((R != 0 && tn == n_tiles_n - 1 ? R / CW : C::NCH) - half + 1) / 2;
      const int64_t r0 = m0 + quarter * 32;          // first row of this warp
      {  // per-row epilogue context, once per tile (64-bit divisions stay out of the inner loops)
        const int64_t gr = r0 + lane;
        const RowCtx rc = ep.row_ctx(gr < p.M ? gr : 0);
        row_res[lane] = rc.res_row * ep.ldr;
        row_gate[lane] = rc.gate ? (int64_t)(rc.gate - ep.gate_mod) : -1;
        if constexpr (HD != 0) {
          if (p.sc.n_dst > 0) {                      // destination row of the scattered qkv output
            const int64_t g = gr < p.M ? gr : 0, sq = g / p.sc.rows_per_seq;
            row_res[lane] = (sq * p.sc.seq_rows + p.sc.row_base + (g - sq * p.sc.rows_per_seq)) * p.sc.dst_ld;
          }
        }
      }
      if constexpr (CW == 32) {
        // one bias element per lane and chunk: all of them requested before the first is converted (a load, its
        // conversion and its store per chunk exposed the L2 latency CPW times per tile)
        bf16 bv[CPW];
#pragma unroll
        for (int ci = 0; ci < CPW; ++ci)
          if (ep.bias && ci < n_ci) bv[ci] = ep.bias[n0 + (half + 2 * ci) * CW + lane];
#pragma unroll
        for (int ci = 0; ci < CPW; ++ci)
          if (ci < n_ci) sbias[ci * CW + lane] = ep.bias ? ep.bias_scale * __bfloat162float(bv[ci]) : 0.f;
      } else {
#pragma unroll
        for (int ci = 0; ci < CPW; ++ci) {
          const int c0 = n0 + (half + 2 * ci) * CW;
          if (ci < n_ci)
            for (int jj = lane; jj < CW; jj += 32)
              sbias[ci * CW + jj] = ep.bias ? ep.bias_scale * __bfloat162float(ep.bias[c0 + jj]) : 0.f;
        }
      }
      __syncwarp();
      const bool gate_uniform = FIXED_U && has_gate && row_gate[0] == row_gate[31];
      // residual / gate registers of one chunk (lane: rows rl = idx / UPR, unit u = idx % UPR, idx = lane + 32 it)
      uint4 rres[NIT];
      float4 rg0, rg1;
      auto prefetch = [&](int ci) {
        const int c0 = n0 + (half + 2 * ci) * CW;
        if (has_res) {
#pragma unroll
          for (int it = 0; it < NIT; ++it) {
            const int idx = lane + 32 * it, rl = idx / UPR, u = idx % UPR;
            if (r0 + rl < p.M) rres[it] = *reinterpret_cast<const uint4*>(ep.res + row_res[rl] + c0 + u * 8);
          }
        }
        if (gate_uniform) {
          const float* g = ep.gate_mod + row_gate[0] + c0 + (lane % UPR) * 8;
          rg0 = *reinterpret_cast<const float4*>(g);
          rg1 = *reinterpret_cast<const float4*>(g + 4);
        }
      };
      prefetch(0);
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * ACC_STRIDE + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int ci = 0; ci < n_ci; ++ci) {
        const int ch = half + 2 * ci;
        const int c0 = n0 + ch * CW;
        uint32_t raw[CW];
        tmem_ld_x32(t_row + ch * CW, raw);
        if constexpr (CW >= 64) tmem_ld_x32(t_row + ch * CW + 32, raw + 32);
        if constexpr (CW == 72) tmem_ld_x8(t_row + ch * CW + 64, raw + 64);
        if constexpr (HD != 0) {
          // ---- the qkv flavour: a chunk is one head.  Its own lean path: the general one below kept 18 row addresses
          // live across the chunk loop and spilled them, and with the shared-memory carve-out at its maximum the L1 that
          // backs local memory is a few KB — every reload was an L2 round trip, 28 % of the kernel's stall samples,
          // and the epilogue warps, not the tensor pipe, set the pace of this flavour (ncu, profiles/) ----
          tmem_ld_wait();
          if (ci == n_ci - 1) {                        // accumulator stage drained: the MMA warp may reuse it
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&tempty[acc]);
          }
          const float* sb = sbias + ci * CW;
          float ss = 0.f;
#pragma unroll
          for (int jj = 0; jj < CW; jj += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(sb + jj);
            const float x0 = __uint_as_float(raw[jj + 0]) + b4.x, x1 = __uint_as_float(raw[jj + 1]) + b4.y;
            const float x2 = __uint_as_float(raw[jj + 2]) + b4.z, x3 = __uint_as_float(raw[jj + 3]) + b4.w;
            ss += x0 * x0;
            ss += x1 * x1;
            ss += x2 * x2;
            ss += x3 * x3;
            raw[jj + 0] = __float_as_uint(x0);
            raw[jj + 1] = __float_as_uint(x1);
            raw[jj + 2] = __float_as_uint(x2);
            raw[jj + 3] = __float_as_uint(x3);
          }
          const int wsel = c0 < p.n_q ? 0 : (c0 < p.n_q + p.n_k ? 1 : 2);
          uint8_t* srow = stg + lane * C::PITCH;
          if (wsel < 2) {
            // LlamaRMSNorm: weight * (x * rsqrt(mean x^2 + eps)).to(bf16), the product rounded to bf16 once (the exact
            // product of two bf16 fits fp32, so one HMUL2.BF16 gives what fp32 multiply + round gave)
            const float r = rsqrtf(ss / CW + p.eps);
            const uint32_t* w2 = s_rms + wsel * (HD / 2);
#pragma unroll
            for (int jj = 0; jj < CW; jj += 8) {
              const uint4 wv = *reinterpret_cast<const uint4*>(w2 + jj / 2);
              const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 hv = __floats2bfloat162_rn(__uint_as_float(raw[jj + 2 * e]) * r,
                                                          __uint_as_float(raw[jj + 2 * e + 1]) * r);
                hv = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&ww[e]), hv);
                o[e] = *reinterpret_cast<uint32_t*>(&hv);
              }
              *reinterpret_cast<uint4*>(srow + jj * 2) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          } else {
#pragma unroll
            for (int jj = 0; jj < CW; jj += 8) {
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 hv = __floats2bfloat162_rn(__uint_as_float(raw[jj + 2 * e]), __uint_as_float(raw[jj + 2 * e + 1]));
                o[e] = *reinterpret_cast<uint32_t*>(&hv);
              }
              *reinterpret_cast<uint4*>(srow + jj * 2) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          }
          __syncwarp();
          // column phase: RPI whole rows per turn (lane = row in turn x 16-byte unit), one running pointer per lane
          constexpr int RPI = 32 / UPR, TURNS = (32 + RPI - 1) / RPI;
          const int rl0 = lane / UPR, u = lane - rl0 * UPR;
          if (lane < RPI * UPR) {
            const bool scat = p.sc.n_dst > 0;
            bf16* base;
            if (scat) {                                // this chunk (= one head) has one destination
              const int part = c0 / p.sc.part_w, cc = c0 - part * p.sc.part_w, g = cc / p.sc.grp;
              base = p.sc.dst[g] + part * p.sc.grp + (cc - g * p.sc.grp) + u * 8;
            } else {
              base = p.out + (r0 + rl0) * p.ldo + c0 + u * 8;
            }
#pragma unroll 1
            for (int t = 0; t < TURNS; ++t) {
              const int rl = t * RPI + rl0;
              if (rl < 32 && r0 + rl < p.M) {
                const uint4 val = *reinterpret_cast<const uint4*>(stg + rl * C::PITCH + u * 16);
                bf16* optr = scat ? base + row_res[rl] : base + (int64_t)t * RPI * p.ldo;
                *reinterpret_cast<uint4*>(optr) = val;
              }
            }
          }
          __syncwarp();
          continue;
        }
        // this chunk's residual / gate registers; the next chunk's loads go out before the TMEM data is consumed
        uint4 cres[NIT];
#pragma unroll
        for (int it = 0; it < NIT; ++it) cres[it] = rres[it];
        const float4 cg0 = rg0, cg1 = rg1;
        if (ci + 1 < n_ci) prefetch(ci + 1);
        tmem_ld_wait();
        if (ci == n_ci - 1) {                        // accumulator stage drained: the MMA warp may reuse it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(&tempty[acc]);
        }
        float v[CW];
        const float* sb = sbias + ci * CW;
#pragma unroll
        for (int jj = 0; jj < CW; jj += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(sb + jj);
          v[jj + 0] = __uint_as_float(raw[jj + 0]) + b4.x;
          v[jj + 1] = __uint_as_float(raw[jj + 1]) + b4.y;
          v[jj + 2] = __uint_as_float(raw[jj + 2]) + b4.z;
          v[jj + 3] = __uint_as_float(raw[jj + 3]) + b4.w;
        }
        if (ep.act == 1) {
#pragma unroll
          for (int jj = 0; jj < CW; ++jj) v[jj] = gelu_tanh_fast(v[jj]);
        }
        // row phase -> staging (bf16, like the reference rounds a projection before gate/residual)
        uint8_t* srow = stg + lane * C::PITCH;
#pragma unroll
        for (int jj = 0; jj < CW; jj += 8) {
          Vec16<bf16> pk;
          pk.pack(v + jj);
          *reinterpret_cast<uint4*>(srow + jj * 2) = pk.raw;
        }
        __syncwarp();
        // column phase: coalesced 16-byte accesses: gate / residual / store.  The staged vectors of all turns are read
        // first (the row guard kept the compiler from hoisting them: four exposed LDS latencies per chunk)
        uint4 spk[NIT];
#pragma unroll
        for (int it = 0; it < NIT; ++it) {
          const int idx = lane + 32 * it;
          spk[it] = *reinterpret_cast<const uint4*>(stg + (idx / UPR) * C::PITCH + (idx % UPR) * 16);
        }
#pragma unroll
        for (int it = 0; it < NIT; ++it) {
          const int idx = lane + 32 * it, rl = idx / UPR, u = idx % UPR;
          const int64_t gr = r0 + rl;
          if (gr < p.M) {
            Vec16<bf16> pk;
            pk.raw = spk[it];
            const int col = c0 + u * 8;
            if (has_gate || has_res) {
              float f[8];
              pk.unpack(f);
              if (has_gate) {
                float4 g0 = cg0, g1 = cg1;
                if (!gate_uniform) {
                  const int64_t goff = row_gate[rl];
                  g0 = *reinterpret_cast<const float4*>(ep.gate_mod + goff + col);
                  g1 = *reinterpret_cast<const float4*>(ep.gate_mod + goff + col + 4);
                }
                f[0] *= g0.x; f[1] *= g0.y; f[2] *= g0.z; f[3] *= g0.w;
                f[4] *= g1.x; f[5] *= g1.y; f[6] *= g1.z; f[7] *= g1.w;
              }
              if (has_res) {
                Vec16<bf16> rv;
                rv.raw = cres[it];
                float rf[8];
                rv.unpack(rf);
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] += rf[e];
              }
              pk.pack(f);
            }
            *reinterpret_cast<uint4*>(p.out + gr * p.ldo + col) = pk.raw;
          }
        }
        __syncwarp();
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  cluster_sync();                                        // the peer's smem / TMEM stay valid until both are done
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, TMEM_COLS);
  }
}

// ---- host side: tensor maps ---------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
  });
  return fn;
}
}  // namespace

// 2-D bf16 row-major [rows, cols] with leading dimension ld (elements); box = box_rows x 64 cols, SW128
int mdv2_make_tmap_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                      uint32_t box_cols) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    mdv2_set_error("cuTensorMapEncodeTiled unavailable");
    return MDV2_ECUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdv2_set_error("cuTensorMapEncodeTiled failed (%d) ptr=%p rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r, ptr,
                   (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
    return MDV2_ECUDA;
  }
  return MDV2_OK;
}

// 3-D bf16 view (d, head, row) of a row-major activation whose heads are D-wide column groups:
// box = 64 x 1 x box_rows, SW128; columns beyond D are zero-filled by the TMA unit.
int mdv2_make_tmap_3d_heads(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    mdv2_set_error("cuTensorMapEncodeTiled unavailable");
    return MDV2_ECUDA;
  }
  cuuint64_t dims[3] = {D, H, rows};
  cuuint64_t strides[2] = {(cuuint64_t)D * 2, ld * 2};
  cuuint32_t box[3] = {64, 1, box_rows};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdv2_set_error("cuTensorMapEncodeTiled(3d) failed (%d) ptr=%p rows=%llu H=%u D=%u ld=%llu", (int)r, ptr,
                   (unsigned long long)rows, H, D, (unsigned long long)ld);
    return MDV2_ECUDA;
  }
  return MDV2_OK;
}

// The 16-wide tail chunk (columns 64 .. 79, those past D zero-filled) of the same 3-D view as 32-byte rows, SW32.
int mdv2_make_tmap_3d_heads_sw32(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t H, uint32_t D, uint64_t ld,
                                 uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    mdv2_set_error("cuTensorMapEncodeTiled unavailable");
    return MDV2_ECUDA;
  }
  cuuint64_t dims[3] = {D, H, rows};
  cuuint64_t strides[2] = {(cuuint64_t)D * 2, ld * 2};
  cuuint32_t box[3] = {16, 1, box_rows};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdv2_set_error("cuTensorMapEncodeTiled(3d sw32) failed (%d) ptr=%p rows=%llu H=%u D=%u ld=%llu", (int)r, ptr,
                   (unsigned long long)rows, H, D, (unsigned long long)ld);
    return MDV2_ECUDA;
  }
  return MDV2_OK;
}

// 4-D bf16 view (d, head, row in sequence, sequence) of a row-major [n_seq·L, ld] activation, for TMA STORES of
// box_rows output rows of one head: rows past the end of a sequence are clipped by the TMA unit.  No swizzle unless the
// head is exactly 128 bytes wide (D = 64), where the staged rows are SW128 so that row-per-thread writes spread over
// the banks.
int mdv2_make_tmap_4d_store(CUtensorMap* out, void* ptr, uint64_t n_seq, uint64_t L, uint32_t H, uint32_t D, uint64_t ld,
                            uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    mdv2_set_error("cuTensorMapEncodeTiled unavailable");
    return MDV2_ECUDA;
  }
  cuuint64_t dims[4] = {D, H, L, n_seq};
  cuuint64_t strides[3] = {(cuuint64_t)D * 2, ld * 2, L * ld * 2};
  cuuint32_t box[4] = {D, 1, box_rows, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   D == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdv2_set_error("cuTensorMapEncodeTiled(4d store) failed (%d) ptr=%p n_seq=%llu L=%llu H=%u D=%u ld=%llu", (int)r, ptr,
                   (unsigned long long)n_seq, (unsigned long long)L, H, D, (unsigned long long)ld);
    return MDV2_ECUDA;
  }
  return MDV2_OK;
}

namespace {
template <int BN, int HD, int MC>
int launch_mc(const CUtensorMap& ta, const CUtensorMap& tw, const CUtensorMap& twr, const Params& p, cudaStream_t st) {
  using C = Cfg<BN, HD>;
  auto kern = gemm_sm100_kernel<BN, HD, MC>;
  static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();
  static int max_clusters[64];
  int dev = 0;
  cudaGetDevice(&dev);
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2 * MC;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) {
      mdv2_set_error("gemm_sm100: cannot set %d B dynamic smem: %s", C::SMEM, cudaGetErrorString(e));
      return MDV2_ECUDA;
    }
    int n = mdv2_num_sms() / (2 * MC);
    if (MC > 1) {                                      // clusters of four do not tile every GPC: ask the driver
      cfg.gridDim = dim3(2 * MC * n);
      int got = 0;
      if (cudaOccupancyMaxActiveClusters(&got, kern, &cfg) == cudaSuccess && got > 0) n = got;
      else cudaGetLastError();
    }
    max_clusters[dev & 63] = n;
    configured = true;
  }
  const int64_t tiles = ((p.M + 2 * BM - 1) / (2 * BM)) * (p.N / BN / MC + ((MC == 1 && p.N % BN) ? 1 : 0));
  int64_t clusters = max_clusters[dev & 63];
  static const int forced_pairs = [] {                 // tuning hook (tools/gemm_pairs_probe.py)
    const char* e = getenv("MDV2_GEMM_PAIRS");
    return e ? atoi(e) : 0;
  }();
  if (forced_pairs > 0 && forced_pairs / MC < clusters) clusters = forced_pairs / MC > 0 ? forced_pairs / MC : 1;
  cfg.gridDim = dim3((unsigned)(2 * MC * (tiles < clusters ? tiles : clusters)));
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tw, twr, p);
  if (e != cudaSuccess) {
    mdv2_set_error("gemm_sm100: launch failed: %s", cudaGetErrorString(e));
    return MDV2_ECUDA;
  }
  MDV2_CHECK_LAUNCH("gemm_sm100");
  // one name per tile flavour: bench.py lists every instantiation as its own kernel (they are: the launch list shows
  // three gemm_sm100_kernel<BN, HD> entries), so that the dominant KERNEL of the step is named, not a family
  static const std::string flavour = "gemm_sm100<" + std::to_string(BN) + "," + std::to_string(HD) + ">";
  mdv2_note_kernel(flavour.c_str());
  return MDV2_OK;
}

// pairs per cluster: one.  MDV2_GEMM_MC=2 selects the A-multicast clusters of four where the N tiles pair up — measured
// (profiles/yardstick_gemm_r02.txt): +7 % per SM, but the driver places only 33 such clusters (132 of 148 SMs: GPCs of
// 18 SMs do not tile by four), so every flavour is 3-5 % slower overall (the qkv flavour 22 %)
template <int BN, int HD>
int launch(const CUtensorMap& ta, const CUtensorMap& ta_mc, const CUtensorMap& tw, const CUtensorMap& twr, const Params& p,
           cudaStream_t st) {
  static const int mc = [] {
    const char* e = getenv("MDV2_GEMM_MC");
    return e ? atoi(e) : 1;
  }();
  if (mc == 2 && p.N % BN == 0 && (p.N / BN) % 2 == 0) return launch_mc<BN, HD, 2>(ta_mc, tw, tw, p, st);
  return launch_mc<BN, HD, 1>(ta, tw, twr, p, st);
}

int pick_bn(int N, int K, const RmsArgs* rms) {
  if (rms && rms->wq) {
    static const int heads3 = [] {                     // MDV2_GEMM_QKV216=0: two heads per tile (BN = 144)
      const char* e = getenv("MDV2_GEMM_QKV216");
      return e ? atoi(e) : 1;
    }();
    if (rms->head_dim == 72 && heads3 && N % 216 == 0) return 216;
    if (rms->head_dim == 72 && N % 144 == 0) return 144;
    if (rms->head_dim == 64 && N % 128 == 0) return 128;
    return 0;
  }
  static const int forced = [] {                       // tuning hook for tools/bench_gemm.py
    const char* e = getenv("MDV2_GEMM_BN");
    return e ? atoi(e) : 0;
  }();
  if ((forced == 256 || forced == 192 || forced == 128 || forced == 64) && N % forced == 0) return forced;
  if (N % 256 == 0) return 256;
  static const int tail = [] {                         // MDV2_GEMM_TAIL=0: no narrow last tile (N = 1152 as 6 x 192)
    const char* e = getenv("MDV2_GEMM_TAIL");
    return e ? atoi(e) : 1;
  }();
  // long-K GEMMs whose N leaves a narrow last tile (fc2: N = 1152, K = 4608) take six equal 192-wide tiles: a tile
  // lasts 36 us there, the pair that draws the half-width tile runs 18 us ahead of the four that share its A row block,
  // and A left L2 in between — 8.0 GB read from DRAM for 3.9 algorithmic (ncu), against 1.4x with equal tiles; after the
  // issue-chain fix the 192-wide tile is no slower there (1430 vs 1406 TF/s).  At K = 1152 the drift stays inside the L2
  // window (1.59 of 1.55 GB) and 256-wide tiles + one narrow tile stay ahead
  if (tail && N > 256 && N % 256 == 128 && N % 192 == 0 && K >= 2304) return 192;
  if (tail && N > 256 && N % 256 == 128) return 256;   // 256-wide tiles + one 128-wide last tile
  if (N % 192 == 0) return 192;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  return 0;
}
}  // namespace

bool mdv2_gemm_sm100_supported(int64_t M, int N, int K, int64_t lda, int64_t ldw, int64_t ldo, const RmsArgs* rms) {
  if (M <= 0 || K % BK != 0 || K <= 0) return false;
  if (lda % 8 != 0 || ldw % 8 != 0 || ldo % 8 != 0 || N % 8 != 0) return false;
  if (rms && rms->wq && (rms->n_q % rms->head_dim != 0 || rms->n_k % rms->head_dim != 0)) return false;
  return pick_bn(N, K, rms) != 0;
}

int mdv2_gemm_sm100(const void* a, int64_t lda, const void* w, int64_t ldw, void* out, int64_t ldo, int64_t M, int N,
                    int K, const Epilogue<bf16>& ep, const RmsArgs* rms, cudaStream_t st, const QkvScatter* scatter) {
  MDV2_REQUIRE(mdv2_gemm_sm100_supported(M, N, K, lda, ldw, ldo, rms), "gemm_sm100: unsupported shape M=%lld N=%d K=%d",
               (long long)M, N, K);
  MDV2_REQUIRE(((uintptr_t)a % 16 == 0) && ((uintptr_t)w % 16 == 0) && ((uintptr_t)out % 16 == 0) &&
                   (!ep.res || ((uintptr_t)ep.res % 16 == 0 && ep.ldr % 8 == 0)),
               "gemm_sm100: operands must be 16-byte aligned");
  const int bn = pick_bn(N, K, rms);
  CUtensorMap ta, ta_mc, tw;
  int rc = mdv2_make_tmap_2d(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = mdv2_make_tmap_2d(&ta_mc, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM / 2, BK);     // 64-row boxes (multicast)
  if (rc) return rc;
  rc = mdv2_make_tmap_2d(&tw, w, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)(bn == 216 ? 224 : bn) / 2, BK);
  if (rc) return rc;
  CUtensorMap twr = tw;                                  // the narrow last column tile (N % bn columns): its own box height
  if (N % bn) {
    rc = mdv2_make_tmap_2d(&twr, w, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)(N % bn) / 2, BK);
    if (rc) return rc;
  }
  Params p;
  p.sc.n_dst = 0;
  if (scatter != nullptr) {
    const int hd = rms ? rms->head_dim : 0;
    MDV2_REQUIRE(rms && rms->wq && (bn == 216 || bn == 144 || bn == 128) && N % 3 == 0, "gemm_sm100: scattered output needs the RMSNorm qkv flavour");
    const int W = N / 3;
    MDV2_REQUIRE(scatter->n_dst >= 1 && scatter->n_dst <= 8 && W % scatter->n_dst == 0 && (W / scatter->n_dst) % hd == 0 &&
                     rms->n_q == W && rms->n_k == W && scatter->rows_per_seq > 0 && M % scatter->rows_per_seq == 0 &&
                     scatter->rank >= 0 && scatter->rank < scatter->n_dst,
                 "gemm_sm100: bad scatter geometry");
    p.sc.n_dst = scatter->n_dst;
    p.sc.grp = W / scatter->n_dst;
    p.sc.part_w = W;
    p.sc.rows_per_seq = scatter->rows_per_seq;
    p.sc.dst_ld = 3 * (int64_t)p.sc.grp;
    p.sc.seq_rows = (int64_t)scatter->n_dst * scatter->rows_per_seq;
    p.sc.row_base = (int64_t)scatter->rank * scatter->rows_per_seq;
    for (int i = 0; i < 8; ++i) p.sc.dst[i] = i < scatter->n_dst ? (bf16*)scatter->dst[i] : nullptr;
    for (int i = 0; i < scatter->n_dst; ++i)
      MDV2_REQUIRE(scatter->dst[i] != nullptr && (uintptr_t)scatter->dst[i] % 16 == 0, "gemm_sm100: bad scatter destination");
  }
  p.M = M; p.N = N; p.K = K; p.ldo = ldo; p.out = (bf16*)out; p.ep = ep;
  p.rms_q = rms ? (const bf16*)rms->wq : nullptr;
  p.rms_k = rms ? (const bf16*)rms->wk : nullptr;
  p.n_q = rms ? rms->n_q : 0;
  p.n_k = rms ? rms->n_k : 0;
  p.eps = rms ? rms->eps : 0.f;
  switch (bn) {
    case 256: return launch<256, 0>(ta, ta_mc, tw, twr, p, st);
    case 192: return launch<192, 0>(ta, ta_mc, tw, twr, p, st);
    case 64: return launch<64, 0>(ta, ta_mc, tw, twr, p, st);
    case 216: return launch<216, 72>(ta, ta_mc, tw, twr, p, st);
    case 144: return launch<144, 72>(ta, ta_mc, tw, twr, p, st);
    case 128:
      if (rms && rms->wq) return launch<128, 64>(ta, ta_mc, tw, twr, p, st);
      return launch<128, 0>(ta, ta_mc, tw, twr, p, st);
  }
  MDV2_REQUIRE(false, "gemm_sm100: no tile for N=%d", N);
}
