/* mdv2.h — C ABI of libmdv2.so, the sm_100a kernels behind the MagicDriveSTDiT3 drop-in.
 *
 * The reference (HerrWangkv/MagicDrive-V2) is pure PyTorch and has no FFI of its own; each entry
 * point below replaces the library call(s) the reference makes at the cited lines (paths relative
 * to the reference root; "stdit3.py" = magicdrivedit/models/magicdrive/magicdrive_stdit3.py,
 * "blocks.py" = magicdrivedit/models/layers/blocks.py).
 *
 * Conventions: plain pointers and sizes only; every pointer is a DEVICE pointer unless noted;
 * `dtype`: 0 = fp32, 1 = bf16 (activations, weights and outputs share it; modulation tables and
 * RoPE tables are always fp32).  `stream` is a cudaStream_t passed as void*.  Functions launch
 * asynchronously on `stream`, never synchronise, never allocate persistent device memory, and
 * return 0 on success or a negative MDV2_E* code (message via mdv2_last_error()).
 *
 * Token geometry ("geo"): rows of an activation matrix are ordered (β, t, s) with β = bi·NC + v;
 * row r -> modulation row (r / S / T / NC)·T + (r / S) % T of a table [b·T, K, C] (fp32).
 */
#ifndef MDV2_H
#define MDV2_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MDV2_OK 0
#define MDV2_EINVAL (-1)   /* unsupported shape / argument */
#define MDV2_ECUDA (-2)    /* CUDA runtime / driver error at launch */
#define MDV2_EARCH (-3)    /* device is not sm_100 */

#define MDV2_BACKEND_AUTO 0
#define MDV2_BACKEND_SIMT 1     /* CUDA-core kernels (fp32 validation mode, odd shapes) */
#define MDV2_BACKEND_TCGEN05 2  /* tcgen05 / TMEM / TMA kernels (bf16) */

int mdv2_version(void);
const char* mdv2_last_error(void);   /* host pointer, thread-local */
const char* mdv2_last_kernel(void);  /* kernel family launched by the last successful call on this thread */

/* LayerNorm(eps, no affine) -> x·(1+scale)+shift with per-(bi,t) rows of `mod`.
 * Replaces nn.LayerNorm/apex FusedLayerNorm + t2i_modulate + t_mask_select:
 * blocks.py:71-80,92-93; stdit3.py:148-156,213-226,288-301,337-340; blocks.py:893-903. */
int mdv2_ln_modulate(const void* x, void* out, const float* mod, int64_t n_rows, int C, int K,
                     int k_shift, int k_scale, int T, int S, int NC, float eps, int dtype,
                     void* stream);

/* adaLN modulation tables of all blocks in one pass: out[n][bt][k][c] (fp32) = round_dtype(tables[n][k][c] +
 * tsel[bt][k][c or 0]); tables = the blocks' stacked scale_shift_table parameters [n, K, C], tsel = the x_mask-selected
 * timestep vector of every (b, t) frame, [BT, K, Ct] with Ct = C or 1 (the scalar modulation of the cross-view branch).
 * Replaces the per-block `self.scale_shift_table[None] + t.reshape(B, 6, -1)` (+ the t0 copy and t_mask_select) of
 * MultiViewSTDiT3Block.forward, stdit3.py:213-221,288-296, and `scale_shift_table_mva[None] + t_mlp[:, :3]`, :303-305. */
int mdv2_adaln_tables(const void* tables, const void* tsel, float* out, int n, int BT, int K, int C, int Ct, int dtype,
                      void* stream);

/* out[M,N] = epilogue(a[M,K] · w[N,K]^T):  v = acc + bias_scale·bias ; optional per-head RMSNorm
 * (columns [0,n_q) with rms_q, [n_q,n_q+n_k) with rms_k, head_dim wide) ; optional tanh-GELU ;
 * optional gate (fp32 table row, index k_gate) ; optional residual (row r, or mapped row
 * (r / (res_mod·res_div))·res_mod + r % res_mod, or r % res_mod when res_div == 0).
 * out may alias res.  Replaces every nn.Linear / F.linear / timm Mlp / LlamaRMSNorm / gate+residual
 * on the path: blocks.py:54-68,183,188,410-456,658-661; stdit3.py:256-262,280,325-356,1145-1157. */
int mdv2_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias, void* out,
              int64_t ldo, int64_t M, int N, int K, int act, float bias_scale,
              const float* gate_mod, int mod_K, int k_gate, const void* res, int64_t ldr,
              int64_t res_mod, int64_t res_div, int T, int S, int NC, const void* rms_q,
              const void* rms_k, int head_dim, int n_q, int n_k, float eps, int dtype, int backend,
              void* stream);

/* Softmax attention over n_seq sequences: query rows [i·Lq,(i+1)·Lq) of q, key/value rows
 * [j·Lk,(j+1)·Lk) with j = i (kv_map == NULL) or j = kv_map[i·n_sets + set]; the outputs of the
 * n_sets K/V sets are summed (cross-view attention).  Heads are D-wide column groups.
 * Replaces flash_attn_func / xformers / the eager softmax path and the cross-view gather+sum:
 * blocks.py:456-504,607-613,710-742; stdit3.py:158-190,303-322. */
int mdv2_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                   int64_t ldv, void* out, int64_t ldo, int n_seq, int Lq, int Lk, int H, int D,
                   float scale, const int32_t* kv_map, int n_sets, int dtype, int backend,
                   void* stream);

/* Temporal self-attention of every (β, s) column over its T frames, with interleaved-pair RoPE on
 * q and k (cos/sin: fp32 [T, D]); qkv rows are (β,t,s) with q|k|v column blocks of width H·D.
 * Replaces the two transposes + Attention.forward + RotaryEmbedding: stdit3.py:232-234,
 * blocks.py:198-278. */
int mdv2_temporal_attention(const void* qkv, int64_t ld, const float* cos_t, const float* sin_t,
                            void* out, int64_t ldo, int B, int T, int S, int H, int D, float scale,
                            int dtype, void* stream);

/* Patch gather for PatchEmbed3D with patch (1,2,2): x[Bx, Cin·NCi, T, Hx, Wx] (channel = c·NCi+v)
 * -> out rows (β = bx·NCi+v, t, s in [s0, s0+Sl)) of Cin·4 values ordered (c, dh, dw); zero beyond
 * Hx/Wx (the H/W padding of blocks.py:132-147 and stdit3.py:1101-1104; view split stdit3.py:1041). */
int mdv2_patchify(const void* x, void* out, int Bx, int Cin, int NCi, int T, int Hx, int Wx, int Hp,
                  int Wp, int s0, int Sl, int dtype, void* stream);

/* unpatchify + crop + fp32 cast + view re-interleave: y[(b·NC+v), t, Hp·Wp, (dh, dw, c)] ->
 * out fp32 [b, c·NC+v, Tx, Hx, Wx].  stdit3.py:1212-1244. */
int mdv2_unpatchify(const void* y, float* out, int b, int NC, int Cout, int T, int Hp, int Wp, int Tx,
                    int Hx, int Wx, int dtype, void* stream);

/* dst[b][a][:] = src[a][b][:] for a contiguous inner run of inner_bytes (multiple of 16): the pack / unpack
 * re-orderings around the sequence-parallel all-to-alls.  Replaces the tensor_split + contiguous + cat copies of
 * magicdrivedit/acceleration/communications.py:11-21 and the head/sequence rearranges of blocks.py:571-638. */
int mdv2_swap01(const void* src, void* dst, int64_t A, int64_t B, int64_t inner_bytes, void* stream);

/* Strided scatter of contiguous runs to up to 8 destination buffers (the ranks of a sequence-parallel group,
 * mapped through NVLink peer memory): for every index (i0,i1,i2,i3) < sizes,
 *   dst_ptrs[i_{peer_dim}] [dst_base + sum_{k != peer_dim} i_k·dst_strides[k] : +run_bytes]
 *     = src[sum_k i_k·src_strides[k] : +run_bytes]            (strides and dst_base in bytes, all multiples of 16).
 * One launch packs, transfers and un-packs a Ulysses all-to-all: it replaces the two all_to_all calls plus the
 * split / cat / rearrange copies of magicdrivedit/acceleration/communications.py:11-66 and blocks.py:571-638.
 * The caller orders it against the peers (a group barrier after the launch, before the data is consumed). */
int mdv2_scatter_runs(const void* src, void* const* dst_ptrs, int n_dst, int peer_dim, const int64_t* sizes,
                      const int64_t* src_strides, const int64_t* dst_strides, int64_t run_bytes, int64_t dst_base,
                      void* stream);

/* qkv projection + per-head RMSNorm of q and k (as mdv2_gemm with rms_q / rms_k) whose epilogue stores head group g of
 * each of q | k | v (N = 3·W columns, groups of W / n_dst columns) straight into dst_ptrs[g] — the ranks of a
 * sequence-parallel group, mapped through NVLink peer memory — in the layout mdv2_attention reads there:
 *   dst_ptrs[g] [seq][rank·rows_per_seq + r][q|k|v][W / n_dst],   seq = row / rows_per_seq, r = row % rows_per_seq.
 * The GEMM and the head-scatter / sequence-gather all-to-all are one kernel: the NVLink stores ride under the MMAs.
 * Replaces qkv Linear + q_norm / k_norm + all_to_all (+ split / cat / rearrange) of SeqParallelMultiHeadAttention,
 * magicdrivedit/models/layers/blocks.py:554-600 and acceleration/communications.py:11-66.  bf16, sm_100a only; the
 * caller orders it against the peers (group barrier behind the launch). */
int mdv2_gemm_qkv_scatter(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias, int64_t M, int N,
                          int K, const void* rms_q, const void* rms_k, int head_dim, float eps,
                          void* const* dst_ptrs, int n_dst, int rows_per_seq, int rank, void* stream);

/* mdv2_attention whose epilogue stores the final (normalised, summed over K/V sets) output row s of sequence i straight
 * into the rank that owns it under sequence parallelism:  dst_ptrs[s / rows_per_rank] [(i·rows_per_rank +
 * s % rows_per_rank)·dst_ld + dst_col + h·D : + D]  (Lq = n_dst·rows_per_rank; dst_col = this rank's head-group column).
 * `partial` [n_seq·Lq, ldp] only carries the running sum of earlier K/V sets (may be NULL when n_sets == 1).
 * Attention and the sequence-scatter / head-gather all-to-all are one kernel.  Replaces flash_attn_func + all_to_all
 * (+ rearranges) of SeqParallelMultiHeadAttention, magicdrivedit/models/layers/blocks.py:607-638 and
 * acceleration/communications.py:11-66.  bf16, sm_100a only; the caller orders it against the peers. */
int mdv2_attention_scatter(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                           void* partial, int64_t ldp, int n_seq, int Lq, int Lk, int H, int D, float scale,
                           const int32_t* kv_map, int n_sets, void* const* dst_ptrs, int n_dst, int rows_per_rank,
                           int64_t dst_ld, int dst_col, void* stream);

/* mdv2_attention_scatter for callers that cannot allocate (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The running sum of the earlier K/V sets (`partial` of mdv2_attention_scatter, [n_seq·Lq, H·D] bf16) lives in a
 * file-scope device array of MDV2_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit; the
 * caller still passes inputs and destinations.  Returns MDV2_EINVAL when n_seq·Lq·H·D·2 bytes exceed the array.
 * Calls no cudaMalloc / cudaFree / allocating routine.  All calls share the array (it keeps the last call's
 * contents; the kernel overwrites what it reads, nothing needs clearing): calls that overlap on the GPU — two
 * streams, or two threads — race on it; serialise them on one stream.  bf16, sm_100a only. */
#define MDV2_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
int64_t mdv2_static_scratch_bytes(void);
int mdv2_attention_scatter_static(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                  int n_seq, int Lq, int Lk, int H, int D, float scale, const int32_t* kv_map,
                                  int n_sets, void* const* dst_ptrs, int n_dst, int rows_per_rank, int64_t dst_ld,
                                  int dst_col, void* stream);

/* Tail of one RFLOW sampling step (magicdrivedit/schedulers/rf/__init__.py:100-109) plus the head of the next
 * (:95 and the x.to(dtype) of magicdrive_stdit3.py:1041), fused into one pass:
 *   v = pred[n:] + guidance * (pred[:n] - pred[n:]);  z_out = z + v * dt[b]        (fp32, each op rounded separately)
 *   if keep: z_out = keep[b, t] ? z_out : x0           (frame mask `mask_t_upper` [n, T] as bytes; x0 has z's dtype)
 *   if z_in:  z_in[0:n] = z_in[n:2n] = cast(z_out)     (zin_dtype 0 = fp32, 1 = bf16: the next forward's input)
 * pred is the model's fp32 output [2n, C, T, hw]; z is [n, C, T, hw] of z_dtype (0 fp32, 1 bf16); dt is a device
 * vector of n floats (already divided by num_timesteps).  z_out may alias z when z is fp32. */
int mdv2_cfg_euler_step(const float* pred, const void* z, int z_dtype, const float* dt, float guidance,
                        const void* x0, const uint8_t* keep, float* z_out, void* z_in, int zin_dtype, int n, int C,
                        int T, int64_t hw, void* stream);

#ifdef __cplusplus
}
#endif
#endif
