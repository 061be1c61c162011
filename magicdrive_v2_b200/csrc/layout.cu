// layout.cu — patch gather / scatter and the temporal-attention micro-kernel.
#include <cstdlib>

#include "common.cuh"
#include "sm100_ptx.cuh"

// ------------------------------------------------------------------------------------------------
// patchify: one thread per output element (coalesced stores)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void patchify_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t total, int Cin,
                                int NCi, int Tn, int Hx, int Wx, int Wp, int s0, int Sl) {
  const int K4 = Cin * 4;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(idx % K4);
    int64_t row = idx / K4;
    const int sl = (int)(row % Sl);
    row /= Sl;
    const int t = (int)(row % Tn);
    const int beta = (int)(row / Tn);
    const int bx = beta / NCi, v = beta % NCi;
    const int c = e >> 2, dh = (e >> 1) & 1, dw = e & 1;
    const int s = s0 + sl;
    const int h = 2 * (s / Wp) + dh, w = 2 * (s % Wp) + dw;
    T val = from_f<T>(0.f);
    if (h < Hx && w < Wx) {
      int64_t ch = (int64_t)bx * Cin * NCi + (int64_t)c * NCi + v;
      val = x[((ch * Tn + t) * Hx + h) * (int64_t)Wx + w];
    }
    out[idx] = val;
  }
}

extern "C" int mdv2_patchify(const void* x, void* out, int Bx, int Cin, int NCi, int T, int Hx, int Wx,
                             int Hp, int Wp, int s0, int Sl, int dtype, void* stream) {
  MDV2_REQUIRE(s0 >= 0 && Sl >= 0 && s0 + Sl <= Hp * Wp, "patchify: token slice out of range");
  MDV2_REQUIRE(2 * Hp >= Hx && 2 * Wp >= Wx, "patchify: patch grid smaller than the input");
  const int64_t total = (int64_t)Bx * NCi * T * Sl * Cin * 4;
  if (total == 0) return MDV2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = (int64_t)mdv2_num_sms() * 32;
  if (blocks > cap) blocks = cap;
  if (dtype == 0)
    patchify_kernel<float><<<(unsigned)blocks, threads, 0, st>>>((const float*)x, (float*)out, total, Cin, NCi, T, Hx, Wx, Wp, s0, Sl);
  else if (dtype == 1)
    patchify_kernel<bf16><<<(unsigned)blocks, threads, 0, st>>>((const bf16*)x, (bf16*)out, total, Cin, NCi, T, Hx, Wx, Wp, s0, Sl);
  else
    MDV2_REQUIRE(false, "patchify: bad dtype");
  MDV2_CHECK_LAUNCH("patchify");
  return MDV2_OK;
}

// ------------------------------------------------------------------------------------------------
// unpatchify + crop + fp32 + view interleave: one thread per output element
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void unpatchify_kernel(const T* __restrict__ y, float* __restrict__ out, int64_t total, int NC,
                                  int Cout, int Tn, int Hp, int Wp, int Tx, int Hx, int Wx) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = idx;
    const int w = (int)(r % Wx); r /= Wx;
    const int h = (int)(r % Hx); r /= Hx;
    const int t = (int)(r % Tx); r /= Tx;
    const int ch = (int)(r % (Cout * NC));
    const int b = (int)(r / (Cout * NC));
    const int c = ch / NC, v = ch % NC;
    const int64_t row = (((int64_t)(b * NC + v) * Tn + t) * Hp + (h >> 1)) * Wp + (w >> 1);
    const int feat = (((h & 1) << 1) | (w & 1)) * Cout + c;
    out[idx] = to_f(y[row * (4 * Cout) + feat]);
  }
}

extern "C" int mdv2_unpatchify(const void* y, float* out, int b, int NC, int Cout, int T, int Hp, int Wp,
                               int Tx, int Hx, int Wx, int dtype, void* stream) {
  MDV2_REQUIRE(Tx <= T && Hx <= 2 * Hp && Wx <= 2 * Wp, "unpatchify: crop larger than the token grid");
  const int64_t total = (int64_t)b * NC * Cout * Tx * Hx * Wx;
  if (total == 0) return MDV2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = (int64_t)mdv2_num_sms() * 32;
  if (blocks > cap) blocks = cap;
  if (dtype == 0)
    unpatchify_kernel<float><<<(unsigned)blocks, threads, 0, st>>>((const float*)y, out, total, NC, Cout, T, Hp, Wp, Tx, Hx, Wx);
  else if (dtype == 1)
    unpatchify_kernel<bf16><<<(unsigned)blocks, threads, 0, st>>>((const bf16*)y, out, total, NC, Cout, T, Hp, Wp, Tx, Hx, Wx);
  else
    MDV2_REQUIRE(false, "unpatchify: bad dtype");
  MDV2_CHECK_LAUNCH("unpatchify");
  return MDV2_OK;
}

// ------------------------------------------------------------------------------------------------
// temporal attention: one warp per (β, s, head); lane owns the RoPE pairs p = lane + 32·j.
// K/V of the T frames stay in registers; strided (β,:,s) access, no transposes materialised.
// ------------------------------------------------------------------------------------------------
template <typename T> struct Pair;
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return *reinterpret_cast<const float2*>(p); }
  static __device__ __forceinline__ void st(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
};
template <> struct Pair<bf16> {
  static __device__ __forceinline__ float2 ld(const bf16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  }
  static __device__ __forceinline__ void st(bf16* p, float2 v) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(v.x, v.y);
  }
};

template <typename T, int TMAX, int PPL>
__global__ void __launch_bounds__(128) temporal_attn_kernel(const T* __restrict__ qkv, int64_t ld,
                                                            const float* __restrict__ cos_t,
                                                            const float* __restrict__ sin_t,
                                                            T* __restrict__ out, int64_t ldo, int64_t n_warps,
                                                            int Tn, int S, int H, int D, float scale) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (wid >= n_warps) return;
  const int h = (int)(wid % H);
  const int64_t bs = wid / H;
  const int s = (int)(bs % S);
  const int64_t beta = bs / S;
  const int C = H * D;
  const int np = D / 2;
  float2 kk[TMAX][PPL], vv[TMAX][PPL];
#pragma unroll
  for (int t = 0; t < TMAX; ++t) {
    if (t < Tn) {
      const T* rowp = qkv + ((beta * Tn + t) * S + s) * ld + (int64_t)h * D;
#pragma unroll
      for (int j = 0; j < PPL; ++j) {
        const int p = lane + 32 * j;
        if (p < np) {
          float2 kv = Pair<T>::ld(rowp + C + 2 * p);
          const float c0 = cos_t[t * D + 2 * p], c1 = cos_t[t * D + 2 * p + 1];
          const float s0 = sin_t[t * D + 2 * p], s1 = sin_t[t * D + 2 * p + 1];
          kk[t][j] = make_float2(kv.x * c0 - kv.y * s0, kv.y * c1 + kv.x * s1);
          vv[t][j] = Pair<T>::ld(rowp + 2 * C + 2 * p);
        } else {
          kk[t][j] = make_float2(0.f, 0.f);
          vv[t][j] = make_float2(0.f, 0.f);
        }
      }
    }
  }
  for (int tq = 0; tq < Tn; ++tq) {
    const T* rowp = qkv + ((beta * Tn + tq) * S + s) * ld + (int64_t)h * D;
    float2 qq[PPL];
#pragma unroll
    for (int j = 0; j < PPL; ++j) {
      const int p = lane + 32 * j;
      if (p < np) {
        float2 qv = Pair<T>::ld(rowp + 2 * p);
        const float c0 = cos_t[tq * D + 2 * p], c1 = cos_t[tq * D + 2 * p + 1];
        const float s0 = sin_t[tq * D + 2 * p], s1 = sin_t[tq * D + 2 * p + 1];
        qq[j] = make_float2((qv.x * c0 - qv.y * s0) * scale, (qv.y * c1 + qv.x * s1) * scale);
      } else {
        qq[j] = make_float2(0.f, 0.f);
      }
    }
    float sc[TMAX];
    float mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < TMAX; ++t) {
      if (t < Tn) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < PPL; ++j) a += qq[j].x * kk[t][j].x + qq[j].y * kk[t][j].y;
        sc[t] = warp_sum(a);
        mx = fmaxf(mx, sc[t]);
      }
    }
    float den = 0.f;
#pragma unroll
    for (int t = 0; t < TMAX; ++t) {
      if (t < Tn) {
        sc[t] = __expf(sc[t] - mx);
        den += sc[t];
      }
    }
    const float inv = 1.f / den;
    T* orow = out + ((beta * Tn + tq) * S + s) * ldo + (int64_t)h * D;
#pragma unroll
    for (int j = 0; j < PPL; ++j) {
      const int p = lane + 32 * j;
      if (p < np) {
        float2 o = make_float2(0.f, 0.f);
#pragma unroll
        for (int t = 0; t < TMAX; ++t) {
          if (t < Tn) {
            o.x += sc[t] * vv[t][j].x;
            o.y += sc[t] * vv[t][j].y;
          }
        }
        Pair<T>::st(orow + 2 * p, make_float2(o.x * inv, o.y * inv));
      }
    }
  }
}


// bf16 fast path (T <= 8, D % 8 == 0, D <= 128): a 16-lane group per (β, s, head), each lane owning one
// 16-byte vector (8 head dims = 4 RoPE pairs) of q/k/v.  Every global access is a 16-byte vector and all
// 3·T loads of a lane are issued before the first use, so each SM keeps tens of KB in flight.
template <int TMAX>
__global__ void __launch_bounds__(256) temporal_attn_vec_kernel(const bf16* __restrict__ qkv, int64_t ld,
                                                                const float* __restrict__ cos_t,
                                                                const float* __restrict__ sin_t,
                                                                bf16* __restrict__ out, int64_t ldo,
                                                                int64_t n_units, int S, int H, int D,
                                                                float scale) {
  constexpr int Tn = TMAX;                               // exact frame count: every t-loop fully unrolls
  const int l = threadIdx.x & 15;
  const int64_t unit = (int64_t)blockIdx.x * 16 + (threadIdx.x >> 4);
  const bool active = unit < n_units && l * 8 < D;
  const int64_t u = unit < n_units ? unit : n_units - 1;
  const int h = (int)(u % H);
  const int64_t bs = u / H;
  const int s = (int)(bs % S);
  const int64_t beta = bs / S;
  const int C = H * D;
  const int d0 = active ? l * 8 : 0;
  const bf16* base = qkv + ((beta * Tn) * S + s) * ld + (int64_t)h * D + d0;
  const int64_t tstride = (int64_t)S * ld;
  uint4 qraw[TMAX], kraw[TMAX], vraw[TMAX];
#pragma unroll
  for (int t = 0; t < TMAX; ++t) {
    if (t < Tn && active) {
      const bf16* rp = base + t * tstride;
      qraw[t] = *reinterpret_cast<const uint4*>(rp);
      kraw[t] = *reinterpret_cast<const uint4*>(rp + C);
      vraw[t] = *reinterpret_cast<const uint4*>(rp + 2 * C);
    } else {
      qraw[t] = kraw[t] = vraw[t] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  auto rope8 = [&](const uint4& raw, int t, float mul, float* o) {
    Vec16<bf16> vv;
    vv.raw = raw;
    float f[8];
    vv.unpack(f);
    const float4 c0 = *reinterpret_cast<const float4*>(cos_t + t * D + d0);
    const float4 c1 = *reinterpret_cast<const float4*>(cos_t + t * D + d0 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(sin_t + t * D + d0);
    const float4 s1 = *reinterpret_cast<const float4*>(sin_t + t * D + d0 + 4);
    const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
      o[i] = (f[i] * cs[i] - f[i + 1] * sn[i]) * mul;
      o[i + 1] = (f[i + 1] * cs[i + 1] + f[i] * sn[i + 1]) * mul;
    }
  };
  float kk[TMAX][8], vv[TMAX][8];
#pragma unroll
  for (int t = 0; t < TMAX; ++t) {
    if (t < Tn) {
      rope8(kraw[t], t, 1.f, kk[t]);
      Vec16<bf16> tv;
      tv.raw = vraw[t];
      tv.unpack(vv[t]);
    }
  }
#pragma unroll
  for (int tq = 0; tq < TMAX; ++tq) {
    if (tq < Tn) {
      float q[8];
      rope8(qraw[tq], tq, scale, q);
      float sc[TMAX];
      float mx = -INFINITY;
#pragma unroll
      for (int t = 0; t < TMAX; ++t) {
        if (t < Tn) {
          float a = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) a += q[i] * kk[t][i];
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
          sc[t] = a;
          mx = fmaxf(mx, a);
        }
      }
      float den = 0.f;
#pragma unroll
      for (int t = 0; t < TMAX; ++t) {
        if (t < Tn) {
          sc[t] = __expf(sc[t] - mx);
          den += sc[t];
        }
      }
      const float inv = 1.f / den;
      float o8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int t = 0; t < TMAX; ++t) {
        if (t < Tn) {
#pragma unroll
          for (int i = 0; i < 8; ++i) o8[i] += sc[t] * vv[t][i];
        }
      }
      if (active) {
#pragma unroll
        for (int i = 0; i < 8; ++i) o8[i] *= inv;
        Vec16<bf16> ov;
        ov.pack(o8);
        *reinterpret_cast<uint4*>(out + ((beta * Tn + tq) * S + s) * ldo + (int64_t)h * D + d0) = ov.raw;
      }
    }
  }
}

// warp-level bf16 tensor-core tiles for the long-clip temporal attention: 17 x 17 x 72 per (column, head) is far below a
// tcgen05 tile (M = 128), so this is the one place the library uses mma.sync — operands stay bf16 in shared memory and
// the bf16 -> fp32 unpacking that made up ~45 % of the scalar version's instructions disappears.
__device__ __forceinline__ void mma_bf16_k16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_bf16_k8(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// One head of one token column, one warp: scores = q k^T (bf16-rounded), fp32 softmax over the Tn keys, probabilities
// rounded to bf16, out = P v written over the head's k slice.  Tn <= 32.  Fragment layouts of m16n8k16 / m16n8k8
// (g = lane / 4, i = lane % 4): A rows g, g+8, column pairs 2i (+8); B column g, row pairs 2i (+8); C rows g, g+8,
// columns 2i, 2i+1.  The C fragments of two adjacent key tiles are exactly the A fragment of P for the second product.
__device__ __forceinline__ void temporal_head_mma(uint8_t* rows, int pitch, int part_bytes, int Tn, int D) {
  const int lane = threadIdx.x & 31, g = lane >> 2, i2 = (lane & 3) * 2;
  const int last = Tn - 1;
  const int n_mt = (Tn + 15) >> 4;                      // query tiles of 16 (1 or 2)
  float c[2][4][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) c[mt][nt][e] = 0.f;
  // rows beyond the clip are clamped to the last frame: their scores are masked / their outputs never stored
  const uint8_t* qrow[2][2];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
    qrow[mt][0] = rows + min(mt * 16 + g, last) * pitch;
    qrow[mt][1] = rows + min(mt * 16 + g + 8, last) * pitch;
  }
  const uint8_t* krow[4];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) krow[nt] = rows + min(nt * 8 + g, last) * pitch + part_bytes;
  const int n_nt = (Tn + 7) >> 3;                       // key tiles of 8 that hold at least one frame
  int k0 = 0;
  for (; k0 + 16 <= D; k0 += 16) {
    const int off = (k0 + i2) * 2;
    uint32_t a[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      a[mt][0] = *reinterpret_cast<const uint32_t*>(qrow[mt][0] + off);
      a[mt][1] = *reinterpret_cast<const uint32_t*>(qrow[mt][1] + off);
      a[mt][2] = *reinterpret_cast<const uint32_t*>(qrow[mt][0] + off + 16);
      a[mt][3] = *reinterpret_cast<const uint32_t*>(qrow[mt][1] + off + 16);
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      if (nt < n_nt) {
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(krow[nt] + off);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(krow[nt] + off + 16);
        mma_bf16_k16(c[0][nt], a[0], b0, b1);
        if (n_mt > 1) mma_bf16_k16(c[1][nt], a[1], b0, b1);
      }
    }
  }
  if (k0 < D) {                                         // D = 16 m + 8: one k8 step for the tail
    const int off = (k0 + i2) * 2;
    uint32_t a[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      a[mt][0] = *reinterpret_cast<const uint32_t*>(qrow[mt][0] + off);
      a[mt][1] = *reinterpret_cast<const uint32_t*>(qrow[mt][1] + off);
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      if (nt < n_nt) {
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(krow[nt] + off);
        mma_bf16_k8(c[0][nt], a[0][0], a[0][1], b0);
        if (n_mt > 1) mma_bf16_k8(c[1][nt], a[1][0], a[1][1], b0);
      }
    }
  }
  // softmax per query row: a row lives in the four lanes of a quad (elements 0,1 -> row g; 2,3 -> row g + 8)
  uint32_t pa[2][2][4];                                 // P as A fragments: [query tile][key block of 16][4]
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float mx = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int tk = nt * 8 + i2 + e;
          float& v = c[mt][nt][half * 2 + e];
          v = tk < Tn ? round_bf16(v) : -INFINITY;      // q @ k^T leaves the GEMM in bf16
          mx = fmaxf(mx, v);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      float den = 0.f;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          float& v = c[mt][nt][half * 2 + e];
          v = __expf(v - mx);                           // masked keys: exp(-inf) = 0
          den += v;
        }
      den += __shfl_xor_sync(0xffffffffu, den, 1);
      den += __shfl_xor_sync(0xffffffffu, den, 2);
      const float inv = 1.f / den;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
        pa[mt][nt >> 1][(nt & 1) * 2 + half] = pack_bf16x2(c[mt][nt][half * 2] * inv, c[mt][nt][half * 2 + 1] * inv);
    }
  }
  __syncwarp();
  // out = P v: B fragment = v[key pair 2i (+8)][dim g] -> two 16-bit loads per register
  const uint8_t* vbase = rows + 2 * part_bytes + g * 2;
  const int n_kb = (Tn + 15) >> 4;                      // key blocks of 16
  for (int d0 = 0; d0 < D; d0 += 8) {
    float o[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[mt][e] = 0.f;
#pragma unroll
    for (int kb = 0; kb < 2; ++kb) {
      if (kb < n_kb) {
        uint32_t b[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int t0 = kb * 16 + r * 8 + i2;
          const uint32_t lo = *reinterpret_cast<const uint16_t*>(vbase + min(t0, last) * pitch + d0 * 2);
          const uint32_t hi = *reinterpret_cast<const uint16_t*>(vbase + min(t0 + 1, last) * pitch + d0 * 2);
          b[r] = lo | (hi << 16);
        }
        mma_bf16_k16(o[0], pa[0][kb], b[0], b[1]);
        if (n_mt > 1) mma_bf16_k16(o[1], pa[1][kb], b[0], b[1]);
      }
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int tq = mt * 16 + half * 8 + g;
        if (tq < Tn)
          *reinterpret_cast<uint32_t*>(rows + tq * pitch + part_bytes + (d0 + i2) * 2) =
              pack_bf16x2(o[mt][half * 2], o[mt][half * 2 + 1]);
      }
    }
  }
}

// Scores + softmax of ONE (column, head) with T <= 8 frames on the tensor cores: a single m16n8 tile whose rows t_q
// and columns t_k >= T are padding (their fragment registers are zero / never loaded).  Lane (g, c) = (lane / 4,
// lane % 4) loads q[g] and k[g] words of a 16-wide k step (rows 20 banks apart: conflict-free) and ends up with
// S[g][2c], S[g][2c + 1]; the row softmax runs inside the quad.  P[t_q][0 .. 8) goes, as fp32 and rounded to bf16 like
// the reference (blocks.py:255-268), over the first 32 bytes of the q slice of row t_q — where phase C reads it.
// Fewer instructions than one thread per (head, t_q) (which unpacks every k vector T times), but a longer dependent
// chain per warp: measured 16 % slower at T = 5 — opt-in, MDV2_TEMPORAL_MMA=2.
template <int TN>
__device__ __forceinline__ void temporal_scores_mma_t8(uint8_t* rows, int pitch, int part_bytes, int D) {
  const int lane = threadIdx.x & 31, g = lane >> 2, i2 = (lane & 3) * 2;
  const bool row_ok = g < TN;
  uint8_t* qrow = rows + (row_ok ? g : 0) * pitch;
  const uint8_t* krow = qrow + part_bytes;
  float c[4] = {0.f, 0.f, 0.f, 0.f};
  int k0 = 0;
  for (; k0 + 16 <= D; k0 += 16) {
    const int off = (k0 + i2) * 2;
    uint32_t a[4] = {0u, 0u, 0u, 0u};
    uint32_t b0 = 0u, b1 = 0u;
    if (row_ok) {
      a[0] = *reinterpret_cast<const uint32_t*>(qrow + off);
      a[2] = *reinterpret_cast<const uint32_t*>(qrow + off + 16);
      b0 = *reinterpret_cast<const uint32_t*>(krow + off);
      b1 = *reinterpret_cast<const uint32_t*>(krow + off + 16);
    }
    mma_bf16_k16(c, a, b0, b1);
  }
  if (k0 < D) {                                         // D = 16 m + 8: one k8 step for the tail
    const int off = (k0 + i2) * 2;
    uint32_t a0 = 0u, b0 = 0u;
    if (row_ok) {
      a0 = *reinterpret_cast<const uint32_t*>(qrow + off);
      b0 = *reinterpret_cast<const uint32_t*>(krow + off);
    }
    mma_bf16_k8(c, a0, 0u, b0);
  }
  float v0 = i2 < TN ? round_bf16(c[0]) : -INFINITY;    // q @ k^T leaves the GEMM in bf16
  float v1 = i2 + 1 < TN ? round_bf16(c[1]) : -INFINITY;
  float mx = fmaxf(v0, v1);
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
  v0 = __expf(v0 - mx);                                 // padding keys: exp(-inf) = 0
  v1 = __expf(v1 - mx);
  float den = v0 + v1;
  den += __shfl_xor_sync(0xffffffffu, den, 1);
  den += __shfl_xor_sync(0xffffffffu, den, 2);
  const float inv = 1.f / den;
  __syncwarp();                                         // every lane's q words are in registers / consumed
  if (row_ok && (TN > 4 || i2 < 4))
    *reinterpret_cast<float2*>(qrow + i2 * 4) = make_float2(round_bf16(v0 * inv), round_bf16(v1 * inv));
}

// ------------------------------------------------------------------------------------------------
// bf16 production path (T <= 8, D % 8 == 0): shared-memory staged temporal attention.
// A CTA owns TA_G token columns (β, s).  The T rows q|k|v of a column (3C contiguous bf16 each) arrive by
// bulk async copies (one elected thread, one mbarrier) — three resident CTAs per SM keep ~140 KB of reads
// in flight while the others compute — and four short phases run out of shared memory:
//   A  RoPE (interleaved pairs) on q and k in place, rounded to bf16 like the reference; q also x scale
//   B  thread = (column, head, t_q): T scores, fp32 softmax, probabilities rounded to bf16 (reference order
//      of roundings, blocks.py:255-268) and parked in the first bytes of the thread's own q slice
//   C  thread = (column, head, 8-dim slice): out[t_q] = sum_k P[t_q][k] v[k] for all t_q, written over k
//   D  bulk async stores of the T output rows
// Row pitch = 3C·2 + 80 bytes: the 20-word skew makes the (head, t_q) threads of phase B hit eight distinct
// 16-byte bank groups per quarter warp (bank group = (head + 5·t_q) mod 8 = 5·thread mod 8).
// The strided gather / scatter never materialises the reference's two transposes (stdit3.py:232-234).
// ------------------------------------------------------------------------------------------------
constexpr int TA_G = 2, TA_THREADS = 256, TA_PAD = 80;
// geometry of the short-clip kernel (tuning hooks for tools/gpu_temporal_variants.sh)
#ifndef MDV2_TS_G
#define MDV2_TS_G 2
#define MDV2_TS_THREADS 256
#define MDV2_TS_MINB 3
#endif
constexpr int TS_G = MDV2_TS_G, TS_THREADS = MDV2_TS_THREADS, TS_MINB = MDV2_TS_MINB;
constexpr int TS_SMEM_MAX = TS_MINB == 3 ? 75 * 1024 : (227 * 1024) / TS_MINB - 1024;

template <int TN>
__global__ void __launch_bounds__(TS_THREADS, TS_MINB) temporal_attn_smem_kernel(const bf16* __restrict__ qkv, int64_t ld,
                                                                          const float* __restrict__ cos_t,
                                                                          const float* __restrict__ sin_t,
                                                                          bf16* __restrict__ out, int64_t ldo,
                                                                          int64_t n_cols, int S, int H, int D,
                                                                          float scale, int use_mma) {
  using namespace sm100;
  extern __shared__ __align__(128) uint8_t ta_smem[];
  const int C = H * D;
  const int row_bytes = 3 * C * 2, pitch = row_bytes + TA_PAD;
  uint64_t* bar = reinterpret_cast<uint64_t*>(ta_smem + TS_G * TN * pitch);
  const int tid = threadIdx.x;
  const int64_t col0 = (int64_t)blockIdx.x * TS_G;
  const int ng = (int)((n_cols - col0) < TS_G ? (n_cols - col0) : TS_G);
  auto grow = [&](int g, int t) {                       // global row of (column col0+g, frame t)
    const int64_t c = col0 + g, beta = c / S, s = c - beta * S;
    return (beta * TN + t) * S + s;
  };
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    mbar_expect_tx(bar, (uint32_t)(ng * TN * row_bytes));
    for (int g = 0; g < ng; ++g)
      for (int t = 0; t < TN; ++t)
        bulk_g2s(ta_smem + (g * TN + t) * pitch, qkv + grow(g, t) * ld, (uint32_t)row_bytes, bar);
  }
  __syncthreads();                                      // barrier init visible to the waiters
  mbar_wait(bar, 0);

  const int vpr = C / 8;                                // 16-byte vectors per q (or k, v) row part
  const int vph = D / 8;                                // ... per head
  // ---- A: RoPE on q and k ----
  // A thread keeps ONE 16-byte vector position of the q|k part of a row (byte offset u·16) and walks the frames and
  // columns: the head-dimension offset is divided out once per thread and the cos / sin vectors are loaded once per
  // frame.  (One flat index over (column, frame, part, vector) cost 181 instructions per vector, 59 % of the kernel's
  // instructions, most of them the four runtime divisions and the angle loads.)
  for (int u = tid; u < 2 * vpr; u += TS_THREADS) {
    const bool is_q = u < vpr;
    const int d0 = ((is_q ? u : u - vpr) % vph) * 8;
#pragma unroll
    for (int t = 0; t < TN; ++t) {
      const float4 c0 = *reinterpret_cast<const float4*>(cos_t + t * D + d0);
      const float4 c1 = *reinterpret_cast<const float4*>(cos_t + t * D + d0 + 4);
      const float4 s0 = *reinterpret_cast<const float4*>(sin_t + t * D + d0);
      const float4 s1 = *reinterpret_cast<const float4*>(sin_t + t * D + d0 + 4);
      const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
      const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
      for (int g = 0; g < TS_G; ++g) {
        if (g < ng) {
          uint4* ptr = reinterpret_cast<uint4*>(ta_smem + (g * TN + t) * pitch + u * 16);
          Vec16<bf16> x;
          x.raw = *ptr;
          float f[8], o[8];
          x.unpack(f);
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            o[e] = f[e] * cs[e] - f[e + 1] * sn[e];
            o[e + 1] = f[e + 1] * cs[e + 1] + f[e] * sn[e + 1];
          }
          x.pack(o);
          if (is_q) {                                   // q = (rope(q) in bf16) * scale, in bf16
            x.unpack(f);
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] *= scale;
            x.pack(f);
          }
          *ptr = x.raw;
        }
      }
    }
  }
  __syncthreads();
  if (use_mma == 1) {
    // ---- B + C on the tensor cores (mma.sync tiles, one warp per (column, head); see temporal_head_mma) ----
    for (int idx = tid >> 5; idx < ng * H; idx += TS_THREADS / 32) {
      const int h = idx % H, g = idx / H;
      temporal_head_mma(ta_smem + g * TN * pitch + h * D * 2, pitch, C * 2, TN, D);
    }
  } else {
  // ---- B: scores + softmax; P[t_q][0..TN) -> first bytes of the q slice of row t_q ----
  if (use_mma == 2) {                                   // one warp per (column, head): one m16n8 tile (opt-in)
    for (int idx = tid >> 5; idx < ng * H; idx += TS_THREADS / 32) {
      const int h = idx % H, g = idx / H;
      temporal_scores_mma_t8<TN>(ta_smem + g * TN * pitch + h * D * 2, pitch, C * 2, D);
    }
  } else if (tid < ng * H * TN) {                       // default: one thread per (column, head, t_q)
    const int tq = tid % TN, h = (tid / TN) % H, g = tid / (TN * H);
    const uint8_t* qs = ta_smem + (g * TN + tq) * pitch + h * D * 2;
    // dot products on packed fp32 pairs (FFMA2: even / odd dimensions accumulate separately, added at the end)
    float2 acc2[TN];
#pragma unroll
    for (int t = 0; t < TN; ++t) acc2[t] = make_float2(0.f, 0.f);
    for (int j = 0; j < vph; ++j) {
      Vec16<bf16> qv;
      qv.raw = *reinterpret_cast<const uint4*>(qs + j * 16);
      float q[8];
      qv.unpack(q);
#pragma unroll
      for (int t = 0; t < TN; ++t) {
        Vec16<bf16> kv;
        kv.raw = *reinterpret_cast<const uint4*>(ta_smem + (g * TN + t) * pitch + C * 2 + h * D * 2 + j * 16);
        float k[8];
        kv.unpack(k);
#pragma unroll
        for (int e = 0; e < 8; e += 2)
          acc2[t] = sm100::ffma2(make_float2(q[e], q[e + 1]), make_float2(k[e], k[e + 1]), acc2[t]);
      }
    }
    float acc[TN];
#pragma unroll
    for (int t = 0; t < TN; ++t) acc[t] = acc2[t].x + acc2[t].y;
    float mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < TN; ++t) {
      acc[t] = __bfloat162float(__float2bfloat16_rn(acc[t]));      // q @ k^T leaves the GEMM in bf16
      mx = fmaxf(mx, acc[t]);
    }
    float den = 0.f;
#pragma unroll
    for (int t = 0; t < TN; ++t) {
      acc[t] = __expf(acc[t] - mx);
      den += acc[t];
    }
    const float inv = 1.f / den;
    float pr[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) pr[t] = t < TN ? __bfloat162float(__float2bfloat16_rn(acc[t] * inv)) : 0.f;
    float4* pd = reinterpret_cast<float4*>(const_cast<uint8_t*>(qs));
    pd[0] = make_float4(pr[0], pr[1], pr[2], pr[3]);
    if (TN > 4) pd[1] = make_float4(pr[4], pr[5], pr[6], pr[7]);
  }
  __syncthreads();
  // ---- C: out[t_q][8 dims] = sum_k P[t_q][k] v[k][8 dims], written over the k part of row t_q ----
  for (int i = tid; i < ng * H * vph; i += TS_THREADS) {
    const int j = i % vph, h = (i / vph) % H, g = i / (vph * H);
    float v[TN][8];
#pragma unroll
    for (int t = 0; t < TN; ++t) {
      Vec16<bf16> vv;
      vv.raw = *reinterpret_cast<const uint4*>(ta_smem + (g * TN + t) * pitch + 2 * C * 2 + h * D * 2 + j * 16);
      vv.unpack(v[t]);
    }
#pragma unroll
    for (int tq = 0; tq < TN; ++tq) {
      const float4* pp = reinterpret_cast<const float4*>(ta_smem + (g * TN + tq) * pitch + h * D * 2);
      const float4 p0 = pp[0];
      float pr[8] = {p0.x, p0.y, p0.z, p0.w, 0.f, 0.f, 0.f, 0.f};
      if (TN > 4) {
        const float4 p1 = pp[1];
        pr[4] = p1.x; pr[5] = p1.y; pr[6] = p1.z; pr[7] = p1.w;
      }
      float2 o2[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) o2[e] = make_float2(0.f, 0.f);
#pragma unroll
      for (int t = 0; t < TN; ++t) {
        const float2 pt = make_float2(pr[t], pr[t]);
#pragma unroll
        for (int e = 0; e < 4; ++e) o2[e] = sm100::ffma2(pt, make_float2(v[t][2 * e], v[t][2 * e + 1]), o2[e]);
      }
      const float o[8] = {o2[0].x, o2[0].y, o2[1].x, o2[1].y, o2[2].x, o2[2].y, o2[3].x, o2[3].y};
      Vec16<bf16> ov;
      ov.pack(o);
      *reinterpret_cast<uint4*>(ta_smem + (g * TN + tq) * pitch + C * 2 + h * D * 2 + j * 16) = ov.raw;
    }
  }
  }
  fence_async_smem();                                   // generic-proxy writes -> visible to the bulk stores
  __syncthreads();
  // ---- D: T output rows per column ----
  if (tid == 0) {
    for (int g = 0; g < ng; ++g)
      for (int t = 0; t < TN; ++t)
        bulk_s2g(out + grow(g, t) * ldo, ta_smem + (g * TN + t) * pitch + C * 2, (uint32_t)(C * 2));
    bulk_commit();
    bulk_wait_read0();                                  // shared memory must outlive the reads
  }
}

// ------------------------------------------------------------------------------------------------
// bf16 production path for LONG clips (8 < T <= 32, e.g. 65 pixel frames -> T = 17): same four phases, but a CTA
// owns ONE token column and a GROUP of HG heads, so the staged q|k|v slices of all T frames stay near 60 KB and
// three CTAs fit an SM (a whole column at T = 17 is 118 KB: one CTA per SM, nothing in flight while it computes).
// Shared-memory row t = [q slice | k slice | v slice] of HG·D bf16 each (three bulk copies per frame, 1152 B at
// HG = 8, D = 72), pitch skewed by 80 B like above.
//   A  RoPE on q and k in place (+ scale on q)
//   B  item = (head, t_q, t_k): one score -> bf16, into a [HG][T][ceil4(T)] fp32 table behind the rows; then
//      item = (head, t_q): fp32 softmax over its table row, probabilities rounded to bf16 in place.  (One thread per
//      (head, t_q) doing all T dot products kept 136 of 256 threads busy on a ~3000-instruction serial chain:
//      3.6 ms per launch at T = 17.)
//   C  item = (t_q, head, 8-dim slice): out = sum_k P[t_q][k] v[k], written over the k slice of row t_q
//   D  one bulk store per frame
// The first version sent these shapes to the warp-per-(column, head) kernel above: 23.7 ms per launch at T = 17
// (127 GB/s), 38 % of the 65-frame step.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TA_THREADS, 3) temporal_attn_smem_long_kernel(const bf16* __restrict__ qkv, int64_t ld,
                                                                               const float* __restrict__ cos_t,
                                                                               const float* __restrict__ sin_t,
                                                                               bf16* __restrict__ out, int64_t ldo,
                                                                               int64_t n_cols, int Tn, int S, int H,
                                                                               int D, int HG, float scale, int use_mma) {
  using namespace sm100;
  extern __shared__ __align__(128) uint8_t ta_smem[];
  const int C = H * D, W = HG * D;                      // elements per q (k, v) slice of the head group
  const int part_bytes = W * 2, pitch = 3 * part_bytes + TA_PAD;
  uint64_t* bar = reinterpret_cast<uint64_t*>(ta_smem + Tn * pitch);
  const int tid = threadIdx.x;
  const int n_groups = H / HG;
  const int64_t col = (int64_t)blockIdx.x / n_groups;
  const int hg = (int)((int64_t)blockIdx.x - col * n_groups);
  const int64_t beta = col / S, s = col - beta * S;
  const int64_t row0 = beta * Tn * S + s;               // global row of frame 0; frame t is row0 + t·S
  if (tid < 32) {                                       // warp 0: lane t issues the three copies of frame t
    if (tid == 0) {
      mbar_init(bar, 1);
      fence_mbar_init();
      mbar_expect_tx(bar, (uint32_t)(Tn * 3 * part_bytes));
    }
    __syncwarp();
    for (int t = tid; t < Tn; t += 32) {
      const bf16* src = qkv + (row0 + (int64_t)t * S) * ld + (int64_t)hg * W;
#pragma unroll
      for (int part = 0; part < 3; ++part)
        bulk_g2s(ta_smem + t * pitch + part * part_bytes, src + (int64_t)part * C, (uint32_t)part_bytes, bar);
    }
  }
  __syncthreads();                                      // barrier init visible to the waiters
  mbar_wait(bar, 0);

  const int vpp = W / 8;                                // 16-byte vectors per slice
  const int vph = D / 8;                                // ... per head
  // ---- A: RoPE on q and k: a thread owns one 16-byte vector position (part, v) and walks the frames ----
  for (int pv = tid; pv < 2 * vpp; pv += TA_THREADS) {
    const int v = pv % vpp, part = pv / vpp;
    const int d0 = (v % vph) * 8;
    for (int t = 0; t < Tn; ++t) {
      uint4* ptr = reinterpret_cast<uint4*>(ta_smem + t * pitch + part * part_bytes + v * 16);
      Vec16<bf16> x;
      x.raw = *ptr;
      float f[8], o[8];
      x.unpack(f);
      const float4 c0 = *reinterpret_cast<const float4*>(cos_t + t * D + d0);
      const float4 c1 = *reinterpret_cast<const float4*>(cos_t + t * D + d0 + 4);
      const float4 s0 = *reinterpret_cast<const float4*>(sin_t + t * D + d0);
      const float4 s1 = *reinterpret_cast<const float4*>(sin_t + t * D + d0 + 4);
      const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
      const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
      for (int e = 0; e < 8; e += 2) {
        o[e] = f[e] * cs[e] - f[e + 1] * sn[e];
        o[e + 1] = f[e + 1] * cs[e + 1] + f[e] * sn[e + 1];
      }
      x.pack(o);
      if (part == 0) {                                  // q = (rope(q) in bf16) * scale, in bf16
        x.unpack(f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] *= scale;
        x.pack(f);
      }
      *ptr = x.raw;
    }
  }
  __syncthreads();
  if (use_mma) {
    // ---- B + C on the tensor cores: one warp per head ----
    for (int h = tid >> 5; h < HG; h += TA_THREADS / 32) temporal_head_mma(ta_smem + h * D * 2, pitch, part_bytes, Tn, D);
    fence_async_smem();
    __syncthreads();
    if (tid < 32) {                                     // lane t stores frame t (a bulk group is per thread)
      for (int t = tid; t < Tn; t += 32)
        bulk_s2g(out + (row0 + (int64_t)t * S) * ldo + (int64_t)hg * W, ta_smem + t * pitch + part_bytes,
                 (uint32_t)part_bytes);
      bulk_commit();
      bulk_wait_read0();                                // shared memory must outlive the reads
    }
    return;
  }
  // ---- scalar fallback (MDV2_TEMPORAL_LONG_SCALAR=1) ----
  // ---- B1: item = (head, t_q, t_k): one score, rounded to bf16 like the reference's q @ k^T ----
  const int Tp = (Tn + 3) & ~3;
  float* sc = reinterpret_cast<float*>(ta_smem + Tn * pitch + 16);      // [HG][Tn][Tp] scores, then probabilities
  for (int i = tid; i < HG * Tn * Tn; i += TA_THREADS) {
    const int tk = i % Tn, tq = (i / Tn) % Tn, h = i / (Tn * Tn);
    const uint8_t* qs = ta_smem + tq * pitch + h * D * 2;
    const uint8_t* ks = ta_smem + tk * pitch + part_bytes + h * D * 2;
    float acc = 0.f;
    for (int j = 0; j < vph; ++j) {
      Vec16<bf16> qv, kv;
      qv.raw = *reinterpret_cast<const uint4*>(qs + j * 16);
      kv.raw = *reinterpret_cast<const uint4*>(ks + j * 16);
      float q[8], k[8];
      qv.unpack(q);
      kv.unpack(k);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc += q[e] * k[e];
    }
    sc[(h * Tn + tq) * Tp + tk] = __bfloat162float(__float2bfloat16_rn(acc));
  }
  __syncthreads();
  // ---- B2: item = (head, t_q): fp32 softmax over its row, probabilities rounded to bf16, zero padding ----
  for (int i = tid; i < HG * Tn; i += TA_THREADS) {
    float* row = sc + i * Tp;                           // i = head·Tn + t_q
    float mx = -INFINITY;
    for (int t = 0; t < Tn; ++t) mx = fmaxf(mx, row[t]);
    float den = 0.f;
    for (int t = 0; t < Tn; ++t) {
      const float e = __expf(row[t] - mx);
      row[t] = e;
      den += e;
    }
    const float inv = 1.f / den;
    for (int t = 0; t < Tn; ++t) row[t] = __bfloat162float(__float2bfloat16_rn(row[t] * inv));
    for (int t = Tn; t < Tp; ++t) row[t] = 0.f;
  }
  __syncthreads();
  // ---- C: out[t_q][8 dims] = sum_k P[t_q][k] v[k][8 dims], written over the k slice of row t_q ----
  for (int i = tid; i < Tn * HG * vph; i += TA_THREADS) {
    const int j = i % vph, h = (i / vph) % HG, tq = i / (vph * HG);
    const float4* pp = reinterpret_cast<const float4*>(sc + (h * Tn + tq) * Tp);
    const uint8_t* vbase = ta_smem + 2 * part_bytes + h * D * 2 + j * 16;
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int t4 = 0; t4 * 4 < Tn; ++t4) {
      const float4 p4 = pp[t4];
      const float pr[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
      for (int e4 = 0; e4 < 4; ++e4) {
        const int t = t4 * 4 + e4;
        if (t < Tn) {
          Vec16<bf16> vv;
          vv.raw = *reinterpret_cast<const uint4*>(vbase + t * pitch);
          float v[8];
          vv.unpack(v);
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] += pr[e4] * v[e];
        }
      }
    }
    Vec16<bf16> ov;
    ov.pack(o);
    *reinterpret_cast<uint4*>(ta_smem + tq * pitch + part_bytes + h * D * 2 + j * 16) = ov.raw;
  }
  fence_async_smem();                                   // generic-proxy writes -> visible to the bulk stores
  __syncthreads();
  // ---- D: one output slice per frame ----
  if (tid == 0) {
    for (int t = 0; t < Tn; ++t)
      bulk_s2g(out + (row0 + (int64_t)t * S) * ldo + (int64_t)hg * W, ta_smem + t * pitch + part_bytes,
               (uint32_t)part_bytes);
    bulk_commit();
    bulk_wait_read0();                                  // shared memory must outlive the reads
  }
}


template <typename T>
static int launch_temporal(const void* qkv, int64_t ld, const float* c, const float* s, void* out, int64_t ldo,
                           int B, int Tn, int S, int H, int D, float scale, cudaStream_t st) {
  const int64_t n_warps = (int64_t)B * S * H;
  if constexpr (sizeof(T) == 2) {
    const int C = H * D;
    const size_t smem = (size_t)TS_G * Tn * (3 * C * 2 + TA_PAD) + 16;
    static const bool legacy = getenv("MDV2_TEMPORAL_LEGACY") != nullptr;       // tuning hook (tools/bench_elem.py)
    if (!legacy && Tn <= 8 && D % 8 == 0 && D * 2 >= 32 && ld % 8 == 0 && ldo % 8 == 0 && ((uintptr_t)qkv % 16 == 0) &&
        ((uintptr_t)out % 16 == 0) && ((uintptr_t)c % 16 == 0) && ((uintptr_t)s % 16 == 0) && smem <= (size_t)TS_SMEM_MAX &&
        TS_G * H * Tn <= TS_THREADS) {
      const int64_t n_cols = (int64_t)B * S;
      const unsigned nb = (unsigned)((n_cols + TS_G - 1) / TS_G);
      // A/B hook: 0 scalar scores (default), 1 tensor-core scores AND P·V (temporal_head_mma), 2 tensor-core scores.
      // Measured at T = 5, 67 200 columns: 0.874 / 1.202 / 1.016 ms — one warp per head is a chain of five dependent
      // mma.sync on one accumulator; the scalar phase keeps 160 threads busy instead
      static const int short_mma = [] {
        const char* e = getenv("MDV2_TEMPORAL_MMA");
        return e ? atoi(e) : 0;
      }();
#define TS_CASE(TT)                                                                                               \
  case TT: {                                                                                                      \
    static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();                                                                               \
    if (!configured) {                                                                                            \
      cudaFuncSetAttribute(temporal_attn_smem_kernel<TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM_MAX); \
      configured = true;                                                                                          \
    }                                                                                                             \
    temporal_attn_smem_kernel<TT><<<nb, TS_THREADS, smem, st>>>((const bf16*)qkv, ld, c, s, (bf16*)out, ldo, n_cols, \
                                                               S, H, D, scale, short_mma);                        \
  } break
      switch (Tn) {
        TS_CASE(1); TS_CASE(2); TS_CASE(3); TS_CASE(4); TS_CASE(5); TS_CASE(6); TS_CASE(7); TS_CASE(8);
      }
#undef TS_CASE
      MDV2_CHECK_LAUNCH("temporal_attention");
      return MDV2_OK;
    }
    if (!legacy && Tn > 8 && Tn <= 128 && D % 8 == 0 && ld % 8 == 0 && ldo % 8 == 0 && ((uintptr_t)qkv % 16 == 0) &&
        ((uintptr_t)out % 16 == 0) && ((uintptr_t)c % 16 == 0) && ((uintptr_t)s % 16 == 0)) {
      // largest head group (a divisor of H, at most 8 heads) whose staged frames + score table leave room for three
      // CTAs per SM
      // (the warp-level MMA phases hold a 32 x 32 score tile per head; longer clips — 129 frames -> T = 33, the
      // reference's "full" 241-frame config -> T = 61 — take the scalar phases, which loop over T)
      static const bool scalar_env = getenv("MDV2_TEMPORAL_LONG_SCALAR") != nullptr;     // A/B hook
      const bool scalar = scalar_env || Tn > 32;
      const int Tp = (Tn + 3) & ~3;
      auto smem_long = [&](int g) {
        return (size_t)Tn * (3 * g * D * 2 + TA_PAD) + 16 + (scalar ? (size_t)g * Tn * Tp * 4 : 0);
      };
      // (very long clips, T ~ 100: not even one head fits 75 KB — one head per CTA, up to 200 KB, one or two CTAs per SM)
      int HG = 0;
      for (int g = H < 8 ? H : 8; g >= 1; --g)
        if (H % g == 0 && smem_long(g) <= 75 * 1024) {
          HG = g;
          break;
        }
      if (!HG && smem_long(1) <= 200 * 1024) HG = 1;
      const int64_t n_ctas = (int64_t)B * S * (HG ? H / HG : 0);
      if (HG && n_ctas <= 0x7fffffff) {
        static PerDeviceFlag configured_flag; bool& configured = configured_flag.here();
        if (!configured) {
          cudaFuncSetAttribute(temporal_attn_smem_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
          configured = true;
        }
        temporal_attn_smem_long_kernel<<<(unsigned)n_ctas, TA_THREADS, smem_long(HG), st>>>(
            (const bf16*)qkv, ld, c, s, (bf16*)out, ldo, (int64_t)B * S, Tn, S, H, D, HG, scale, scalar ? 0 : 1);
        MDV2_CHECK_LAUNCH("temporal_attention");
        return MDV2_OK;
      }
    }
    if (Tn <= 8 && D % 8 == 0 && D <= 128 && ld % 8 == 0 && ldo % 8 == 0 && ((uintptr_t)qkv % 16 == 0) &&
        ((uintptr_t)out % 16 == 0)) {
      const unsigned nb = (unsigned)((n_warps + 15) / 16);
#define TV_CASE(TT)                                                                                              \
  case TT:                                                                                                       \
    temporal_attn_vec_kernel<TT><<<nb, 256, 0, st>>>((const bf16*)qkv, ld, c, s, (bf16*)out, ldo, n_warps, S, H, D, \
                                                     scale);                                                     \
    break
      switch (Tn) {
        TV_CASE(1); TV_CASE(2); TV_CASE(3); TV_CASE(4); TV_CASE(5); TV_CASE(6); TV_CASE(7); TV_CASE(8);
      }
#undef TV_CASE
      MDV2_CHECK_LAUNCH("temporal_attention");
      return MDV2_OK;
    }
  }
  MDV2_REQUIRE(Tn <= 64, "temporal_attention: T=%d not taken by the staged kernels (alignment / size) and beyond the "
               "register kernel's 64 frames", Tn);
  dim3 grid((unsigned)((n_warps + 3) / 4)), block(128);
  const int ppl = (D / 2 + 31) / 32;
#define TA_CASE(TM, PP)                                                                                \
  temporal_attn_kernel<T, TM, PP><<<grid, block, 0, st>>>((const T*)qkv, ld, c, s, (T*)out, ldo, n_warps, \
                                                          Tn, S, H, D, scale)
  if (ppl == 1) {
    if (Tn <= 8) TA_CASE(8, 1); else if (Tn <= 32) TA_CASE(32, 1); else TA_CASE(64, 1);
  } else {
    if (Tn <= 8) TA_CASE(8, 2); else if (Tn <= 32) TA_CASE(32, 2); else TA_CASE(64, 2);
  }
#undef TA_CASE
  MDV2_CHECK_LAUNCH("temporal_attention");
  return MDV2_OK;
}

extern "C" int mdv2_temporal_attention(const void* qkv, int64_t ld, const float* cos_t, const float* sin_t,
                                       void* out, int64_t ldo, int B, int T, int S, int H, int D, float scale,
                                       int dtype, void* stream) {
  if ((int64_t)B * S * H == 0) return MDV2_OK;
  // bf16: T <= 8 staged kernel, 8 < T <= 32 warp-MMA kernel, 32 < T <= 128 the same staging with scalar phases;
  // fp32 (validation mode) and the shapes the staged kernels do not take: the register kernel, T <= 64
  MDV2_REQUIRE(T >= 1 && T <= (dtype == 1 ? 128 : 64), "temporal_attention: T=%d unsupported (1..%d)", T,
               dtype == 1 ? 128 : 64);
  MDV2_REQUIRE(D % 2 == 0 && D <= 128, "temporal_attention: head_dim=%d unsupported", D);
  MDV2_REQUIRE(ld % 2 == 0 && ldo % 2 == 0, "temporal_attention: odd leading dimension");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) return launch_temporal<float>(qkv, ld, cos_t, sin_t, out, ldo, B, T, S, H, D, scale, st);
  if (dtype == 1) return launch_temporal<bf16>(qkv, ld, cos_t, sin_t, out, ldo, B, T, S, H, D, scale, st);
  MDV2_REQUIRE(false, "temporal_attention: bad dtype");
}

// ------------------------------------------------------------------------------------------------
// swap01: dst[b][a][:] = src[a][b][:] with a contiguous inner run of `inner_bytes` (multiple of 16).
// The four re-orderings around the sequence-parallel all-to-alls are all of this form:
//   qkv [n·3][p][HpD]      -> send  [p][n·3][HpD]        (heads scatter)
//   recv [p][β·T][Sl·3·HpD] -> full [β·T][p][Sl·3·HpD]    (sequence gather)
//   o    [β·T][p][Sl·HpD]   -> send2 [p][β·T][Sl·HpD]
//   recv2 [p][n][HpD]       -> ao    [n][p][HpD]          (heads gather)
// Replaces the tensor_split / cat / contiguous copies of communications.py:11-21.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) swap01_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                                     uint32_t A, uint32_t B, uint32_t inner, int64_t total) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / inner;                   // destination row = b·A + a
    const uint32_t v = (uint32_t)(idx - row * inner);
    const uint32_t b = (uint32_t)(row / A), a = (uint32_t)(row - (int64_t)b * A);
    dst[idx] = src[((int64_t)a * B + b) * inner + v];
  }
}

extern "C" int mdv2_swap01(const void* src, void* dst, int64_t A, int64_t B, int64_t inner_bytes, void* stream) {
  if (A * B * inner_bytes == 0) return MDV2_OK;
  MDV2_REQUIRE(A > 0 && B > 0 && inner_bytes > 0 && inner_bytes % 16 == 0, "swap01: inner run must be a multiple of 16 bytes");
  MDV2_REQUIRE(((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 16 == 0), "swap01: operands must be 16-byte aligned");
  MDV2_REQUIRE(A < (1ll << 31) && B < (1ll << 31) && inner_bytes / 16 < (1ll << 31), "swap01: dimension too large");
  const int64_t inner = inner_bytes / 16, total = A * B * inner;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)mdv2_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  swap01_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, (uint4*)dst, (uint32_t)A,
                                                                    (uint32_t)B, (uint32_t)inner, total);
  MDV2_CHECK_LAUNCH("swap01");
  return MDV2_OK;
}

// ------------------------------------------------------------------------------------------------
// scatter_runs: see include/mdv2.h.  One 16-byte vector per thread and grid-stride step; the run index is
// decomposed with 32-bit divisions (a few ALU ops per 16 bytes against an NVLink- or HBM-bound copy).
// ------------------------------------------------------------------------------------------------
struct ScatterArgs {
  uint4* dst[8];
  uint32_t size[4];
  int64_t sstr[4], dstr[4];                            // in 16-byte units
  int64_t dst_base;
  uint32_t run;                                        // 16-byte vectors per run
  int peer_dim;
};

__global__ void __launch_bounds__(256) scatter_runs_kernel(const uint4* __restrict__ src, const ScatterArgs a,
                                                           int64_t total) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / a.run;
    const uint32_t v = (uint32_t)(idx - r * a.run);
    // r = ((i0·size1 + i1)·size2 + i2)·size3 + i3
    const int64_t r2 = r / a.size[3];
    const uint32_t i3 = (uint32_t)(r - r2 * a.size[3]);
    const int64_t r1 = r2 / a.size[2];
    const uint32_t i2 = (uint32_t)(r2 - r1 * a.size[2]);
    const uint32_t i0 = (uint32_t)(r1 / a.size[1]);
    const uint32_t i1 = (uint32_t)(r1 - (int64_t)i0 * a.size[1]);
    const uint32_t ii[4] = {i0, i1, i2, i3};
    int64_t so = v, dof = a.dst_base + v;
    uint32_t peer = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      so += (int64_t)ii[k] * a.sstr[k];
      if (k == a.peer_dim) peer = ii[k];
      else dof += (int64_t)ii[k] * a.dstr[k];
    }
    a.dst[peer][dof] = src[so];
  }
}

extern "C" int mdv2_scatter_runs(const void* src, void* const* dst_ptrs, int n_dst, int peer_dim, const int64_t* sizes,
                                 const int64_t* src_strides, const int64_t* dst_strides, int64_t run_bytes,
                                 int64_t dst_base, void* stream) {
  MDV2_REQUIRE(n_dst >= 1 && n_dst <= 8, "scatter_runs: 1..8 destinations");
  MDV2_REQUIRE(peer_dim >= 0 && peer_dim < 4 && sizes[peer_dim] == n_dst, "scatter_runs: peer dimension must span the destinations");
  MDV2_REQUIRE(run_bytes > 0 && run_bytes % 16 == 0 && dst_base % 16 == 0 && ((uintptr_t)src % 16 == 0),
               "scatter_runs: runs must be 16-byte aligned");
  ScatterArgs a;
  int64_t total = run_bytes / 16;
  for (int k = 0; k < 4; ++k) {
    MDV2_REQUIRE(sizes[k] >= 1 && sizes[k] < (1ll << 31), "scatter_runs: bad size");
    MDV2_REQUIRE(src_strides[k] % 16 == 0 && dst_strides[k] % 16 == 0, "scatter_runs: strides must be multiples of 16 bytes");
    a.size[k] = (uint32_t)sizes[k];
    a.sstr[k] = src_strides[k] / 16;
    a.dstr[k] = dst_strides[k] / 16;
    total *= sizes[k];
  }
  for (int i = 0; i < 8; ++i) {
    a.dst[i] = (uint4*)(i < n_dst ? dst_ptrs[i] : nullptr);
    MDV2_REQUIRE(i >= n_dst || (dst_ptrs[i] != nullptr && (uintptr_t)dst_ptrs[i] % 16 == 0), "scatter_runs: bad destination");
  }
  a.dst_base = dst_base / 16;
  a.run = (uint32_t)(run_bytes / 16);
  a.peer_dim = peer_dim;
  if (total == 0) return MDV2_OK;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)mdv2_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  scatter_runs_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, a, total);
  MDV2_CHECK_LAUNCH("scatter_runs");
  return MDV2_OK;
}
