// sampler.cu — the per-step tail of RFLOW.sample (reference: magicdrivedit/schedulers/rf/__init__.py:93-113):
//   pred_cond, pred_uncond = pred.chunk(2)                      (:100)
//   v   = pred_uncond + g * (pred_cond - pred_uncond)           (:101)
//   z'  = z + v * dt[:, None, None, None, None]                 (:104-106)
//   z'  = where(mask_t_upper[:, None, :, None, None], z', x0)   (:108-109, only with a frame mask)
//   z_in = cat([z', z'])  (+ the model's own x.to(dtype))       (:95 of the NEXT step, magicdrive_stdit3.py:1041)
// One HBM pass instead of six eager ones.  Every product and sum is rounded separately (__fmul_rn / __fadd_rn:
// no FMA contraction), so the fp32 result is bit-identical to the eager sequence.
#include "common.cuh"

template <typename TZ, int V>
__global__ void __launch_bounds__(256) cfg_euler_kernel(const float* __restrict__ pred, const TZ* __restrict__ z,
                                                        const float* __restrict__ dt, float g,
                                                        const TZ* __restrict__ x0, const uint8_t* __restrict__ keep,
                                                        float* __restrict__ z_out, bf16* __restrict__ zin_bf16,
                                                        float* __restrict__ zin_f32, int64_t per_sample, int64_t hw,
                                                        int T, int64_t total) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * V;
  for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V; e < total; e += stride) {
    const int64_t b = e / per_sample;
    const float d = dt[b];
    float pc[V], pu[V], zz[V], o[V];
    if constexpr (V == 4) {
      const float4 c4 = *reinterpret_cast<const float4*>(pred + e);
      const float4 u4 = *reinterpret_cast<const float4*>(pred + total + e);
      pc[0] = c4.x; pc[1] = c4.y; pc[2] = c4.z; pc[3] = c4.w;
      pu[0] = u4.x; pu[1] = u4.y; pu[2] = u4.z; pu[3] = u4.w;
    } else {
      pc[0] = pred[e];
      pu[0] = pred[total + e];
    }
#pragma unroll
    for (int i = 0; i < V; ++i) zz[i] = to_f(z[e + i]);
    bool kp = true;
    if (keep != nullptr) kp = keep[b * T + (e / hw) % T] != 0;   // V consecutive elements share (b, t): hw % V == 0
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float v = __fadd_rn(pu[i], __fmul_rn(g, __fsub_rn(pc[i], pu[i])));
      o[i] = __fadd_rn(zz[i], __fmul_rn(v, d));
      if (!kp) o[i] = to_f(x0[e + i]);
    }
    if constexpr (V == 4) {
      *reinterpret_cast<float4*>(z_out + e) = make_float4(o[0], o[1], o[2], o[3]);
      if (zin_f32 != nullptr) {
        *reinterpret_cast<float4*>(zin_f32 + e) = make_float4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<float4*>(zin_f32 + total + e) = make_float4(o[0], o[1], o[2], o[3]);
      }
      if (zin_bf16 != nullptr) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o[0], o[1]), hi = __floats2bfloat162_rn(o[2], o[3]);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(zin_bf16 + e) = pk;
        *reinterpret_cast<uint2*>(zin_bf16 + total + e) = pk;
      }
    } else {
      z_out[e] = o[0];
      if (zin_f32 != nullptr) zin_f32[e] = zin_f32[total + e] = o[0];
      if (zin_bf16 != nullptr) zin_bf16[e] = zin_bf16[total + e] = __float2bfloat16_rn(o[0]);
    }
  }
}

template <typename TZ>
static int launch_cfg_euler(const float* pred, const void* z, const float* dt, float g, const void* x0,
                            const uint8_t* keep, float* z_out, void* z_in, int zin_dtype, int n, int C, int T,
                            int64_t hw, cudaStream_t st) {
  const int64_t per_sample = (int64_t)C * T * hw, total = per_sample * n;
  bf16* zb = zin_dtype == 1 ? (bf16*)z_in : nullptr;
  float* zf = zin_dtype == 0 ? (float*)z_in : nullptr;
  const bool vec = hw % 4 == 0 && ((uintptr_t)pred % 16 == 0) && ((uintptr_t)z % 16 == 0) &&
                   ((uintptr_t)z_out % 16 == 0) && ((uintptr_t)z_in % 16 == 0) && ((uintptr_t)x0 % 16 == 0);
  const int V = vec ? 4 : 1;
  const int64_t threads = (total + V - 1) / V;
  const unsigned grid = (unsigned)std::min<int64_t>((threads + 255) / 256, 148 * 16);
  if (vec)
    cfg_euler_kernel<TZ, 4><<<grid, 256, 0, st>>>(pred, (const TZ*)z, dt, g, (const TZ*)x0, keep, z_out, zb, zf,
                                                   per_sample, hw, T, total);
  else
    cfg_euler_kernel<TZ, 1><<<grid, 256, 0, st>>>(pred, (const TZ*)z, dt, g, (const TZ*)x0, keep, z_out, zb, zf,
                                                   per_sample, hw, T, total);
  MDV2_CHECK_LAUNCH("cfg_euler_step");
  return MDV2_OK;
}

extern "C" int mdv2_cfg_euler_step(const float* pred, const void* z, int z_dtype, const float* dt, float guidance,
                                   const void* x0, const uint8_t* keep, float* z_out, void* z_in, int zin_dtype,
                                   int n, int C, int T, int64_t hw, void* stream) {
  if (n == 0 || C == 0 || T == 0 || hw == 0) return MDV2_OK;
  MDV2_REQUIRE(pred != nullptr && z != nullptr && dt != nullptr && z_out != nullptr, "cfg_euler_step: null operand");
  MDV2_REQUIRE((keep == nullptr) == (x0 == nullptr), "cfg_euler_step: keep mask and x0 come together");
  MDV2_REQUIRE(z_in == nullptr || zin_dtype == 0 || zin_dtype == 1, "cfg_euler_step: bad z_in dtype %d", zin_dtype);
  if (z_in == nullptr) zin_dtype = -1;
  cudaStream_t st = (cudaStream_t)stream;
  if (z_dtype == 0) return launch_cfg_euler<float>(pred, z, dt, guidance, x0, keep, z_out, z_in, zin_dtype, n, C, T, hw, st);
  if (z_dtype == 1) return launch_cfg_euler<bf16>(pred, z, dt, guidance, x0, keep, z_out, z_in, zin_dtype, n, C, T, hw, st);
  MDV2_REQUIRE(false, "cfg_euler_step: bad z dtype %d", z_dtype);
}
