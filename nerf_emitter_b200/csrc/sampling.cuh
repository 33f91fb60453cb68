// sampling.cuh - warp-cooperative per-ray proposal loop (levels 0 and 1) shared by the forward and
// backward kernels.  One warp owns one ray; lane L owns samples L, L+32, L+64, ... so that
// neighbouring lanes look up neighbouring points of the same ray (same / adjacent hash cells on the
// coarse levels -> the 8-byte gathers of a warp coalesce into few 32-byte sectors).
//
// Code-size note (ncu, round 1): the first version unrolled the 8 samples per lane and instantiated
// the level twice; 18 k SASS instructions, 91 % of warp stalls were instruction-cache misses
// (stall_no_inst).  The loops over samples are therefore NOT unrolled and both levels run through one
// copy of the code with per-level parameters; per-sample values live in the warp's shared scratch.
//
// Follows ProposalNetworkSampler.generate_ray_samples (NS/model_components/ray_samplers.py:584-626),
// UniformLinDispPiecewiseSampler (:231-254, :82-134), PDFSampler (:282-380),
// HashMLPDensityField.get_density (NS/fields/density_fields.py:95-122) and RaySamples.get_weights
// (NS/cameras/rays.py:128-150).
#pragma once
#include <cuda_fp16.h>
#include <cuda/std/type_traits>
#include "common.cuh"

namespace nemit {

// proposal MLP in shared memory: row j = {W0[j][0..9], b0[j], w1[j]} (12 floats, 16 rows) + b1
struct PropMlpSmem {
  float row[kPropHidden][12];
  float b1;
  float pad[3];
};

__device__ __forceinline__ void load_prop_mlp(PropMlpSmem* dst, const float* __restrict__ w0,
                                              const float* __restrict__ b0, const float* __restrict__ w1,
                                              const float* __restrict__ b1, int tid, int nthreads) {
  for (int i = tid; i < kPropHidden * 12; i += nthreads) {
    int j = i / 12, k = i % 12;
    float v;
    if (k < kPropEnc) v = w0[j * kPropEnc + k];
    else if (k == kPropEnc) v = b0[j];
    else v = w1[j];
    dst->row[j][k] = v;
  }
  if (tid == 0) dst->b1 = b1[0];
}

// hash-encode one point on a 5-level proposal grid: all 40 gathers are issued before the first blend
__device__ __forceinline__ void prop_encode(const float p[3], const float2* __restrict__ table, const GridDev& G,
                                            float enc[kPropEnc]) {
  LevelCell cell[kPropLevels];
  float2 v[kPropLevels][8];
#pragma unroll
  for (int l = 0; l < kPropLevels; ++l) {
    level_cell(p, G.scal[l], G.table_mask, (uint32_t)l * G.table_size, cell[l]);
#pragma unroll
    for (int c = 0; c < 8; ++c) v[l][c] = ldg_f2(table, cell[l].idx[c]);
  }
#pragma unroll
  for (int l = 0; l < kPropLevels; ++l) {
    float2 r = blend_cell(v[l], cell[l]);
    enc[2 * l] = r.x;
    enc[2 * l + 1] = r.y;
  }
}

// raw (pre-activation) density of the proposal MLP
__device__ __forceinline__ float prop_mlp_raw(const float enc[kPropEnc], const PropMlpSmem* W) {
  float raw = W->b1;
#pragma unroll 4
  for (int j = 0; j < kPropHidden; ++j) {
    const float4* r4 = reinterpret_cast<const float4*>(W->row[j]);
    float4 a = r4[0], b = r4[1], c = r4[2];
    float h = c.z;  // b0[j]
    h = fmaf(a.x, enc[0], h); h = fmaf(a.y, enc[1], h); h = fmaf(a.z, enc[2], h); h = fmaf(a.w, enc[3], h);
    h = fmaf(b.x, enc[4], h); h = fmaf(b.y, enc[5], h); h = fmaf(b.z, enc[6], h); h = fmaf(b.w, enc[7], h);
    h = fmaf(c.x, enc[8], h); h = fmaf(c.y, enc[9], h);
    raw = fmaf(c.w, fmaxf(h, 0.f), raw);
  }
  return raw;
}

// density at a sample position (NeRF coordinates): contraction -> hash -> MLP -> trunc_exp * selector
static __device__ __noinline__ float prop_density(float x0, float x1, float x2, const float2* __restrict__ table,
                                           const GridDev& G, const PropMlpSmem* W, const ModelDev& M) {
  float x[3] = {x0, x1, x2}, p[3];
  if (!contract_normalise(x, M, p)) return 0.f;   // reference evaluates at p=0 and multiplies by 0
  float enc[kPropEnc];
  prop_encode(p, table, G, enc);
  return expf(prop_mlp_raw(enc, W));
}

// ---------------------------------------------------------------------------------------------------------
// Fast proposal density (NEMIT_SAMPLER_FAST, the fp16 mode's sampler): same algorithm, packed-half arithmetic.
// The two features of a table entry travel as one half2 through the trilinear blend (7 fused lerps instead of
// 2 x 7 x (2 mul + add)), the 10 -> 16 -> 1 MLP runs on half2 pairs of hidden units (80 HFMA2 instead of 176 FFMA),
// the contraction uses reciprocal approximations.  This is what the reference does under torch.autocast / tcnn
// (fp16 proposal networks); sample indices are then no longer bit-equal to the fp32 CPU path, which is why the
// exact sampler stays the default and the parity mode.
// ---------------------------------------------------------------------------------------------------------
struct PropMlpFastSmem {
  __half2 w0[kPropEnc][8];   // [k][jp] = (W0[2jp][k], W0[2jp+1][k])
  __half2 b0[8];
  __half2 w1[8];
  float b1;
  float pad[3];
};

__device__ __forceinline__ void load_prop_mlp_fast(PropMlpFastSmem* dst, const float* __restrict__ w0,
                                                   const float* __restrict__ b0, const float* __restrict__ w1,
                                                   const float* __restrict__ b1, int tid, int nthreads) {
  for (int i = tid; i < kPropEnc * 8; i += nthreads) {
    int k = i / 8, jp = i % 8;
    dst->w0[k][jp] = __floats2half2_rn(w0[(2 * jp) * kPropEnc + k], w0[(2 * jp + 1) * kPropEnc + k]);
  }
  for (int jp = tid; jp < 8; jp += nthreads) {
    dst->b0[jp] = __floats2half2_rn(b0[2 * jp], b0[2 * jp + 1]);
    dst->w1[jp] = __floats2half2_rn(w1[2 * jp], w1[2 * jp + 1]);
  }
  if (tid == 0) dst->b1 = b1[0];
}

// hash coordinates and fractions as level_cell, without the second rounding of ceil (x1 = x0 + (frac > 0))
__device__ __forceinline__ __half2 lerp_h2(__half2 a_hi, __half2 b_lo, __half2 f) {
  return __hfma2(f, __hsub2(a_hi, b_lo), b_lo);
}
__device__ __forceinline__ __half2 blend_cell_h2(const __half2 h[8], const LevelCell& c) {
  const __half2 fx = __float2half2_rn(c.fx), fy = __float2half2_rn(c.fy), fz = __float2half2_rn(c.fz);
  const __half2 a = lerp_h2(h[0], h[1], fx), b = lerp_h2(h[2], h[3], fx);
  const __half2 cc = lerp_h2(h[4], h[5], fx), d = lerp_h2(h[6], h[7], fx);
  return lerp_h2(lerp_h2(a, b, fy), lerp_h2(cc, d, fy), fz);
}

__device__ __forceinline__ bool contract_normalise_fast(const float x[3], const ModelDev& M, float p[3]) {
  const float mag = fmaxf(fabsf(x[0]), fmaxf(fabsf(x[1]), fabsf(x[2])));
  float f = 1.f;
  if (mag >= 1.f) {
    const float inv = __frcp_rn(mag);
    f = (2.f - inv) * inv;
  }
  bool sel = true;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    p[k] = fmaf(x[k], 0.25f * f, 0.5f);
    sel = sel && (p[k] > 0.f) && (p[k] < 1.f);
  }
  if (M.disable_inside_aabb) {
    bool inside = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) inside = inside && (x[k] > M.aabb_min[k]) && (x[k] < M.aabb_max[k]);
    sel = sel && !inside;
  }
  return sel;
}

// `table` is the packed-half copy of the proposal table (k_prop_tables_half, one __half2 per entry, refreshed by every
// sampler launch): the same cvt.rn the blend applied per corner before, done once per entry instead of 40 times per sample
static __device__ __noinline__ float prop_density_fast(float x0, float x1, float x2, const uint32_t* __restrict__ table,
                                                       const GridDev& G, const PropMlpFastSmem* W, const ModelDev& M) {
  float x[3] = {x0, x1, x2}, p[3];
  if (!contract_normalise_fast(x, M, p)) return 0.f;
  LevelCell cell[kPropLevels];
  __half2 v[kPropLevels][8];
#pragma unroll
  for (int l = 0; l < kPropLevels; ++l) {
    level_cell(p, G.scal[l], G.table_mask, (uint32_t)l * G.table_size, cell[l]);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t u = __ldg(table + cell[l].idx[c]);
      v[l][c] = *reinterpret_cast<const __half2*>(&u);
    }
  }
  __half2 acc[8];
  {
    const uint4* b0 = reinterpret_cast<const uint4*>(W->b0);
    const uint4 t0 = b0[0], t1 = b0[1];
    const uint32_t u[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) acc[jp] = *reinterpret_cast<const __half2*>(&u[jp]);
  }
#pragma unroll
  for (int l = 0; l < kPropLevels; ++l) {
    const __half2 e = blend_cell_h2(v[l], cell[l]);
    const __half2 e0 = __low2half2(e), e1 = __high2half2(e);
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      const uint4* wr = reinterpret_cast<const uint4*>(W->w0[2 * l + f]);
      const uint4 t0 = wr[0], t1 = wr[1];
      const uint32_t u[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
      const __half2 ev = f == 0 ? e0 : e1;
#pragma unroll
      for (int jp = 0; jp < 8; ++jp) acc[jp] = __hfma2(*reinterpret_cast<const __half2*>(&u[jp]), ev, acc[jp]);
    }
  }
  const uint4* w1p = reinterpret_cast<const uint4*>(W->w1);
  const uint4 t0 = w1p[0], t1 = w1p[1];
  const uint32_t u[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
  const __half2 zero = __float2half2_rn(0.f);
  // the hidden -> 1 reduction in fp32: 16 terms of either sign
  float raw = W->b1;
#pragma unroll
  for (int jp = 0; jp < 8; ++jp) {
    const float2 h = __half22float2(__hmax2(acc[jp], zero));
    const float2 w = __half22float2(*reinterpret_cast<const __half2*>(&u[jp]));
    raw = fmaf(h.x, w.x, raw);
    raw = fmaf(h.y, w.y, raw);
  }
  return __expf(raw);
}

// per-warp scratch of the proposal loop
struct WarpScratch {
  float s_edge[kS0 + 4];   // spacing edges of the current level (S+1 used)
  float buf[kS0 + 4];      // delta*sigma -> weights -> padded weights
  float cdf[kS0 + 4];      // S+1 used
  float s_new[kS1 + 4];    // edges produced by the PDF stage (M+1 used)
};

// ATen's CPU float sum over a contiguous row (aten/src/ATen/native/cpu/SumKernel.cpp, row_sum with
// ilp_factor 4 over Vectorized<float> of width V): reproduced so that pdf = w / sum(w) has the same
// bits as torch.sum on the reference's CPU path.  `w` in shared memory, n values.  All lanes return
// the sum.  V == 0 selects a float64 sum rounded once.
__device__ __forceinline__ float row_sum_like_aten(const float* w, int n, int V, int lane) {
  if (V <= 0) {
    double acc = 0.0;
    for (int i = lane; i < n; i += 32) acc += (double)w[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    return (float)acc;
  }
  const int nv = n / V;          // whole vectors
  const int n_ilp = nv / 4;
  float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
  if (lane < V) {
#pragma unroll 1
    for (int i = 0; i < n_ilp; ++i) {
      p0 = add_rn(p0, w[(i * 4 + 0) * V + lane]);
      p1 = add_rn(p1, w[(i * 4 + 1) * V + lane]);
      p2 = add_rn(p2, w[(i * 4 + 2) * V + lane]);
      p3 = add_rn(p3, w[(i * 4 + 3) * V + lane]);
    }
#pragma unroll 1
    for (int i = n_ilp * 4; i < nv; ++i) p0 = add_rn(p0, w[i * V + lane]);
    p0 = add_rn(add_rn(add_rn(p0, p1), p2), p3);
  }
  float acc = 0.f;
#pragma unroll 1
  for (int k = nv * V; k < n; ++k) acc = add_rn(acc, w[k]);   // scalar tail first
#pragma unroll 1
  for (int l = 0; l < V; ++l) acc = add_rn(acc, __shfl_sync(0xffffffffu, p0, l));
  return acc;
}

struct RayCtx {
  float o[3], d[3];
  float s_near, s_far;
};

// One proposal level: density at the S bin midpoints -> weights (in sc->buf[0..S)).
// Edges must be in sc->s_edge[0..S].  S is a multiple of 32.
template <class MLP = PropMlpSmem>
__device__ __forceinline__ void proposal_level_weights(WarpScratch* sc, const RayCtx& R, int S,
                                                       const float2* __restrict__ table, const GridDev& G,
                                                       const MLP* W, const ModelDev& M, int lane,
                                                       float* ds_out = nullptr, const uint32_t* __restrict__ table_h = nullptr,
                                                       const RotArgs* Q = nullptr) {
  constexpr bool kFast = !cuda::std::is_same<MLP, PropMlpSmem>::value;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float t0 = s_to_t(sc->s_edge[i], R.s_near, R.s_far);
    float t1 = s_to_t(sc->s_edge[i + 1], R.s_near, R.s_far);
    float tsum = add_rn(t0, t1);
    float sigma;
    float x[3], dir[3];
    if (Q) rot_sample(*Q, R.o, R.d, tsum, x, dir);      // Rotater.apply_frustums: every level's samples (rays.py:285-286)
    else {
#pragma unroll
      for (int k = 0; k < 3; ++k) x[k] = sample_pos(R.o[k], R.d[k], tsum);
    }
    if constexpr (kFast)
      sigma = prop_density_fast(x[0], x[1], x[2], table_h, G, W, M);
    else
      sigma = prop_density(x[0], x[1], x[2], table, G, W, M);
    sc->buf[i] = mul_rn(sub_rn(t1, t0), sigma);
    if (ds_out) ds_out[i] = sc->buf[i];
  }
  __syncwarp();
  // exclusive prefix accumulated in float64 and rounded to float32 per prefix (ATen's CPU cumsum)
  double carry = 0.0;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float ds = sc->buf[i];
    double inc = warp_incl_scan((double)ds, lane);
    double exc = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) exc = 0.0;
    float prefix = (float)(carry + exc);
    carry += __shfl_sync(0xffffffffu, inc, 31);
    float alpha = sub_rn(1.f, expf(-ds));
    sc->buf[i] = nan_to_num(mul_rn(alpha, expf(-prefix)));
  }
  __syncwarp();
}

// PDFSampler.generate_ray_samples for one ray: S existing bins (edges in sc->s_edge, weights in
// sc->buf), M new samples -> M+1 new edges into sc->s_new.  u_base = linspace(0,1-1/nb,nb) from the
// host.  jit < 0 means deterministic.  inds_out (M+1) optional.
__device__ __forceinline__ void pdf_resample_warp(WarpScratch* sc, int S, int M, const float* __restrict__ u_base,
                                                  float jit, float anneal, int sum_width, int lane,
                                                  int32_t* inds_out) {
  const int nb = M + 1;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float x = sc->buf[i];
    if (anneal != 1.f) x = powf(x, anneal);
    sc->buf[i] = add_rn(x, 0.01f);
  }
  __syncwarp();
  float tot = row_sum_like_aten(sc->buf, S, sum_width, lane);
  const float fill = fmaxf(sub_rn(1e-5f, tot), 0.f);
  const float fill_each = div_rn(fill, (float)S);
  tot = add_rn(tot, fill);
  // cdf = [0, min(1, cumsum(pdf))]: float64 accumulate, each prefix rounded to float32
  double carry = 0.0;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float pdf = div_rn(add_rn(sc->buf[i], fill_each), tot);
    double inc = warp_incl_scan((double)pdf, lane);
    sc->cdf[i + 1] = fminf(1.f, (float)(carry + inc));
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) sc->cdf[0] = 0.f;
  __syncwarp();
  const float half_step = (float)(1.0 / (2.0 * (double)nb));
#pragma unroll 1
  for (int k = lane; k <= M; k += 32) {
    float u = (jit >= 0.f) ? add_rn(u_base[k], div_rn(jit, (float)nb)) : add_rn(u_base[k], half_step);
    int lo = 0, hi = S + 1;          // searchsorted(cdf, u, side="right") over S+1 entries
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (sc->cdf[mid] <= u) lo = mid + 1; else hi = mid;
    }
    if (inds_out) inds_out[k] = lo;
    int i0 = min(max(lo - 1, 0), S), i1 = min(max(lo, 0), S);
    float c0 = sc->cdf[i0], c1 = sc->cdf[i1];
    float e0 = sc->s_edge[i0], e1 = sc->s_edge[i1];
    float t = div_rn(sub_rn(u, c0), sub_rn(c1, c0));
    t = isnan(t) ? 0.f : t;                        // nan_to_num(.., 0) then clip
    t = fminf(fmaxf(t, 0.f), 1.f);
    sc->s_new[k] = add_rn(e0, mul_rn(t, sub_rn(e1, e0)));
  }
  __syncwarp();
}

// initial edges (ray_samplers.py:104-117): lin0 = linspace(0,1,S0+1) from the host; with jitter the
// edges move inside [midpoint below, midpoint above].
__device__ __forceinline__ void initial_edges(WarpScratch* sc, const float* __restrict__ lin0, float jit, int lane) {
#pragma unroll 1
  for (int i = lane; i <= kS0; i += 32) {
    float e = lin0[i];
    if (jit >= 0.f) {
      float lo = (i == 0) ? lin0[0] : div_rn(add_rn(lin0[i], lin0[i - 1]), 2.f);
      float hi = (i == kS0) ? lin0[kS0] : div_rn(add_rn(lin0[i + 1], lin0[i]), 2.f);
      e = add_rn(lo, mul_rn(sub_rn(hi, lo), jit));
    }
    sc->s_edge[i] = e;
  }
  __syncwarp();
}

}  // namespace nemit
