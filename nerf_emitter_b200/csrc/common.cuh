// common.cuh - shared device helpers of libnerfemitter (sm_100a).
//
// Arithmetic that feeds the PDF resampling (spacing maps, sample positions, contraction, hash
// coordinates) is written with explicit round-to-nearest intrinsics (__fmul_rn / __fadd_rn /
// __fdiv_rn): nvcc must not contract a*b+c into an FMA there, because the parity gate asks for
// searchsorted indices that are bit-equal to the reference's torch-CPU arithmetic, which rounds
// after every operation.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/nemit.h"

#define NEMIT_PRIME_Y 2654435761u
#define NEMIT_PRIME_Z 805459861u

namespace nemit {

constexpr int kS0 = 256;         // proposal samples level 0
constexpr int kS1 = 96;          // proposal samples level 1
constexpr int kS2 = 48;          // nerf samples
constexpr int kPropLevels = 5;
constexpr int kPropEnc = 10;     // 5 levels x 2 features
constexpr int kPropHidden = 16;
constexpr int kMainLevels = 16;
constexpr int kMainEnc = 32;
constexpr int kHidden = 64;
constexpr int kGeo = 15;
constexpr int kApp = 32;
constexpr int kAppMean = -1;   // RayArgs::app_row: mean of the embedding rows
constexpr int kAppZero = -2;   // RayArgs::app_row: zero appearance vector
constexpr int kHeadIn = 63;      // 16 SH + 15 geo + 32 appearance

struct GridDev {
  int num_levels;
  uint32_t table_mask;           // T-1 (T is a power of two)
  uint32_t table_size;
  float scal[NEMIT_MAX_LEVELS];
};

// Everything a kernel needs that is constant for a handle; passed by value (__grid_constant__).
struct ModelDev {
  GridDev main_grid;
  GridDev prop_grid[2];
  float aabb_min[3], aabb_max[3];
  float scene_scale;
  float near_plane, far_plane;
  float rgb_bias;
  int disable_inside_aabb;
  int cpu_sum_width;
  int num_images;
  float interlevel_mult, distortion_mult;
};

// ---------------------------------------------------------------------------------------------
// exact (non-contracted) fp32 helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }

// NS/model_components/ray_samplers.py:250  spacing_fn
__device__ __forceinline__ float spacing_fwd(float x) {
  return x < 1.f ? div_rn(x, 2.f) : sub_rn(1.f, div_rn(1.f, mul_rn(2.f, x)));
}
// ray_samplers.py:251  spacing_fn_inv
__device__ __forceinline__ float spacing_inv(float s) {
  return s < 0.5f ? mul_rn(2.f, s) : div_rn(1.f, sub_rn(2.f, mul_rn(2.f, s)));
}
// ray_samplers.py:121-122  spacing_to_euclidean_fn: f_inv(x*s_far + (1-x)*s_near)
__device__ __forceinline__ float s_to_t(float s, float s_near, float s_far) {
  return spacing_inv(add_rn(mul_rn(s, s_far), mul_rn(sub_rn(1.f, s), s_near)));
}

// Mitsuba unit cube -> NeRF coordinates (DS/emitters/nerf_op.py:76-89, NS/utils/mi_gl_conversion.py:4-9,40-42):
// o_mi' = o + near*v ; x = (o_mi'*2-1)*scale ; (x,y,z) -> (z,x,y).
__device__ __forceinline__ void load_ray(const float* __restrict__ o, const float* __restrict__ v,
                                         const float* __restrict__ near_per_ray, float near_scalar,
                                         int64_t n, int64_t r, float scene_scale, float oo[3], float dd[3]) {
  float nr = near_per_ray ? near_per_ray[r] : near_scalar;
  // scene_scale == 0 selects NEMIT_RAYS_NERF_N3 (kernels.h ray_scale): RayBundle.origins / .directions, (N,3) rows already
  // in NeRF coordinates.  Both layouts issue the same six loads (strides differ), so no load waits behind a branch.
  const bool nerf = scene_scale == 0.f;
  const int64_t sr = nerf ? 3 : 1, sk = nerf ? 1 : n;
  float vm[3] = {v[r * sr], v[r * sr + sk], v[r * sr + 2 * sk]};
  float x[3] = {o[r * sr], o[r * sr + sk], o[r * sr + 2 * sk]};
  float om[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    x[k] = add_rn(x[k], mul_rn(nr, vm[k]));
    om[k] = mul_rn(sub_rn(mul_rn(x[k], 2.f), 1.f), scene_scale);
  }
  // matmul with the permutation matrix adds exact zeros; the + 0 translation is exact as well
  oo[0] = nerf ? x[0] : om[2]; oo[1] = nerf ? x[1] : om[0]; oo[2] = nerf ? x[2] : om[1];
  dd[0] = nerf ? vm[0] : vm[2]; dd[1] = nerf ? vm[1] : vm[0]; dd[2] = nerf ? vm[2] : vm[1];
}

// NS/cameras/rays.py:49-58: pos = o + d*(t0+t1)/2 (evaluated as ((d*(t0+t1))/2) + o)
__device__ __forceinline__ float sample_pos(float o, float d, float tsum) {
  return add_rn(o, mul_rn(mul_rn(d, tsum), 0.5f));
}

// Rotater.apply_frustums (NS/field_components/rotater.py:125-149, hooked into RayBundle.get_ray_samples at
// NS/cameras/rays.py:285-286): per SAMPLE, if the sample's unrotated position lies inside the rotation bound (sphere around
// the scene centre, or an axis-aligned box), the sample takes the rotated ray's origin and direction
// (batch_affine_left, NS/utils/mi_gl_conversion.py:69-73).  Passed by value inside the kernels' argument blocks.
struct RotArgs {
  int mode;        // 0 = no rotation, 1 = bounding sphere, 2 = axis-aligned box
  float m[12];     // 3 x 4 row-major [R | t] of the call's rotation, NeRF coordinates (Rotater.torch_transforms[id, :3, :4])
  float b[6];      // sphere: centre xyz, radius; box: min xyz, max xyz (inclusive, rotater.py:129-130)
};

// position x and direction dir of the sample at t = tsum / 2 on the ray (o, d); returns true when it was rotated
__device__ __forceinline__ bool rot_sample(const RotArgs& Q, const float o[3], const float d[3], float tsum, float x[3],
                                           float dir[3]) {
#pragma unroll
  for (int k = 0; k < 3; ++k) { x[k] = add_rn(o[k], mul_rn(mul_rn(d[k], tsum), 0.5f)); dir[k] = d[k]; }
  if (Q.mode == 0) return false;
  bool inside;
  if (Q.mode == 1) {
    const float a = x[0] - Q.b[0], b = x[1] - Q.b[1], c = x[2] - Q.b[2];
    inside = sqrtf(a * a + b * b + c * c) < Q.b[3];
  } else {
    inside = Q.b[0] <= x[0] && x[0] <= Q.b[3] && Q.b[1] <= x[1] && x[1] <= Q.b[4] && Q.b[2] <= x[2] && x[2] <= Q.b[5];
  }
  if (!inside) return false;
  float o2[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    o2[i] = Q.m[4 * i] * o[0] + Q.m[4 * i + 1] * o[1] + Q.m[4 * i + 2] * o[2] + Q.m[4 * i + 3];
    dir[i] = Q.m[4 * i] * d[0] + Q.m[4 * i + 1] * d[1] + Q.m[4 * i + 2] * d[2];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) x[k] = add_rn(o2[k], mul_rn(mul_rn(dir[k], tsum), 0.5f));
  return true;
}
// transpose of the rotation: cotangent of a rotated sample's origin / direction back to the ray's
__device__ __forceinline__ void rot_transpose(const RotArgs& Q, const float g[3], float out[3]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) out[i] = Q.m[i] * g[0] + Q.m[4 + i] * g[1] + Q.m[8 + i] * g[2];
}

// SceneContraction(order=inf) + (x+2)/4 + selector (NS/field_components/spatial_distortions.py:67-70,
// NS/fields/density_fields.py:96-106, NS/data/scene_box.py:36-37). Returns selector; p is zeroed when
// the selector is false (density_fields.py:106).
__device__ __forceinline__ bool contract_normalise(const float x[3], const ModelDev& M, float p[3]) {
  float mag = fmaxf(fabsf(x[0]), fmaxf(fabsf(x[1]), fabsf(x[2])));
  bool sel = true;
  if (mag < 1.f) {
#pragma unroll
    for (int k = 0; k < 3; ++k) p[k] = x[k];
  } else {
    float f = sub_rn(2.f, div_rn(1.f, mag));
#pragma unroll
    for (int k = 0; k < 3; ++k) p[k] = mul_rn(f, div_rn(x[k], mag));
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    p[k] = div_rn(add_rn(p[k], 2.f), 4.f);
    sel = sel && (p[k] > 0.f) && (p[k] < 1.f);
  }
  if (M.disable_inside_aabb) {
    bool inside = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) inside = inside && (x[k] > M.aabb_min[k]) && (x[k] < M.aabb_max[k]);
    sel = sel && !inside;
  }
  if (!sel) { p[0] = 0.f; p[1] = 0.f; p[2] = 0.f; }
  return sel;
}

// NS/field_components/encodings.py:322-339: (x*1) ^ (y*2654435761) ^ (z*805459861) mod T. The
// reference multiplies in int64; for T a power of two <= 2^32 the low bits equal the uint32 product.
__device__ __forceinline__ uint32_t hash3(uint32_t x, uint32_t y, uint32_t z, uint32_t mask) {
  return (x ^ (y * NEMIT_PRIME_Y) ^ (z * NEMIT_PRIME_Z)) & mask;
}

// One level of the torch-path hash grid (encodings.py:341-382): 8 gathers + trilinear blend in the
// reference's order (ceil corner weighted by the offset, floor by 1-offset; x, then y, then z).
struct LevelCell {
  uint32_t idx[8];   // order: (cx,cy,cz) (fx,cy,cz) (cx,fy,cz) (fx,fy,cz) (cx,cy,fz) (fx,cy,fz) (cx,fy,fz) (fx,fy,fz)
  float fx, fy, fz;
};

__device__ __forceinline__ void floor_ceil_small(float x, float& fl, uint32_t& i0, uint32_t& i1) {
  const float magic = 12582912.f;                               // 1.5 * 2^23
  const float t = add_rn(x, magic);
  float r = sub_rn(t, magic);                                   // round-to-nearest-even(x), exact
  int ri = __float_as_int(t) - 0x4B400000;
  const bool up = r > x;
  fl = up ? r - 1.f : r;
  i0 = (uint32_t)(up ? ri - 1 : ri);
  i1 = i0 + (x > fl ? 1u : 0u);                                 // ceil == floor on integers
}

// ALU_FLOOR: floor / ceil / float->int through floor_ceil_small instead of the conversion unit.  Measured: the field
// kernels (XU 13 %, ALU 17 % busy) gain 3 % (3.37 -> 3.28 ms per 2^18 rays), the sampler (ALU 42 % busy) loses 6 %,
// so only the field kernels ask for it.  Both variants give bit-identical cells.
// (cell0, when given, receives the integer coordinates of the cell's floor corner)
template <bool ALU_FLOOR = false>
__device__ __forceinline__ void level_cell(const float p[3], float scale, uint32_t mask, uint32_t level_off,
                                           LevelCell& c, uint32_t* cell0 = nullptr) {
  float xs = mul_rn(p[0], scale), ys = mul_rn(p[1], scale), zs = mul_rn(p[2], scale);
  float xf, yf, zf;
  uint32_t x0, y0, z0, x1, y1, z1;
  if (!ALU_FLOOR) {
    xf = floorf(xs); yf = floorf(ys); zf = floorf(zs);
    x0 = (uint32_t)(int)xf; y0 = (uint32_t)(int)yf; z0 = (uint32_t)(int)zf;
    x1 = (uint32_t)(int)ceilf(xs); y1 = (uint32_t)(int)ceilf(ys); z1 = (uint32_t)(int)ceilf(zs);
  } else {
    // without the conversion unit (4 XU operations per axis otherwise): 0 <= xs < 2^22 here (positions are normalised to
    // [0,1], resolutions <= 2048), so xs + 1.5 * 2^23 holds round-to-nearest(xs) in its low mantissa bits; step down when
    // that rounded up.  Bit-identical to floorf / ceilf (checked exhaustively over every float in [0, 4096]).
    floor_ceil_small(xs, xf, x0, x1); floor_ceil_small(ys, yf, y0, y1); floor_ceil_small(zs, zf, z0, z1);
  }
  c.fx = sub_rn(xs, xf); c.fy = sub_rn(ys, yf); c.fz = sub_rn(zs, zf);
  if (cell0) { cell0[0] = x0; cell0[1] = y0; cell0[2] = z0; }
  uint32_t hx0 = x0, hx1 = x1;
  uint32_t hy0 = y0 * NEMIT_PRIME_Y, hy1 = y1 * NEMIT_PRIME_Y;
  uint32_t hz0 = z0 * NEMIT_PRIME_Z, hz1 = z1 * NEMIT_PRIME_Z;
  c.idx[0] = ((hx1 ^ hy1 ^ hz1) & mask) + level_off;
  c.idx[1] = ((hx0 ^ hy1 ^ hz1) & mask) + level_off;
  c.idx[2] = ((hx1 ^ hy0 ^ hz1) & mask) + level_off;
  c.idx[3] = ((hx0 ^ hy0 ^ hz1) & mask) + level_off;
  c.idx[4] = ((hx1 ^ hy1 ^ hz0) & mask) + level_off;
  c.idx[5] = ((hx0 ^ hy1 ^ hz0) & mask) + level_off;
  c.idx[6] = ((hx1 ^ hy0 ^ hz0) & mask) + level_off;
  c.idx[7] = ((hx0 ^ hy0 ^ hz0) & mask) + level_off;
}

__device__ __forceinline__ float lerp_ref(float a_hi, float b_lo, float f) {
  // a*f + b*(1-f), each product and the sum rounded (torch evaluates them as separate ops)
  return add_rn(mul_rn(a_hi, f), mul_rn(b_lo, sub_rn(1.f, f)));
}

__device__ __forceinline__ float2 blend_cell(const float2 v[8], const LevelCell& c) {
  float2 r;
  {
    float a = lerp_ref(v[0].x, v[1].x, c.fx), b = lerp_ref(v[2].x, v[3].x, c.fx);
    float cc = lerp_ref(v[4].x, v[5].x, c.fx), d = lerp_ref(v[6].x, v[7].x, c.fx);
    r.x = lerp_ref(lerp_ref(a, b, c.fy), lerp_ref(cc, d, c.fy), c.fz);
  }
  {
    float a = lerp_ref(v[0].y, v[1].y, c.fx), b = lerp_ref(v[2].y, v[3].y, c.fx);
    float cc = lerp_ref(v[4].y, v[5].y, c.fx), d = lerp_ref(v[6].y, v[7].y, c.fx);
    r.y = lerp_ref(lerp_ref(a, b, c.fy), lerp_ref(cc, d, c.fy), c.fz);
  }
  return r;
}

__device__ __forceinline__ float2 ldg_f2(const float2* __restrict__ t, uint32_t i) { return __ldg(t + i); }

// The 8 corner entries of one cell.  The x-neighbours (ceil x, floor x) of a (y, z) pair hash to indices that differ
// in bit 0 only whenever floor x is even, i.e. they are one aligned 16-byte pair of the table: those lanes fetch the pair
// with ONE 128-bit load, the others with two 64-bit loads, predicated per lane into the same four registers.  An
// L1-missing lane-load costs 1.0 SM-cycle whatever its width and an L1-hitting one 0.34 (tools/gather_bench.cu,
// profiles/r1_gather_bench.txt), so this removes a quarter of the pair cost on half of the pairs.
__device__ __forceinline__ void gather_cell(const float2* __restrict__ t, const LevelCell& c, float2 v[8]) {
#ifdef NEMIT_PLAIN_GATHER
#pragma unroll
  for (int q = 0; q < 8; ++q) v[q] = __ldg(t + c.idx[q]);
  return;
#endif
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t ic = c.idx[2 * q], il = c.idx[2 * q + 1];       // ceil-x corner, floor-x corner
    const uint32_t pair = (ic ^ il) == 1u;
    float2 a, b;                 // pair: (a, b) = (even entry, odd entry); otherwise (floor corner, ceil corner)
    if (pair) {
      const float4 w = __ldg(reinterpret_cast<const float4*>(t + (il & ~1u)));
      a = make_float2(w.x, w.y); b = make_float2(w.z, w.w);
    } else {
      a = __ldg(t + il); b = __ldg(t + ic);
    }
    const float a0 = a.x, a1 = a.y, b0 = b.x, b1 = b.y;
    const bool swap = pair && (il & 1u);
    v[2 * q + 1] = swap ? make_float2(b0, b1) : make_float2(a0, a1);
    v[2 * q] = swap ? make_float2(a0, a1) : make_float2(b0, b1);
  }
}

// torch.nan_to_num defaults: nan -> 0, +inf -> FLT_MAX, -inf -> -FLT_MAX
__device__ __forceinline__ float nan_to_num(float x) {
  if (isnan(x)) return 0.f;
  if (isinf(x)) return x > 0.f ? 3.4028234663852886e38f : -3.4028234663852886e38f;
  return x;
}

__device__ __forceinline__ double warp_incl_scan(double v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    double t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// weights of one level from the per-sample delta*sigma held in registers (sample i = k*32 + lane).
// Exclusive prefix accumulated in float64 and rounded to float32 per prefix, as ATen's CPU cumsum does.
template <int K>
__device__ __forceinline__ void weights_from_ds(const float (&ds)[K], float (&w)[K], int lane) {
  double carry = 0.0;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double inc = warp_incl_scan((double)ds[k], lane);
    double exc = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) exc = 0.0;
    float prefix = (float)(carry + exc);
    carry += __shfl_sync(0xffffffffu, inc, 31);
    float alpha = sub_rn(1.f, expf(-ds[k]));
    w[k] = nan_to_num(mul_rn(alpha, expf(-prefix)));
  }
}

// ---------------------------------------------------------------------------------------------
// derivatives (jvp / backward kernels)
// ---------------------------------------------------------------------------------------------
// d(blend_cell)/d(offset): the reference's a*f + b*(1-f) differentiates to (a - b) per axis.
__device__ __forceinline__ void blend_cell_grad(const float2 v[8], const LevelCell& c, float2& gx, float2& gy, float2& gz) {
  const float fx = c.fx, fy = c.fy, fz = c.fz, ux = 1.f - fx, uy = 1.f - fy, uz = 1.f - fz;
#define NEMIT_BCG(F)                                                                              \
  {                                                                                               \
    float a = v[0].F * fx + v[1].F * ux, b = v[2].F * fx + v[3].F * ux;                           \
    float cc = v[4].F * fx + v[5].F * ux, d = v[6].F * fx + v[7].F * ux;                          \
    gz.F = (a * fy + b * uy) - (cc * fy + d * uy);                                                \
    gy.F = fz * (a - b) + uz * (cc - d);                                                          \
    gx.F = fz * (fy * (v[0].F - v[1].F) + uy * (v[2].F - v[3].F)) +                               \
           uz * (fy * (v[4].F - v[5].F) + uy * (v[6].F - v[7].F));                                \
  }
  NEMIT_BCG(x)
  NEMIT_BCG(y)
#undef NEMIT_BCG
}

// trilinear corner weights in the LevelCell corner order
__device__ __forceinline__ void corner_weights(const LevelCell& c, float w[8]) {
  const float fx = c.fx, fy = c.fy, fz = c.fz, ux = 1.f - fx, uy = 1.f - fy, uz = 1.f - fz;
  w[0] = fx * fy * fz; w[1] = ux * fy * fz; w[2] = fx * uy * fz; w[3] = ux * uy * fz;
  w[4] = fx * fy * uz; w[5] = ux * fy * uz; w[6] = fx * uy * uz; w[7] = ux * uy * uz;
}

// Contraction + (x+2)/4 + selector: vector-Jacobian product g_p -> g_x (spatial_distortions.py:67-70).
// x' = x (m<1) else f(m) x, f = (2-1/m)/m, m = |x_j| the max component.
__device__ __forceinline__ void contract_vjp(const float x[3], bool sel, const float gp[3], float gx[3]) {
  if (!sel) { gx[0] = gx[1] = gx[2] = 0.f; return; }
  float a0 = fabsf(x[0]), a1 = fabsf(x[1]), a2 = fabsf(x[2]);
  float m = fmaxf(a0, fmaxf(a1, a2));
  if (m < 1.f) {
#pragma unroll
    for (int k = 0; k < 3; ++k) gx[k] = 0.25f * gp[k];
    return;
  }
  int j = (a0 >= a1 && a0 >= a2) ? 0 : (a1 >= a2 ? 1 : 2);
  float inv = 1.f / m;
  float f = (2.f - inv) * inv;
  float fp = 2.f * inv * inv * (inv - 1.f);
  float dot = gp[0] * x[0] + gp[1] * x[1] + gp[2] * x[2];
#pragma unroll
  for (int k = 0; k < 3; ++k) gx[k] = 0.25f * f * gp[k];
  gx[j] += 0.25f * dot * fp * (x[j] < 0.f ? -1.f : 1.f);
}

// forward-mode counterpart: tangent of p given the tangent of x
__device__ __forceinline__ void contract_jvp(const float x[3], bool sel, const float dx[3], float dp[3]) {
  if (!sel) { dp[0] = dp[1] = dp[2] = 0.f; return; }
  float a0 = fabsf(x[0]), a1 = fabsf(x[1]), a2 = fabsf(x[2]);
  float m = fmaxf(a0, fmaxf(a1, a2));
  if (m < 1.f) {
#pragma unroll
    for (int k = 0; k < 3; ++k) dp[k] = 0.25f * dx[k];
    return;
  }
  int j = (a0 >= a1 && a0 >= a2) ? 0 : (a1 >= a2 ? 1 : 2);
  float inv = 1.f / m;
  float f = (2.f - inv) * inv;
  float fp = 2.f * inv * inv * (inv - 1.f);
  float dm = (x[j] < 0.f ? -1.f : 1.f) * dx[j];
#pragma unroll
  for (int k = 0; k < 3; ++k) dp[k] = 0.25f * (f * dx[k] + x[k] * fp * dm);
}

__device__ __forceinline__ float warp_incl_scan_f(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// like weights_from_ds but also returns the transmittance T and e = exp(-ds) (alpha = 1 - e)
template <int K>
__device__ __forceinline__ void weights_full(const float (&ds)[K], float (&w)[K], float (&T)[K], float (&e)[K], int lane) {
  double carry = 0.0;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double inc = warp_incl_scan((double)ds[k], lane);
    double exc = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) exc = 0.0;
    float prefix = (float)(carry + exc);
    carry += __shfl_sync(0xffffffffu, inc, 31);
    e[k] = expf(-ds[k]);
    T[k] = expf(-prefix);
    w[k] = nan_to_num(mul_rn(sub_rn(1.f, e[k]), T[k]));
  }
}

// float2 reduction into a gradient table (red.global.add.v2.f32 on sm_90+)
__device__ __forceinline__ void atomic_add_f2(float2* addr, float2 v) { atomicAdd(addr, v); }

}  // namespace nemit
