// field_common.cuh - pieces shared by the main-field kernels (forward, jvp, backward):
// 16-level hash encoding of one point, SH16, the shared-memory register-tiled fp32 GEMM used by the
// exact (fp32) MLP path, and the per-ray compositing scan.
//
// Reference: NerfactoField.get_density / get_outputs (NS/fields/nerfacto_field.py:228-334),
// components_from_spherical_harmonics (NS/utils/math.py:27-90), RaySamples.get_weights
// (NS/cameras/rays.py:128-150), RGBRenderer (NS/model_components/renderers.py:71-121,203-234).
#pragma once
#include "kernels.h"

namespace nemit {

constexpr int kTileM = 128;                 // samples per MLP tile
constexpr int kRaysPerGroup = 8;            // 8 rays x 48 samples = 384 = 3 tiles
constexpr int kGroupSamples = kRaysPerGroup * kS2;
constexpr int kHeadK = 31;                  // SH16 + geo15 (appearance is folded into the bias)

// 16-level encode of one point; results go to column m of a k-major tile X[k][kTileM]
__device__ __forceinline__ void main_encode_to_tile(const float p[3], const float2* __restrict__ table,
                                                    const GridDev& G, float* X, int m) {
#pragma unroll 1
  for (int l0 = 0; l0 < kMainLevels; l0 += 4) {
    LevelCell cell[4];
    float2 v[4][8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      level_cell(p, G.scal[l0 + j], G.table_mask, (uint32_t)(l0 + j) * G.table_size, cell[j]);
#pragma unroll
      for (int c = 0; c < 8; ++c) v[j][c] = ldg_f2(table, cell[j].idx[c]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 r = blend_cell(v[j], cell[j]);
      X[(2 * (l0 + j)) * kTileM + m] = r.x;
      X[(2 * (l0 + j) + 1) * kTileM + m] = r.y;
    }
  }
}

// NS/utils/math.py:52-76, levels = 4
__device__ __forceinline__ void sh16(float x, float y, float z, float c[16]) {
  float xx = x * x, yy = y * y, zz = z * z;
  c[0] = 0.28209479177387814f;
  c[1] = 0.4886025119029199f * y;
  c[2] = 0.4886025119029199f * z;
  c[3] = 0.4886025119029199f * x;
  c[4] = 1.0925484305920792f * x * y;
  c[5] = 1.0925484305920792f * y * z;
  c[6] = 0.9461746957575601f * zz - 0.31539156525251999f;
  c[7] = 1.0925484305920792f * x * z;
  c[8] = 0.5462742152960396f * (xx - yy);
  c[9] = 0.5900435899266435f * y * (3.f * xx - yy);
  c[10] = 2.890611442640554f * x * y * z;
  c[11] = 0.4570457994644658f * y * (5.f * zz - 1.f);
  c[12] = 0.3731763325901154f * z * (5.f * zz - 3.f);
  c[13] = 0.4570457994644658f * x * (5.f * zz - 1.f);
  c[14] = 1.445305721320277f * z * (xx - yy);
  c[15] = 0.5900435899266435f * x * (xx - 3.f * yy);
}

// Y[n][m] = act( bias[n] + sum_k Wt[k][n] * X[k][m] ),  m < 128, n < N.  X, Y k-major tiles in shared
// memory (row stride kTileM), Wt[k][N] in shared memory.  128 threads; thread (tm, tn) owns an
// 8 x TN register tile: 16 FMAs per 128-bit shared load.
template <int K, int N, bool RELU>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ X, const float* __restrict__ Wt,
                                          const float* __restrict__ bias, float* __restrict__ Y, int tid) {
  constexpr int TN = N / 8;            // 8 thread columns
  static_assert(N % 8 == 0 && (TN == 8 || TN == 2), "tile shape");
  const int tm = tid & 15, tn = tid >> 4;
  float acc[8][TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) {
    float b = bias ? bias[tn * TN + j] : 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i][j] = b;
  }
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(X + k * kTileM + tm * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(X + k * kTileM + tm * 8 + 4);
    float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float w[TN];
    if constexpr (TN == 8) {
      const float4 w0 = *reinterpret_cast<const float4*>(Wt + k * N + tn * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(Wt + k * N + tn * 8 + 4);
      w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
    } else {
      const float2 w0 = *reinterpret_cast<const float2*>(Wt + k * N + tn * 2);
      w[0] = w0.x; w[1] = w0.y;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
  }
#pragma unroll
  for (int j = 0; j < TN; ++j) {
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = RELU ? fmaxf(acc[i][j], 0.f) : acc[i][j];
    float* dst = Y + (tn * TN + j) * kTileM + tm * 8;
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// MLP weights staged in shared memory, transposed to k-major
struct FieldWeightsSmem {
  float base_w0[kMainEnc * kHidden];     // [32][64]
  float base_w1[kHidden * 16];           // [64][16]
  float head_w0[32 * kHidden];           // [31][64] (+1 pad row): SH rows 0..15, geo rows 16..30
  float head_w1[kHidden * kHidden];      // [64][64]
  float head_w2[kHidden * 4];            // [64][4] (3 used)
  float base_b0[kHidden];
  float base_b1[16];
  float head_b0_eff[kHidden];            // b0 + W0[:, 31:63] . appearance
  float head_b1[kHidden];
  float head_b2[4];
};

// effective head bias = b0 + W0[:,31:63] . e, with e = embedding row (training) or the mean row (eval)
// (NS/fields/nerfacto_field.py:272-285). Computed per CTA: 64 x 32 FMAs.
// Component i of the appearance vector every sample of the call sees (NS/fields/nerfacto_field.py:272-285):
// app_row >= 0: that embedding row (training mode, or always_use_camera_indices with a Rotater, NS/models/nerfacto.py:206);
// kAppMean: the mean over rows (eval, use_average_appearance_embedding); kAppZero: zeros.  The host resolves the row
// (RayArgs::app_row, api.cu appearance_row).
__device__ __forceinline__ float appearance_component(const float* __restrict__ emb, int num_images, int app_row, int i) {
  if (app_row >= 0) return emb[(int64_t)app_row * kApp + i];
  if (app_row == kAppZero) return 0.f;
  float s = 0.f;
  for (int r = 0; r < num_images; ++r) s += emb[(int64_t)r * kApp + i];
  return s / (float)num_images;
}
// its adjoint: g = d loss / d component i
__device__ __forceinline__ void appearance_component_grad(float* __restrict__ gemb, int num_images, int app_row, int i, float g) {
  if (app_row >= 0) { atomicAdd(gemb + (int64_t)app_row * kApp + i, g); return; }
  if (app_row == kAppZero) return;
  for (int r = 0; r < num_images; ++r) atomicAdd(gemb + (int64_t)r * kApp + i, g / (float)num_images);
}

__device__ __forceinline__ void load_field_weights(FieldWeightsSmem* S, const FieldParams& P, int num_images,
                                                   int app_row, float* app_tmp /*32 floats smem*/,
                                                   int tid, int nthreads) {
  for (int i = tid; i < kHidden * kMainEnc; i += nthreads) { int n = i / kMainEnc, k = i % kMainEnc; S->base_w0[k * kHidden + n] = P.base_w0[i]; }
  for (int i = tid; i < 16 * kHidden; i += nthreads) { int n = i / kHidden, k = i % kHidden; S->base_w1[k * 16 + n] = P.base_w1[i]; }
  for (int i = tid; i < kHidden * kHeadIn; i += nthreads) {
    int n = i / kHeadIn, k = i % kHeadIn;
    if (k < kHeadK) S->head_w0[k * kHidden + n] = P.head_w0[i];
  }
  for (int i = tid; i < kHidden; i += nthreads) S->head_w0[kHeadK * kHidden + i] = 0.f;
  for (int i = tid; i < kHidden * kHidden; i += nthreads) { int n = i / kHidden, k = i % kHidden; S->head_w1[k * kHidden + n] = P.head_w1[i]; }
  for (int i = tid; i < kHidden * 4; i += nthreads) { int k = i / 4, n = i % 4; S->head_w2[i] = n < 3 ? P.head_w2[n * kHidden + k] : 0.f; }
  for (int i = tid; i < kHidden; i += nthreads) { S->base_b0[i] = P.base_b0[i]; S->head_b1[i] = P.head_b1[i]; }
  for (int i = tid; i < 16; i += nthreads) S->base_b1[i] = P.base_b1[i];
  for (int i = tid; i < 4; i += nthreads) S->head_b2[i] = i < 3 ? P.head_b2[i] : 0.f;
  for (int i = tid; i < kApp; i += nthreads) {
    app_tmp[i] = appearance_component(P.appearance, num_images, app_row, i);
  }
  __syncthreads();
  for (int n = tid; n < kHidden; n += nthreads) {
    float b = P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(P.head_w0[n * kHeadIn + kHeadK + k], app_tmp[k], b);
    S->head_b0_eff[n] = b;
  }
  __syncthreads();
}

// Reverse-mode adjoint of RGBRenderer + get_weights + distortion loss for ONE ray, one warp
// (NS/model_components/renderers.py:71-121,203-234; NS/cameras/rays.py:128-150;
// NS/model_components/losses.py:134-153,558-577; objective NS/pipelines/mitsuba_sdf.py:1170-1177).
// t_edge / s_edge: the ray's 49 euclidean / spacing edges; sigma (48) and rgb (48x3) in shared memory
// (sigma is overwritten with the weights).  Writes dL/dsigma, dL/drgb_sample (t^2 scaling applied)
// and the final weights of global ray r (chunk-local index).
// Returns this lane's max |d/d(pre-activation)| over its samples: |dL/drgb| * rgb (SafeExp) and
// |dL/dsigma| * exp(clamp(raw, -15, 15)) (trunc_exp) - the tensor-core backward scales its fp16 operands by it.
// SCAN_DIST: the pairwise distortion term sum_j w_j |u_i - u_j| over the sorted midpoints u through two prefix sums,
// u_i (2 A_i - A) + B - 2 B_i with A_i = sum_{j<i} w_j, B_i = sum_{j<i} w_j u_j (48 terms instead of 48^2; used by
// the tensor-core seed pass; the fp32 kernels keep the reference's O(n^2) summation order).
template <bool SCAN_DIST = false>
__device__ __forceinline__ float composite_adjoint_ray(const ModelDev& M, const FieldArgs& A, const float* t_edge,
                                                       const float* s_edge, float* sigma, const float* rgb,
                                                       int64_t r, int lane) {
  const bool eval = !A.training;
  float ds[2], w[2], T[2], e[2], delta[2], scale[2], sg[2];
  float seed_max = 0.f;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    int si = k * 32 + lane;
    ds[k] = 0.f; delta[k] = 0.f; scale[k] = 0.f; sg[k] = 0.f;
    if (si < kS2) {
      float t0 = t_edge[si], t1 = t_edge[si + 1];
      delta[k] = sub_rn(t1, t0);
      sg[k] = sigma[si];
      ds[k] = mul_rn(delta[k], sigma[si]);
      float tm = (t0 + t1) * 0.5f;
      scale[k] = fminf(fmaxf(tm * tm, 0.f), 1.f);
    }
  }
  weights_full<2>(ds, w, T, e, lane);
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    int si = k * 32 + lane;
    if (si < kS2) sigma[si] = w[k];           // weights visible to all lanes (distortion term)
  }
  __syncwarp();
  float c[2][3], fin[2][3];
  float cr = 0.f, cg = 0.f, cb = 0.f, acc = 0.f;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    int si = k * 32 + lane;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) { c[k][ch] = 0.f; fin[k][ch] = 1.f; }
    if (si < kS2) {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        float v = rgb[si * 3 + ch];
        if (eval) { fin[k][ch] = (isnan(v) || isinf(v)) ? 0.f : 1.f; v = nan_to_num(v); }
        c[k][ch] = v;
      }
      cr = fmaf(w[k], c[k][0], cr); cg = fmaf(w[k], c[k][1], cg); cb = fmaf(w[k], c[k][2], cb);
      acc += w[k];
      if (A.weights2) A.weights2[r * kS2 + si] = w[k];
    }
  }
  cr = warp_sum(cr); cg = warp_sum(cg); cb = warp_sum(cb); acc = warp_sum(acc);
  float last[3];
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    float v = rgb[(kS2 - 1) * 3 + ch];
    last[ch] = eval ? nan_to_num(v) : v;
  }
  const float bg = 1.f - acc;
  const int64_t rg = A.ray_offset + r;
  float gb[3];
  if (A.point_lights) {
    // brightness = sum(rgb * rgb_weights): the seed is the weight vector itself (nerfacto.py:562-568).  The pass also
    // returns the composite (renderer_rgb, clamp / nan_to_num in eval mode as in the forward) and the depth of the sample
    // with the largest weight * luminance(field_rgb) ("contrib" DepthRenderer, renderers.py:389-397; first index on ties)
    gb[0] = A.lum_w[0]; gb[1] = A.lum_w[1]; gb[2] = A.lum_w[2];
    float best = -INFINITY;
    int best_i = kS2;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      int si = k * 32 + lane;
      if (si < kS2) {
        const float lum = rgb[si * 3 + 0] * A.lum_w[0] + rgb[si * 3 + 1] * A.lum_w[1] + rgb[si * 3 + 2] * A.lum_w[2];
        const float cv = w[k] * lum;
        if (cv > best || (cv == best && si < best_i)) { best = cv; best_i = si; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
      if (ob > best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
    }
    if (lane == 0) {
      if (best_i >= kS2) best_i = 0;       // all-NaN contributions
      float pr = cr + last[0] * bg, pg = cg + last[1] * bg, pb = cb + last[2] * bg;
      if (eval) { pr = fmaxf(pr, 0.f); pg = fmaxf(pg, 0.f); pb = fmaxf(pb, 0.f); }
      A.rgb[r * 3 + 0] = pr; A.rgb[r * 3 + 1] = pg; A.rgb[r * 3 + 2] = pb;
      A.depth[r] = (t_edge[best_i] + t_edge[best_i + 1]) * 0.5f;
    }
  } else {
    gb[0] = A.grad_rgb[rg * 3 + 0]; gb[1] = A.grad_rgb[rg * 3 + 1]; gb[2] = A.grad_rgb[rg * 3 + 2];
  }
  if (eval) {   // clamp(min=0) passes the gradient where the composite is >= 0
    if (!(cr + last[0] * bg >= 0.f)) gb[0] = 0.f;
    if (!(cg + last[1] * bg >= 0.f)) gb[1] = 0.f;
    if (!(cb + last[2] * bg >= 0.f)) gb[2] = 0.f;
  }
  // regulariser means are per backward chunk (get_reg_loss_dict is evaluated per chunk)
  const int64_t chunk = A.reg_chunk > 0 ? A.reg_chunk : A.n_total;
  const int64_t c0 = (rg / chunk) * chunk;
  const float cnt = (float)min(chunk, A.n_total - c0);
  const float dm = M.distortion_mult / cnt;
  float gw[2], gww[2];
  float pa[2] = {0.f, 0.f}, pb[2] = {0.f, 0.f}, at = 0.f, bt = 0.f;
  if (SCAN_DIST) {
    float um[2], wv[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      int si = k * 32 + lane;
      um[k] = si < kS2 ? (s_edge[si] + s_edge[si + 1]) * 0.5f : 0.f;
      wv[k] = si < kS2 ? w[k] : 0.f;
    }
    const float ia0 = warp_incl_scan_f(wv[0], lane), ib0 = warp_incl_scan_f(wv[0] * um[0], lane);
    const float ta0 = __shfl_sync(0xffffffffu, ia0, 31), tb0 = __shfl_sync(0xffffffffu, ib0, 31);
    const float ia1 = warp_incl_scan_f(wv[1], lane) + ta0, ib1 = warp_incl_scan_f(wv[1] * um[1], lane) + tb0;
    at = __shfl_sync(0xffffffffu, ia1, 31);
    bt = __shfl_sync(0xffffffffu, ib1, 31);
    pa[0] = ia0 - wv[0]; pb[0] = ib0 - wv[0] * um[0];
    pa[1] = ia1 - wv[1]; pb[1] = ib1 - wv[1] * um[1];
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    int si = k * 32 + lane;
    gw[k] = 0.f; gww[k] = 0.f;
    if (si < kS2) {
      float ui = (s_edge[si] + s_edge[si + 1]) * 0.5f;
      float inter = 0.f;
      if (SCAN_DIST) {
        inter = fmaf(ui, 2.f * pa[k] - at, bt - 2.f * pb[k]);
      } else {
#pragma unroll 4
        for (int j = 0; j < kS2; ++j) {
          float uj = (s_edge[j] + s_edge[j + 1]) * 0.5f;
          inter = fmaf(sigma[j], fabsf(ui - uj), inter);
        }
      }
      float dist = 2.f * inter + 2.f * w[k] * (s_edge[si + 1] - s_edge[si]) * (1.f / 3.f);
      gw[k] = gb[0] * (c[k][0] - last[0]) + gb[1] * (c[k][1] - last[1]) + gb[2] * (c[k][2] - last[2]) + dm * dist;
      gww[k] = gw[k] * w[k];
    }
  }
  // suffix sums of gw*w over later samples
  float inc0 = warp_incl_scan_f(gww[0], lane);
  float tot0 = __shfl_sync(0xffffffffu, inc0, 31);
  float inc1 = warp_incl_scan_f(gww[1], lane) + tot0;
  float total = __shfl_sync(0xffffffffu, inc1, 31);
  float suffix[2] = {total - inc0, total - inc1};
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    int si = k * 32 + lane;
    if (si < kS2) {
      float gds = gw[k] * T[k] * e[k] - suffix[k];
      const float gsig = gds * delta[k] * scale[k];
      // trunc_exp backward exp(clamp(raw, -15, 15)) = clamp(sigma, e^-15, e^15); sigma == 0 <=> selector off
      const float dpre_sigma = sg[k] > 0.f ? gsig * fminf(fmaxf(sg[k], 3.0590232e-7f), 3269017.4f) : 0.f;
      seed_max = fmaxf(seed_max, fabsf(dpre_sigma));
      float wl = w[k] + (si == kS2 - 1 ? bg : 0.f);
      float dpre[3];
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        const float gc = gb[ch] * wl * fin[k][ch] * scale[k];
        if (!A.seeds4) A.g_rgbs[(r * kS2 + si) * 3 + ch] = gc;
        // SafeExp backward: exp(x) below the clamp at 88, 0 above (rgb = exp(min(x, 88)))
        const float c_raw = rgb[si * 3 + ch];
        dpre[ch] = c_raw < 1.6516e38f ? gc * c_raw : 0.f;
        seed_max = fmaxf(seed_max, fabsf(dpre[ch]));
      }
      if (A.seeds4) __stcs(A.seeds4 + r * kS2 + si, make_float4(dpre[0], dpre[1], dpre[2], dpre_sigma));
      else A.g_sigma[r * kS2 + si] = gsig;
    }
  }
  __syncwarp();
  return seed_max;
}

// Tangent GEMM: Yd[n][m] = (MASK ? Yp[n][m] > 0 : 1) * sum_k Wt[k][n] * Xd[k][m]  (no bias; ReLU mask from
// the primal tile Yp).  Same thread mapping as tile_gemm.
template <int K, int N, bool MASK>
__device__ __forceinline__ void tile_gemm_tangent(const float* __restrict__ Xd, const float* __restrict__ Wt,
                                                  const float* __restrict__ Yp, float* __restrict__ Yd, int tid) {
  constexpr int TN = N / 8;
  static_assert(N % 8 == 0 && (TN == 8 || TN == 2), "tile shape");
  const int tm = tid & 15, tn = tid >> 4;
  float acc[8][TN];
#pragma unroll
  for (int j = 0; j < TN; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i][j] = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(Xd + k * kTileM + tm * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(Xd + k * kTileM + tm * 8 + 4);
    float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float w[TN];
    if constexpr (TN == 8) {
      const float4 w0 = *reinterpret_cast<const float4*>(Wt + k * N + tn * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(Wt + k * N + tn * 8 + 4);
      w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
    } else {
      const float2 w0 = *reinterpret_cast<const float2*>(Wt + k * N + tn * 2);
      w[0] = w0.x; w[1] = w0.y;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
  }
#pragma unroll
  for (int j = 0; j < TN; ++j) {
    float o[8];
    const int off = (tn * TN + j) * kTileM + tm * 8;
    if (MASK) {
      const float4 y0 = *reinterpret_cast<const float4*>(Yp + off);
      const float4 y1 = *reinterpret_cast<const float4*>(Yp + off + 4);
      const float y[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = y[i] > 0.f ? acc[i][j] : 0.f;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = acc[i][j];
    }
    *reinterpret_cast<float4*>(Yd + off) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(Yd + off + 4) = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// tangent of sh16 for direction tangent (dx, dy, dz)
__device__ __forceinline__ void sh16_tangent(float x, float y, float z, float dx, float dy, float dz, float t[16]) {
  const float k1 = 0.4886025119029199f, k2 = 1.0925484305920792f;
  float xx = x * x, yy = y * y, zz = z * z;
  float dxx = 2.f * x * dx, dyy = 2.f * y * dy, dzz = 2.f * z * dz;
  t[0] = 0.f;
  t[1] = k1 * dy;
  t[2] = k1 * dz;
  t[3] = k1 * dx;
  t[4] = k2 * (dx * y + x * dy);
  t[5] = k2 * (dy * z + y * dz);
  t[6] = 0.9461746957575601f * dzz;
  t[7] = k2 * (dx * z + x * dz);
  t[8] = 0.5462742152960396f * (dxx - dyy);
  t[9] = 0.5900435899266435f * (dy * (3.f * xx - yy) + y * (3.f * dxx - dyy));
  t[10] = 2.890611442640554f * (dx * y * z + x * dy * z + x * y * dz);
  t[11] = 0.4570457994644658f * (dy * (5.f * zz - 1.f) + y * 5.f * dzz);
  t[12] = 0.3731763325901154f * (dz * (5.f * zz - 3.f) + z * 5.f * dzz);
  t[13] = 0.4570457994644658f * (dx * (5.f * zz - 1.f) + x * 5.f * dzz);
  t[14] = 1.445305721320277f * (dz * (xx - yy) + z * (dxx - dyy));
  t[15] = 0.5900435899266435f * (dx * (xx - 3.f * yy) + x * (dxx - 3.f * dyy));
}

}  // namespace nemit
