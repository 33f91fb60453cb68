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
__device__ __forceinline__ void load_field_weights(FieldWeightsSmem* S, const FieldParams& P, int num_images,
                                                   int training, int camera_idx, float* app_tmp /*32 floats smem*/,
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
    float e;
    if (training) {
      e = P.appearance[(int64_t)camera_idx * kApp + i];
    } else {
      float s = 0.f;
      for (int r = 0; r < num_images; ++r) s += P.appearance[(int64_t)r * kApp + i];
      e = s / (float)num_images;
    }
    app_tmp[i] = e;
  }
  __syncthreads();
  for (int n = tid; n < kHidden; n += nthreads) {
    float b = P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(P.head_w0[n * kHeadIn + kHeadK + k], app_tmp[k], b);
    S->head_b0_eff[n] = b;
  }
  __syncthreads();
}

}  // namespace nemit
