// field.cu - k_field_forward (exact fp32 path): 16-level hash gather -> base MLP 32->64->16 ->
// head MLP 63->64->64->3 -> SafeExp -> weights scan -> composite, for groups of 8 rays (384 samples =
// 3 tiles of 128).  The MLPs run as register-tiled fp32 GEMMs out of shared memory; this is the
// parity (<=1e-3) path, the tensor-core path lives in field_tc.cu.
#include "kernels.h"
#include "field_common.cuh"

namespace nemit {

struct FieldSmem {
  FieldWeightsSmem W;
  float bufA[kHidden * kTileM];          // enc (32 rows) / head input (31 rows) / h3
  float bufB[kHidden * kTileM];          // h1 / h2
  float sigma[kGroupSamples];
  float rgb[kGroupSamples * 3];
  float t_edge[kRaysPerGroup * (kS2 + 1)];
  float ray_o[kRaysPerGroup * 3];
  float ray_d[kRaysPerGroup * 3];
  float app_tmp[kApp];
  float s_edge[kRaysPerGroup * (kS2 + 1)];   // adjoint-seed mode only
};

// MODE 0: forward (rgb, accumulation, weights).  MODE 1: first pass of the backward - same forward,
// then the per-ray composite adjoint (RGB renderer, get_weights, distortion loss, t^2 gradient
// scaling) writes dL/dsigma and dL/drgb_sample for k_field_backward.
template <int MODE>
__global__ void __launch_bounds__(128)
k_field_forward(const __grid_constant__ ModelDev M, const __grid_constant__ FieldArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FieldSmem* sm = reinterpret_cast<FieldSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  load_field_weights(&sm->W, A.P, M.num_images, A.app_row, sm->app_tmp, tid, 128);

  const int64_t n_groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t r0 = g * kRaysPerGroup;
    const int rays_here = (int)min((int64_t)kRaysPerGroup, A.n - r0);
    __syncthreads();
    // stage rays and their 49 edges
    if (tid < kRaysPerGroup) {
      float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
      if (tid < rays_here) load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + r0 + tid,
                                    ray_scale(M, A), o, d);
#pragma unroll
      for (int k = 0; k < 3; ++k) { sm->ray_o[tid * 3 + k] = o[k]; sm->ray_d[tid * 3 + k] = d[k]; }
    }
    for (int i = tid; i < kRaysPerGroup * (kS2 + 1); i += 128) {
      int rr = i / (kS2 + 1);
      sm->t_edge[i] = rr < rays_here ? A.t_edges2[r0 * (kS2 + 1) + i] : 0.f;
      if (MODE == 1) sm->s_edge[i] = rr < rays_here ? A.s_edges2[r0 * (kS2 + 1) + i] : 0.f;
    }
    __syncthreads();

    for (int tile = 0; tile < kGroupSamples / kTileM; ++tile) {
      const int s = tile * kTileM + tid;            // sample within the group
      const int rr = s / kS2, si = s % kS2;
      const float t0 = sm->t_edge[rr * (kS2 + 1) + si], t1 = sm->t_edge[rr * (kS2 + 1) + si + 1];
      const float tsum = add_rn(t0, t1);
      float x[3], p[3], dirq[3];
      rot_sample(A.rot, &sm->ray_o[rr * 3], &sm->ray_d[rr * 3], tsum, x, dirq);      // Rotater.apply_frustums when set for this stage
      const bool sel = contract_normalise(x, M, p) && rr < rays_here;
      main_encode_to_tile(p, A.P.main_table, M.main_grid, sm->bufA, tid);
      __syncthreads();
      tile_gemm<kMainEnc, kHidden, true>(sm->bufA, sm->W.base_w0, sm->W.base_b0, sm->bufB, tid);
      __syncthreads();
      // base layer 2: 16 outputs -> rows 15..30 of bufA (row 15 = raw sigma scratch, rows 16..30 = geo)
      tile_gemm<kHidden, 16, false>(sm->bufB, sm->W.base_w1, sm->W.base_b1, sm->bufA + 15 * kTileM, tid);
      __syncthreads();
      {
        float raw = sm->bufA[15 * kTileM + tid];
        sm->sigma[s] = sel ? expf(raw) : 0.f;       // trunc_exp * selector
        float c[16];
        // SH at (d+1)/2 (NS/fields/base_field.py:146-152)
        sh16((dirq[0] + 1.f) * 0.5f, (dirq[1] + 1.f) * 0.5f, (dirq[2] + 1.f) * 0.5f, c);
#pragma unroll
        for (int k = 0; k < 16; ++k) sm->bufA[k * kTileM + tid] = c[k];
        sm->bufA[31 * kTileM + tid] = 0.f;      // pad row of the K=32 head GEMM
      }
      __syncthreads();
      tile_gemm<32, kHidden, true>(sm->bufA, sm->W.head_w0, sm->W.head_b0_eff, sm->bufB, tid);   // row 31: zero weight
      __syncthreads();
      tile_gemm<kHidden, kHidden, true>(sm->bufB, sm->W.head_w1, sm->W.head_b1, sm->bufA, tid);
      __syncthreads();
      {
        float a0 = sm->W.head_b2[0], a1 = sm->W.head_b2[1], a2 = sm->W.head_b2[2];
#pragma unroll 8
        for (int k = 0; k < kHidden; ++k) {
          float h = sm->bufA[k * kTileM + tid];
          const float4 w = *reinterpret_cast<const float4*>(sm->W.head_w2 + k * 4);
          a0 = fmaf(h, w.x, a0); a1 = fmaf(h, w.y, a1); a2 = fmaf(h, w.z, a2);
        }
        // SafeExp: exp(min(x + bias, 88))
        sm->rgb[s * 3 + 0] = expf(fminf(a0 + M.rgb_bias, 88.f));
        sm->rgb[s * 3 + 1] = expf(fminf(a1 + M.rgb_bias, 88.f));
        sm->rgb[s * 3 + 2] = expf(fminf(a2 + M.rgb_bias, 88.f));
      }
      __syncthreads();
    }

    // composite: warp w handles rays w and w+4; lane handles samples lane and lane+32
    for (int rr = wid; rr < rays_here; rr += 4) {
      if (MODE == 1) {
        composite_adjoint_ray(M, A, &sm->t_edge[rr * (kS2 + 1)], &sm->s_edge[rr * (kS2 + 1)], &sm->sigma[rr * kS2],
                              &sm->rgb[rr * kS2 * 3], r0 + rr, lane);
        continue;
      }
      float ds[2], w[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        ds[k] = 0.f;
        if (si < kS2) {
          float t0 = sm->t_edge[rr * (kS2 + 1) + si], t1 = sm->t_edge[rr * (kS2 + 1) + si + 1];
          ds[k] = mul_rn(sub_rn(t1, t0), sm->sigma[rr * kS2 + si]);
        }
      }
      weights_from_ds<2>(ds, w, lane);
      float cr = 0.f, cg = 0.f, cb = 0.f, acc = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        if (si < kS2) {
          const float* c = &sm->rgb[(rr * kS2 + si) * 3];
          float r = c[0], gg = c[1], b = c[2];
          if (!A.training) { r = nan_to_num(r); gg = nan_to_num(gg); b = nan_to_num(b); }
          cr = fmaf(w[k], r, cr); cg = fmaf(w[k], gg, cg); cb = fmaf(w[k], b, cb);
          acc += w[k];
          if (A.weights2) A.weights2[(r0 + rr) * kS2 + si] = w[k];
        }
      }
      cr = warp_sum(cr); cg = warp_sum(cg); cb = warp_sum(cb); acc = warp_sum(acc);
      if (lane == 0) {
        const float* last = &sm->rgb[(rr * kS2 + kS2 - 1) * 3];
        float lr = last[0], lg = last[1], lb = last[2];
        if (!A.training) { lr = nan_to_num(lr); lg = nan_to_num(lg); lb = nan_to_num(lb); }
        float bg = 1.f - acc;                       // background "last_sample"
        cr += lr * bg; cg += lg * bg; cb += lb * bg;
        if (!A.training) { cr = fmaxf(cr, 0.f); cg = fmaxf(cg, 0.f); cb = fmaxf(cb, 0.f); }
        A.rgb[(r0 + rr) * 3 + 0] = cr; A.rgb[(r0 + rr) * 3 + 1] = cg; A.rgb[(r0 + rr) * 3 + 2] = cb;
        if (A.accumulation) A.accumulation[r0 + rr] = acc;
      }
    }
  }
}

template <int MODE>
static cudaError_t launch_field_mode(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(FieldSmem);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_forward<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  int64_t ctas = groups < 148 * 2 ? groups : 148 * 2;     // persistent: 2 CTAs per SM (smem-limited)
  k_field_forward<MODE><<<(unsigned)ctas, 128, smem, stream>>>(M, A);
  return cudaGetLastError();
}
cudaError_t launch_field_forward(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  return launch_field_mode<0>(M, A, stream);
}
cudaError_t launch_field_adjoint_seed(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  return launch_field_mode<1>(M, A, stream);
}

// ---------------------------------------------------------------------------------------------
// component drop-in: HashEncoding.forward (NS/field_components/encodings.py:341-382)
// ---------------------------------------------------------------------------------------------
__global__ void k_hash_encode(const __grid_constant__ GridDev G, const float2* __restrict__ table,
                              const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float p[3] = {x[i * 3], x[i * 3 + 1], x[i * 3 + 2]};
    for (int l = 0; l < G.num_levels; ++l) {
      LevelCell c;
      level_cell(p, G.scal[l], G.table_mask, (uint32_t)l * G.table_size, c);
      float2 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = ldg_f2(table, c.idx[k]);
      float2 r = blend_cell(v, c);
      out[i * 2 * G.num_levels + 2 * l] = r.x;
      out[i * 2 * G.num_levels + 2 * l + 1] = r.y;
    }
  }
}

cudaError_t launch_hash_encode(const GridDev& G, const float2* table, const float* x, int64_t n, float* out,
                               cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_hash_encode<<<(unsigned)blocks, 256, 0, stream>>>(G, table, x, n, out);
  return cudaGetLastError();
}

}  // namespace nemit
