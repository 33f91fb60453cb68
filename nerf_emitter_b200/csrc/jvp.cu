// jvp.cu - k_field_jvp: forward-mode derivative of the emitter radiance w.r.t. the ray origin and direction
// (exact fp32 path).  Sampling runs on the primal rays; tangents are attached to the 48 final samples only and
// pushed through positions -> contraction -> hash grid -> base MLP -> (SH, geo) -> head MLP -> SafeExp ->
// weights -> composite.  Primal and tangent tiles travel together through the shared-memory GEMMs.
//
// Reference: NeRFEmitterOp.ns_forward_ray (DS/emitters/nerf_emitter_op.py:60-90),
// Model.forward_grad_for_camera_ray_bundle (NS/models/base_model.py:315-344), SdfNerfactoModel.forward_grad
// (NS/models/sdf_nerfacto.py:85-113), _TruncExp.jvp (NS/field_components/activations.py:38-42).
#include "kernels.h"
#include "field_common.cuh"

namespace nemit {

struct JvpSmem {
  FieldWeightsSmem W;
  float bufA[kHidden * kTileM];
  float bufB[kHidden * kTileM];
  float dA[kHidden * kTileM];            // tangent tiles
  float dB[kHidden * kTileM];
  float sigma[kGroupSamples], dsigma[kGroupSamples];
  float rgb[kGroupSamples * 3], drgb[kGroupSamples * 3];
  float t_edge[kRaysPerGroup * (kS2 + 1)];
  float ray_o[kRaysPerGroup * 3], ray_d[kRaysPerGroup * 3];
  float ray_do[kRaysPerGroup * 3], ray_dd[kRaysPerGroup * 3];
  float app_tmp[kApp];
};

// 16-level encode of one point together with its directional derivative along dp
__device__ __forceinline__ void main_encode_dual_to_tile(const float p[3], const float dp[3],
                                                         const float2* __restrict__ table, const GridDev& G,
                                                         float* X, float* Xd, int m) {
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
      float2 gx, gy, gz;
      blend_cell_grad(v[j], cell[j], gx, gy, gz);
      const float sc = G.scal[l0 + j];
      X[(2 * (l0 + j)) * kTileM + m] = r.x;
      X[(2 * (l0 + j) + 1) * kTileM + m] = r.y;
      Xd[(2 * (l0 + j)) * kTileM + m] = sc * (gx.x * dp[0] + gy.x * dp[1] + gz.x * dp[2]);
      Xd[(2 * (l0 + j) + 1) * kTileM + m] = sc * (gx.y * dp[0] + gy.y * dp[1] + gz.y * dp[2]);
    }
  }
}

__global__ void __launch_bounds__(128)
k_field_jvp(const __grid_constant__ ModelDev M, const __grid_constant__ FieldArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  JvpSmem* sm = reinterpret_cast<JvpSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  load_field_weights(&sm->W, A.P, M.num_images, A.app_row, sm->app_tmp, tid, 128);

  const int64_t n_groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t r0 = g * kRaysPerGroup;
    const int rays_here = (int)min((int64_t)kRaysPerGroup, A.n - r0);
    __syncthreads();
    if (tid < kRaysPerGroup) {
      float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f}, dot[3] = {0.f, 0.f, 0.f}, ddt[3] = {0.f, 0.f, 0.f};
      if (tid < rays_here) {
        const int64_t R = A.ray_offset + r0 + tid;
        load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, R, ray_scale(M, A), o, d);
        const float nr = A.near_per_ray ? A.near_per_ray[R] : A.near_scalar;
        float dom[3], dvm[3];
        if (A.nerf_layout) {
          // forward_grad_for_camera_ray_bundle (NS/models/base_model.py:315-344): (N,3) tangents in NeRF coordinates
#pragma unroll
          for (int k = 0; k < 3; ++k) { ddt[k] = A.d_v[R * 3 + k]; dot[k] = A.d_o[R * 3 + k] + nr * ddt[k]; }
        } else {
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            dvm[k] = A.d_v[k * A.n_total + R];
            dom[k] = 2.f * M.scene_scale * (A.d_o[k * A.n_total + R] + nr * dvm[k]);   // nerf_emitter_op.py:69
          }
          dot[0] = dom[2]; dot[1] = dom[0]; dot[2] = dom[1];
          ddt[0] = dvm[2]; ddt[1] = dvm[0]; ddt[2] = dvm[1];
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        sm->ray_o[tid * 3 + k] = o[k]; sm->ray_d[tid * 3 + k] = d[k];
        sm->ray_do[tid * 3 + k] = dot[k]; sm->ray_dd[tid * 3 + k] = ddt[k];
      }
    }
    for (int i = tid; i < kRaysPerGroup * (kS2 + 1); i += 128) {
      int rr = i / (kS2 + 1);
      sm->t_edge[i] = rr < rays_here ? A.t_edges2[r0 * (kS2 + 1) + i] : 0.f;
    }
    __syncthreads();

    for (int tile = 0; tile < kGroupSamples / kTileM; ++tile) {
      const int s = tile * kTileM + tid;
      const int rr = s / kS2, si = s % kS2;
      const float t0 = sm->t_edge[rr * (kS2 + 1) + si], t1 = sm->t_edge[rr * (kS2 + 1) + si + 1];
      const float tsum = add_rn(t0, t1);
      float x[3], p[3], dx[3], dp[3], dirq[3];
      // primal: Rotater.apply_frustums when set; the tangents stay those of the unrotated ray - forward_grad attaches them
      // to the already rotated frustums (NS/models/sdf_nerfacto.py:96-104)
      rot_sample(A.rot, &sm->ray_o[rr * 3], &sm->ray_d[rr * 3], tsum, x, dirq);
#pragma unroll
      for (int k = 0; k < 3; ++k) dx[k] = sm->ray_do[rr * 3 + k] + sm->ray_dd[rr * 3 + k] * (tsum * 0.5f);
      const bool sel = contract_normalise(x, M, p) && rr < rays_here;
      contract_jvp(x, sel, dx, dp);
      main_encode_dual_to_tile(p, dp, A.P.main_table, M.main_grid, sm->bufA, sm->dA, tid);
      __syncthreads();
      tile_gemm<kMainEnc, kHidden, true>(sm->bufA, sm->W.base_w0, sm->W.base_b0, sm->bufB, tid);
      tile_gemm_tangent<kMainEnc, kHidden, true>(sm->dA, sm->W.base_w0, sm->bufB, sm->dB, tid);
      __syncthreads();
      tile_gemm<kHidden, 16, false>(sm->bufB, sm->W.base_w1, sm->W.base_b1, sm->bufA + 15 * kTileM, tid);
      tile_gemm_tangent<kHidden, 16, false>(sm->dB, sm->W.base_w1, nullptr, sm->dA + 15 * kTileM, tid);
      __syncthreads();
      {
        const float raw = sm->bufA[15 * kTileM + tid], draw = sm->dA[15 * kTileM + tid];
        sm->sigma[s] = sel ? expf(raw) : 0.f;
        sm->dsigma[s] = sel ? expf(fminf(fmaxf(raw, -15.f), 15.f)) * draw : 0.f;   // _TruncExp.jvp
        float c[16], ct[16];
        const float qx = (dirq[0] + 1.f) * 0.5f, qy = (dirq[1] + 1.f) * 0.5f, qz = (dirq[2] + 1.f) * 0.5f;
        sh16(qx, qy, qz, c);
        // forward-mode tangents do flow through the no_grad SH
        sh16_tangent(qx, qy, qz, sm->ray_dd[rr * 3 + 0] * 0.5f, sm->ray_dd[rr * 3 + 1] * 0.5f,
                     sm->ray_dd[rr * 3 + 2] * 0.5f, ct);
#pragma unroll
        for (int k = 0; k < 16; ++k) { sm->bufA[k * kTileM + tid] = c[k]; sm->dA[k * kTileM + tid] = ct[k]; }
        sm->bufA[31 * kTileM + tid] = 0.f;
        sm->dA[31 * kTileM + tid] = 0.f;
      }
      __syncthreads();
      tile_gemm<32, kHidden, true>(sm->bufA, sm->W.head_w0, sm->W.head_b0_eff, sm->bufB, tid);
      tile_gemm_tangent<32, kHidden, true>(sm->dA, sm->W.head_w0, sm->bufB, sm->dB, tid);
      __syncthreads();
      tile_gemm<kHidden, kHidden, true>(sm->bufB, sm->W.head_w1, sm->W.head_b1, sm->bufA, tid);
      tile_gemm_tangent<kHidden, kHidden, true>(sm->dB, sm->W.head_w1, sm->bufA, sm->dA, tid);
      __syncthreads();
      {
        float a[3] = {sm->W.head_b2[0], sm->W.head_b2[1], sm->W.head_b2[2]}, da[3] = {0.f, 0.f, 0.f};
#pragma unroll 8
        for (int k = 0; k < kHidden; ++k) {
          const float h = sm->bufA[k * kTileM + tid], dh = sm->dA[k * kTileM + tid];
          const float4 w = *reinterpret_cast<const float4*>(sm->W.head_w2 + k * 4);
          a[0] = fmaf(h, w.x, a[0]); a[1] = fmaf(h, w.y, a[1]); a[2] = fmaf(h, w.z, a[2]);
          da[0] = fmaf(dh, w.x, da[0]); da[1] = fmaf(dh, w.y, da[1]); da[2] = fmaf(dh, w.z, da[2]);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const float xj = a[j] + M.rgb_bias;
          const float v = expf(fminf(xj, 88.f));
          sm->rgb[s * 3 + j] = v;
          sm->drgb[s * 3 + j] = xj < 88.f ? v * da[j] : 0.f;
        }
      }
      __syncthreads();
    }

    // composite with tangents: warp w handles rays w and w+4
    for (int rr = wid; rr < rays_here; rr += 4) {
      float ds[2], dds[2], w[2], T[2], e[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        ds[k] = 0.f; dds[k] = 0.f;
        if (si < kS2) {
          float delta = sub_rn(sm->t_edge[rr * (kS2 + 1) + si + 1], sm->t_edge[rr * (kS2 + 1) + si]);
          ds[k] = mul_rn(delta, sm->sigma[rr * kS2 + si]);
          dds[k] = delta * sm->dsigma[rr * kS2 + si];
        }
      }
      weights_full<2>(ds, w, T, e, lane);
      // exclusive prefix of the tangent of delta*sigma
      float i0 = warp_incl_scan_f(dds[0], lane);
      float tot0 = __shfl_sync(0xffffffffu, i0, 31);
      float i1 = warp_incl_scan_f(dds[1], lane) + tot0;
      float pre[2] = {i0 - dds[0], i1 - dds[1]};
      float c[3] = {0.f, 0.f, 0.f}, dc[3] = {0.f, 0.f, 0.f}, acc = 0.f, dacc = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        if (si < kS2) {
          // w = (1-e) T ; dw = e*dds*T - (1-e)*T*prefix(dds)
          float dw = e[k] * dds[k] * T[k] - w[k] * pre[k];
          if (!isfinite(mul_rn(sub_rn(1.f, e[k]), T[k]))) dw = 0.f;   // nan_to_num cuts the tangent
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            float v = sm->rgb[(rr * kS2 + si) * 3 + ch], dv = sm->drgb[(rr * kS2 + si) * 3 + ch];
            if (!A.training) { if (isnan(v) || isinf(v)) dv = 0.f; v = nan_to_num(v); }
            c[ch] = fmaf(w[k], v, c[ch]);
            dc[ch] += dw * v + w[k] * dv;
          }
          acc += w[k];
          dacc += dw;
        }
      }
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) { c[ch] = warp_sum(c[ch]); dc[ch] = warp_sum(dc[ch]); }
      acc = warp_sum(acc);
      dacc = warp_sum(dacc);
      if (lane == 0) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          float lv = sm->rgb[(rr * kS2 + kS2 - 1) * 3 + ch], dlv = sm->drgb[(rr * kS2 + kS2 - 1) * 3 + ch];
          if (!A.training) { if (isnan(lv) || isinf(lv)) dlv = 0.f; lv = nan_to_num(lv); }
          float comp = c[ch] + lv * (1.f - acc);
          float dcomp = dc[ch] + dlv * (1.f - acc) - lv * dacc;
          if (!A.training && !(comp >= 0.f)) dcomp = 0.f;          // clamp(min=0)
          if (isnan(dcomp)) dcomp = 0.f;                            // nerf_emitter_op.py:87
          A.d_rgb[(r0 + rr) * 3 + ch] = dcomp;
        }
      }
    }
  }
}

cudaError_t launch_field_jvp(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(JvpSmem);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_jvp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  int64_t ctas = groups < 148 ? groups : 148;
  k_field_jvp<<<(unsigned)ctas, 128, smem, stream>>>(M, A);
  return cudaGetLastError();
}

}  // namespace nemit
