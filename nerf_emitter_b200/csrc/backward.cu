// backward.cu - reverse-mode kernels of the emitter path (exact fp32 path):
//   k_field_backward   : main field on the 48 final samples - forward recompute per 128-sample tile, MLP
//                        dgrad/wgrad out of shared memory, hash-table scatter-add, d/d(position) -> per-ray sums
//   k_finalize_ray_grads : per-ray sums -> grad_o / grad_v in Mitsuba coordinates ((3,N) SoA)
// (the proposal-network backward lives in prop_backward.cu, the composite adjoint in field.cu MODE 1)
//
// Reference: SdfNerfactoModel.get_backward_outputs (NS/models/sdf_nerfacto.py:115-168),
// MitsubaSdfPipeline.backward_for_camera_ray_bundle (NS/pipelines/mitsuba_sdf.py:1144-1219),
// NeRFEmitterOp.ns_backward_ray (DS/emitters/nerf_emitter_op.py:92-117), _TruncExp.backward
// (NS/field_components/activations.py:44-48).
#include "kernels.h"
#include "field_common.cuh"

namespace nemit {

// Y[k][m] <- mask(Y[k][m]) * sum_{n<NR} Wt[k*LDW + n] * G[n][m]   for k < KO, in place on Y.
// Wt is the FORWARD k-major weight block (row = input feature k, column = output n), so the dgrad needs
// no second copy of the weights.  128 threads; thread (tm, tk) owns 8 samples x KO/8 rows.
template <int KO, int NR, int LDW, bool MASK>
__device__ __forceinline__ void dgrad_tile(const float* __restrict__ G, const float* __restrict__ Wt,
                                           float* __restrict__ Y, int tid) {
  constexpr int TK = KO / 8;
  static_assert(KO % 8 == 0 && NR % 4 == 0, "tile shape");
  const int tm = tid & 15, tk = tid >> 4;
  float acc[TK][8];
#pragma unroll
  for (int j = 0; j < TK; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
#pragma unroll 2
  for (int n = 0; n < NR; n += 4) {
    float g[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 a0 = *reinterpret_cast<const float4*>(G + (n + q) * kTileM + tm * 8);
      const float4 a1 = *reinterpret_cast<const float4*>(G + (n + q) * kTileM + tm * 8 + 4);
      g[q][0] = a0.x; g[q][1] = a0.y; g[q][2] = a0.z; g[q][3] = a0.w;
      g[q][4] = a1.x; g[q][5] = a1.y; g[q][6] = a1.z; g[q][7] = a1.w;
    }
#pragma unroll
    for (int j = 0; j < TK; ++j) {
      const float4 w = *reinterpret_cast<const float4*>(Wt + (tk * TK + j) * LDW + n);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[j][i] = fmaf(w.x, g[0][i], acc[j][i]);
        acc[j][i] = fmaf(w.y, g[1][i], acc[j][i]);
        acc[j][i] = fmaf(w.z, g[2][i], acc[j][i]);
        acc[j][i] = fmaf(w.w, g[3][i], acc[j][i]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < TK; ++j) {
    float* dst = Y + (tk * TK + j) * kTileM + tm * 8;
    float o[8];
    if (MASK) {
      const float4 y0 = *reinterpret_cast<const float4*>(dst);
      const float4 y1 = *reinterpret_cast<const float4*>(dst + 4);
      const float y[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = y[i] > 0.f ? acc[j][i] : 0.f;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = acc[j][i];
    }
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// dW[k*LDW + n] += sum_m X[k][m] * G[n][m]  (k < K, n < N), accumulators in shared memory (each entry owned
// by one thread).  Each lane walks m in a rotated order so that lanes reading different rows hit
// different banks.
template <int K, int N, int LDW, int TK, int TN>
__device__ __forceinline__ void wgrad_tile(const float* __restrict__ X, const float* __restrict__ G,
                                           float* __restrict__ dW, int tid) {
  constexpr int KB = K / TK, NB = N / TN;
  static_assert(KB * NB == 128, "one (k,n) block per thread");
  const int kb = tid % KB, nb = tid / KB;
  const int rot = 4 * ((kb + nb) & 31);
  float acc[TK][TN];
#pragma unroll
  for (int a = 0; a < TK; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) acc[a][b] = 0.f;
#pragma unroll 2
  for (int mi = 0; mi < kTileM; mi += 4) {
    const int m = (mi + rot) & (kTileM - 1);
    float4 x[TK], g[TN];
#pragma unroll
    for (int a = 0; a < TK; ++a) x[a] = *reinterpret_cast<const float4*>(X + (kb * TK + a) * kTileM + m);
#pragma unroll
    for (int b = 0; b < TN; ++b) g[b] = *reinterpret_cast<const float4*>(G + (nb * TN + b) * kTileM + m);
#pragma unroll
    for (int a = 0; a < TK; ++a)
#pragma unroll
      for (int b = 0; b < TN; ++b) {
        acc[a][b] = fmaf(x[a].x, g[b].x, acc[a][b]);
        acc[a][b] = fmaf(x[a].y, g[b].y, acc[a][b]);
        acc[a][b] = fmaf(x[a].z, g[b].z, acc[a][b]);
        acc[a][b] = fmaf(x[a].w, g[b].w, acc[a][b]);
      }
  }
#pragma unroll
  for (int a = 0; a < TK; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) dW[(kb * TK + a) * LDW + nb * TN + b] += acc[a][b];
}

// db[n] += sum_m G[n][m] for n < N (N <= 128): thread n sums its row in a lane-rotated order
template <int N>
__device__ __forceinline__ void rowsum_tile(const float* __restrict__ G, float* __restrict__ db, int tid) {
  if (tid < N) {
    float s = 0.f;
#pragma unroll 8
    for (int mi = 0; mi < kTileM; ++mi) s += G[tid * kTileM + ((mi + tid) & (kTileM - 1))];
    db[tid] += s;
  }
}

struct BwdSmem {
  FieldWeightsSmem W;                 // forward (k-major) weights, also used transposed by the dgrads
  FieldWeightsSmem dW;                // gradient accumulators in the same layout (head_b0_eff slot = d head_b0)
  float enc[kMainEnc * kTileM];       // hash encoding            -> d/d enc
  float h1[kHidden * kTileM];         // base hidden              -> its gradient
  float in[32 * kTileM];              // head input: SH rows 0..15, geo rows 16..30, pad row 31; rows 15..30 double
                                      // as the base-MLP output (row 15 = raw density) and later its gradient
  float h2[kHidden * kTileM];
  float h3[kHidden * kTileM];
  float gout[4 * kTileM];             // d/d(head output), row 3 = 0
  float app_tmp[kApp];
};

__global__ void __launch_bounds__(128)
k_field_backward(const __grid_constant__ ModelDev M, const __grid_constant__ FieldBwdArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BwdSmem* sm = reinterpret_cast<BwdSmem*>(smem_raw);
  const int tid = threadIdx.x;
  load_field_weights(&sm->W, A.P, M.num_images, A.app_row, sm->app_tmp, tid, 128);
  {
    float* z = reinterpret_cast<float*>(&sm->dW);
    for (int i = tid; i < (int)(sizeof(FieldWeightsSmem) / sizeof(float)); i += 128) z[i] = 0.f;
  }
  __syncthreads();

  const int64_t n_samples = A.n * kS2;
  const int64_t n_tiles = (n_samples + kTileM - 1) / kTileM;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t s = tile * kTileM + tid;
    const bool valid = s < n_samples;
    const int64_t rr = valid ? s / kS2 : 0;
    const int si = valid ? (int)(s % kS2) : 0;
    float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f}, t0 = 0.f, t1 = 0.f;
    if (valid) {
      load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + rr, ray_scale(M, A), o, d);
      t0 = A.t_edges2[rr * (kS2 + 1) + si];
      t1 = A.t_edges2[rr * (kS2 + 1) + si + 1];
    }
    const float tsum = add_rn(t0, t1);
    float x[3], p[3], dirq[3];
    // Rotater.apply_frustums only where the reference's reverse pass sees rotated final samples: the light-point-cloud
    // pass (calc_brightness_grad, NS/models/nerfacto.py:547-556); get_backward_outputs puts the unrotated rays back
    // (NS/models/sdf_nerfacto.py:129-131).  The host sets A.rot accordingly.
    const bool rotated = rot_sample(A.rot, o, d, tsum, x, dirq);
    const bool sel = contract_normalise(x, M, p);
    // ---- forward recompute, activations kept ----------------------------------------------
    main_encode_to_tile(p, A.P.main_table, M.main_grid, sm->enc, tid);
    __syncthreads();
    tile_gemm<kMainEnc, kHidden, true>(sm->enc, sm->W.base_w0, sm->W.base_b0, sm->h1, tid);
    __syncthreads();
    tile_gemm<kHidden, 16, false>(sm->h1, sm->W.base_w1, sm->W.base_b1, sm->in + 15 * kTileM, tid);
    __syncthreads();
    const float raw = sm->in[15 * kTileM + tid];
    {
      float c[16];
      sh16((dirq[0] + 1.f) * 0.5f, (dirq[1] + 1.f) * 0.5f, (dirq[2] + 1.f) * 0.5f, c);
#pragma unroll
      for (int k = 0; k < 16; ++k) sm->in[k * kTileM + tid] = c[k];
      sm->in[31 * kTileM + tid] = 0.f;
    }
    __syncthreads();
    tile_gemm<32, kHidden, true>(sm->in, sm->W.head_w0, sm->W.head_b0_eff, sm->h2, tid);
    __syncthreads();
    tile_gemm<kHidden, kHidden, true>(sm->h2, sm->W.head_w1, sm->W.head_b1, sm->h3, tid);
    __syncthreads();
    {
      float a0 = sm->W.head_b2[0], a1 = sm->W.head_b2[1], a2 = sm->W.head_b2[2];
#pragma unroll 8
      for (int k = 0; k < kHidden; ++k) {
        float h = sm->h3[k * kTileM + tid];
        const float4 w = *reinterpret_cast<const float4*>(sm->W.head_w2 + k * 4);
        a0 = fmaf(h, w.x, a0); a1 = fmaf(h, w.y, a1); a2 = fmaf(h, w.z, a2);
      }
      float a[3] = {a0 + M.rgb_bias, a1 + M.rgb_bias, a2 + M.rgb_bias};
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        // SafeExp backward: exp(clamp(x, max=88)) -> gradient exp(x) below the clamp, 0 above
        float g = valid ? A.g_rgbs[s * 3 + j] : 0.f;
        sm->gout[j * kTileM + tid] = (a[j] < 88.f) ? g * expf(a[j]) : 0.f;
      }
      sm->gout[3 * kTileM + tid] = 0.f;
    }
    __syncthreads();
    // ---- head MLP backward -----------------------------------------------------------------
    wgrad_tile<kHidden, 4, 4, 2, 1>(sm->h3, sm->gout, sm->dW.head_w2, tid);
    rowsum_tile<4>(sm->gout, sm->dW.head_b2, tid);
    __syncthreads();
    dgrad_tile<kHidden, 4, 4, true>(sm->gout, sm->W.head_w2, sm->h3, tid);
    __syncthreads();
    wgrad_tile<kHidden, kHidden, kHidden, 4, 8>(sm->h2, sm->h3, sm->dW.head_w1, tid);
    rowsum_tile<kHidden>(sm->h3, sm->dW.head_b1, tid);
    __syncthreads();
    dgrad_tile<kHidden, kHidden, kHidden, true>(sm->h3, sm->W.head_w1, sm->h2, tid);
    __syncthreads();
    wgrad_tile<32, kHidden, kHidden, 4, 4>(sm->in, sm->h2, sm->dW.head_w0, tid);
    rowsum_tile<kHidden>(sm->h2, sm->dW.head_b0_eff, tid);
    __syncthreads();
    // geo-feature gradient into rows 16..31 (row 31 has zero weights); SH carries no reverse-mode gradient
    dgrad_tile<16, kHidden, kHidden, false>(sm->h2, sm->W.head_w0 + 16 * kHidden, sm->in + 16 * kTileM, tid);
    {
      // trunc_exp backward: exp(clamp(raw, -15, 15)); the selector multiplies the density
      float gs = (valid && sel) ? A.g_sigma[s] : 0.f;
      sm->in[15 * kTileM + tid] = gs * expf(fminf(fmaxf(raw, -15.f), 15.f));
    }
    __syncthreads();
    // ---- base MLP backward -----------------------------------------------------------------
    wgrad_tile<kHidden, 16, 16, 2, 4>(sm->h1, sm->in + 15 * kTileM, sm->dW.base_w1, tid);
    rowsum_tile<16>(sm->in + 15 * kTileM, sm->dW.base_b1, tid);
    __syncthreads();
    dgrad_tile<kHidden, 16, 16, true>(sm->in + 15 * kTileM, sm->W.base_w1, sm->h1, tid);
    __syncthreads();
    wgrad_tile<kMainEnc, kHidden, kHidden, 4, 4>(sm->enc, sm->h1, sm->dW.base_w0, tid);
    rowsum_tile<kHidden>(sm->h1, sm->dW.base_b0, tid);
    __syncthreads();
    dgrad_tile<kMainEnc, kHidden, kHidden, false>(sm->h1, sm->W.base_w0, sm->enc, tid);
    __syncthreads();
    // ---- hash grid backward: table scatter-add and d/d(position) --------------------------
    if (valid) {
      float gp[3] = {0.f, 0.f, 0.f};
#pragma unroll 1
      for (int l0 = 0; l0 < kMainLevels; l0 += 4) {
        LevelCell cell[4];
        float2 v[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          level_cell(p, M.main_grid.scal[l0 + j], M.main_grid.table_mask, (uint32_t)(l0 + j) * M.main_grid.table_size, cell[j]);
          if (sel) {
#pragma unroll
            for (int c = 0; c < 8; ++c) v[j][c] = ldg_f2(A.P.main_table, cell[j].idx[c]);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 g = make_float2(sm->enc[(2 * (l0 + j)) * kTileM + tid], sm->enc[(2 * (l0 + j) + 1) * kTileM + tid]);
          if (g.x == 0.f && g.y == 0.f) continue;
          float cw[8];
          corner_weights(cell[j], cw);
          const float gxs = g.x * A.loss_scale, gys = g.y * A.loss_scale;
          if (A.G.main_table) {        // null in the ray-gradient-only pass (nemit_point_lights)
#pragma unroll
            for (int c = 0; c < 8; ++c) atomic_add_f2(A.G.main_table + cell[j].idx[c], make_float2(cw[c] * gxs, cw[c] * gys));
          }
          if (sel) {
            float2 dx, dy, dz;
            blend_cell_grad(v[j], cell[j], dx, dy, dz);
            const float sc = M.main_grid.scal[l0 + j];
            gp[0] = fmaf(sc, g.x * dx.x + g.y * dx.y, gp[0]);
            gp[1] = fmaf(sc, g.x * dy.x + g.y * dy.y, gp[1]);
            gp[2] = fmaf(sc, g.x * dz.x + g.y * dz.y, gp[2]);
          }
        }
      }
      if (sel) {
        float gx[3];
        contract_vjp(x, sel, gp, gx);
        if (rotated) {            // x = R (o + d t) + t0: the cotangent goes back through R^T
          float gr[3];
          rot_transpose(A.rot, gx, gr);
          gx[0] = gr[0]; gx[1] = gr[1]; gx[2] = gr[2];
        }
        const float tm = tsum * 0.5f;
        float* rg = A.ray_grad + rr * 6;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          atomicAdd(rg + k, gx[k]);
          atomicAdd(rg + 3 + k, gx[k] * tm);
        }
      }
    }
    __syncthreads();
  }

  // ---- flush this CTA's weight-gradient accumulators (torch layouts, GradScaler scale applied) ------
  __syncthreads();
  if (!A.G.main_table) return;
  const float ls = A.loss_scale;
  const FieldWeightsSmem& D = sm->dW;
  for (int i = tid; i < kHidden * kMainEnc; i += 128) { int n = i / kMainEnc, k = i % kMainEnc; atomicAdd(A.G.base_w0 + i, ls * D.base_w0[k * kHidden + n]); }
  for (int i = tid; i < 16 * kHidden; i += 128) { int n = i / kHidden, k = i % kHidden; atomicAdd(A.G.base_w1 + i, ls * D.base_w1[k * 16 + n]); }
  for (int i = tid; i < kHidden * kHeadIn; i += 128) {
    int n = i / kHeadIn, k = i % kHeadIn;
    // appearance columns: the input is the same 32-vector for every sample -> rank-1 with d head_b0
    float g = k < kHeadK ? D.head_w0[k * kHidden + n] : D.head_b0_eff[n] * sm->app_tmp[k - kHeadK];
    atomicAdd(A.G.head_w0 + i, ls * g);
  }
  for (int i = tid; i < kHidden * kHidden; i += 128) { int n = i / kHidden, k = i % kHidden; atomicAdd(A.G.head_w1 + i, ls * D.head_w1[k * kHidden + n]); }
  for (int i = tid; i < 3 * kHidden; i += 128) { int n = i / kHidden, k = i % kHidden; atomicAdd(A.G.head_w2 + i, ls * D.head_w2[k * 4 + n]); }
  for (int i = tid; i < kHidden; i += 128) {
    atomicAdd(A.G.base_b0 + i, ls * D.base_b0[i]);
    atomicAdd(A.G.head_b0 + i, ls * D.head_b0_eff[i]);
    atomicAdd(A.G.head_b1 + i, ls * D.head_b1[i]);
  }
  for (int i = tid; i < 16; i += 128) atomicAdd(A.G.base_b1 + i, ls * D.base_b1[i]);
  for (int i = tid; i < 3; i += 128) atomicAdd(A.G.head_b2 + i, ls * D.head_b2[i]);
  // appearance embedding (NS/fields/nerfacto_field.py:272-285): row / mean / zeros as resolved by the host (appearance_component)
  for (int k = tid; k < kApp; k += 128) {
    float g = 0.f;
    for (int n = 0; n < kHidden; ++n) g = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + k], D.head_b0_eff[n], g);
    g *= ls;
    appearance_component_grad(A.G.appearance, M.num_images, A.app_row, k, g);
  }
}

cudaError_t launch_field_backward(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(BwdSmem);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t tiles = (A.n * kS2 + kTileM - 1) / kTileM;
  int64_t ctas = tiles < 148 ? tiles : 148;      // persistent, one CTA per SM (shared-memory limited)
  k_field_backward<<<(unsigned)ctas, 128, smem, stream>>>(M, A);
  return cudaGetLastError();
}

// per-ray sums -> gradients w.r.t. the op's inputs (DS/emitters/nerf_op.py:76-89 transposed,
// nerf_emitter_op.py:112 change of variables, mitsuba_sdf.py:1206-1214 non-finite scrub)
__global__ void k_finalize_ray_grads(const __grid_constant__ ModelDev M, const __grid_constant__ FinalizeArgs A) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < A.n; r += (int64_t)gridDim.x * blockDim.x) {
    const float* g = A.ray_grad + r * 6;
    const int64_t R = A.ray_offset + r;
    const float nr = A.near_per_ray ? A.near_per_ray[R] : A.near_scalar;
    if (A.nerf_layout) {
      // backward_for_camera_ray_bundle leaves (N,3) gradients in ray_bundle.origins.grad / .directions.grad
      // (NS/pipelines/mitsuba_sdf.py:1144-1219), NeRF coordinates, inf/nan scrubbed (:1206-1214)
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        float a = g[k], b = g[3 + k];
        if (!isfinite(a)) a = 0.f;
        if (!isfinite(b)) b = 0.f;
        A.grad_o[R * 3 + k] = a;
        if (A.grad_v) A.grad_v[R * 3 + k] = b + nr * a;
      }
      continue;
    }
    // nerf (x,y,z) = mitsuba (z,x,y):  g_mi[0] = g_nerf[1], g_mi[1] = g_nerf[2], g_mi[2] = g_nerf[0]
    float go[3] = {g[1], g[2], g[0]}, gv[3] = {g[4], g[5], g[3]};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float a = 2.f * M.scene_scale * go[k];
      float b = gv[k];
      if (!isfinite(a)) a = 0.f;
      if (!isfinite(b)) b = 0.f;
      A.grad_o[k * A.n_total + R] = a - nr * b;
      if (A.grad_v) A.grad_v[k * A.n_total + R] = b;
    }
  }
}

cudaError_t launch_finalize_ray_grads(const ModelDev& M, const FinalizeArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  int64_t blocks = (A.n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_finalize_ray_grads<<<(unsigned)blocks, 256, 0, stream>>>(M, A);
  return cudaGetLastError();
}

}  // namespace nemit
