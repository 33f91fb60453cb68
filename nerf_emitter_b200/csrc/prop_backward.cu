// prop_backward.cu - k_prop_backward: gradient of the interlevel (proposal) loss w.r.t. the two proposal
// networks, one ray per warp.  The proposal densities are recomputed in fp32 at the sampling pass's own edges, then per
// level: outer-measure bound of the final histogram by the proposal histogram, its adjoint into the
// proposal weights, the weights -> density adjoint, the 10->16->1 MLP backward and the hash-table
// scatter-add.  Rays are detached on this path (NS/models/sdf_nerfacto.py:121-126): no ray gradients.
//
// Reference: interlevel_loss / lossfun_outer / outer (NS/model_components/losses.py:52-131),
// HashMLPDensityField.get_density (NS/fields/density_fields.py:95-122), get_weights
// (NS/cameras/rays.py:128-150), objective NS/pipelines/mitsuba_sdf.py:1170-1177.
#include "sampling.cuh"
#include "kernels.h"

namespace nemit {

constexpr int kPbWarps = 8;
constexpr int kStageStride = 45;     // per-sample staging row: enc[10] | 1 | gh[16] | g_raw | relu(h)[16]
constexpr int kPropParams = 193;     // 160 W0 + 16 b0 + 16 W1 + 1 b1
constexpr int kAccPerLane = 7;       // ceil(193 / 32)
constexpr float kEpsLoss = 1.0e-7f;  // NS/model_components/losses.py:42

struct PbWarp {
  WarpScratch sc;
  float ds[kS0 + 4];       // delta*sigma per sample
  float gw[kS0 + 4];       // difference array -> dL/dw -> dL/dsigma
  float c2[kS2 + 4];       // spacing edges of the final samples
  float w2[kS2];           // final weights
  float stage[32 * kStageStride];
};

struct PbSmem {
  PropMlpSmem mlp[2];
  float red[2][kPropParams + 3];
  PbWarp warp[kPbWarps];
};

// first index in a[0..n) with a[idx] > v  (torch.searchsorted(side="right"))
__device__ __forceinline__ int ss_right(const float* a, int n, float v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (a[mid] <= v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// hidden activations (pre-ReLU) and raw density of the proposal MLP
__device__ __forceinline__ float prop_mlp_hidden(const float enc[kPropEnc], const PropMlpSmem* W, float h[kPropHidden]) {
  float raw = W->b1;
#pragma unroll
  for (int j = 0; j < kPropHidden; ++j) {
    const float4* r4 = reinterpret_cast<const float4*>(W->row[j]);
    float4 a = r4[0], b = r4[1], c = r4[2];
    float v = c.z;
    v = fmaf(a.x, enc[0], v); v = fmaf(a.y, enc[1], v); v = fmaf(a.z, enc[2], v); v = fmaf(a.w, enc[3], v);
    v = fmaf(b.x, enc[4], v); v = fmaf(b.y, enc[5], v); v = fmaf(b.z, enc[6], v); v = fmaf(b.w, enc[7], v);
    v = fmaf(c.x, enc[8], v); v = fmaf(c.y, enc[9], v);
    h[j] = v;
    raw = fmaf(c.w, fmaxf(v, 0.f), raw);
  }
  return raw;
}

// adjoint of one level for one ray.  In: edges sc.s_edge[0..S], weights sc.buf[0..S), ds[0..S).
// acc: this lane's share of the 193 parameter-gradient sums.
static __device__ __noinline__ void prop_level_backward(PbWarp* pw, const RayCtx& R, int S, const float2* __restrict__ table,
                                                 float2* __restrict__ g_table, const GridDev& G,
                                                 const PropMlpSmem* W, const ModelDev& M, float coef, float loss_scale,
                                                 float (&acc)[kAccPerLane], int lane, const RotArgs& Q) {
  WarpScratch* sc = &pw->sc;
  // cumulative proposal weights cy[0..S] (losses.py:66) and a zeroed difference array
  float carry = 0.f;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float inc = warp_incl_scan_f(sc->buf[i], lane);
    sc->cdf[i + 1] = carry + inc;
    carry += __shfl_sync(0xffffffffu, inc, 31);
    pw->gw[i] = 0.f;
  }
  if (lane == 0) { sc->cdf[0] = 0.f; pw->gw[S] = 0.f; }
  __syncwarp();
  // w_outer per final interval and d loss / d w_outer scattered as +g at lo, -g after hi
  bool any = false;
#pragma unroll 1
  for (int i = lane; i < kS2; i += 32) {
    int lo = ss_right(sc->s_edge, S, pw->c2[i]) - 1;
    lo = min(max(lo, 0), S - 1);
    int hi = ss_right(sc->s_edge + 1, S, pw->c2[i + 1]);
    hi = min(max(hi, 0), S - 1);
    float wo = sc->cdf[hi + 1] - sc->cdf[lo];
    float w = pw->w2[i];
    float diff = fmaxf(w - wo, 0.f);
    if (diff > 0.f) {
      float g = -2.f * diff / (w + kEpsLoss) * coef;
      atomicAdd(&pw->gw[lo], g);
      atomicAdd(&pw->gw[hi + 1], -g);
      any = true;
    }
  }
  if (!__any_sync(0xffffffffu, any)) return;      // the proposal bounds the final histogram: zero gradient
  __syncwarp();
  // dL/dw_j = prefix sum of the difference array; also inclusive scan of gw*w for the transmittance term
  float c1 = 0.f, c2s = 0.f;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float inc = warp_incl_scan_f(pw->gw[i], lane);
    float gwi = c1 + inc;
    c1 += __shfl_sync(0xffffffffu, inc, 31);
    float inc2 = warp_incl_scan_f(gwi * sc->buf[i], lane);
    sc->cdf[i] = c2s + inc2;                       // inclusive prefix of gw*w
    c2s += __shfl_sync(0xffffffffu, inc2, 31);
    pw->gw[i] = gwi;
  }
  const float total = c2s;
  __syncwarp();
  // weights -> delta*sigma -> sigma adjoint (rays.py:128-150)
  double carry_d = 0.0;
#pragma unroll 1
  for (int i = lane; i < S; i += 32) {
    float ds = pw->ds[i];
    double inc = warp_incl_scan((double)ds, lane);
    double exc = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) exc = 0.0;
    float T = expf(-(float)(carry_d + exc));
    carry_d += __shfl_sync(0xffffffffu, inc, 31);
    float gds = pw->gw[i] * T * expf(-ds) - (total - sc->cdf[i]);
    float t0 = s_to_t(sc->s_edge[i], R.s_near, R.s_far), t1 = s_to_t(sc->s_edge[i + 1], R.s_near, R.s_far);
    pw->gw[i] = gds * sub_rn(t1, t0);              // dL/dsigma_i
  }
  __syncwarp();
  // per sample: recompute the field, MLP backward, table scatter; parameter sums through the staging tile
#pragma unroll 1
  for (int i0 = 0; i0 < S; i0 += 32) {
    const int i = i0 + lane;
    float gs = pw->gw[i];
    float t0 = s_to_t(sc->s_edge[i], R.s_near, R.s_far), t1 = s_to_t(sc->s_edge[i + 1], R.s_near, R.s_far);
    float tsum = add_rn(t0, t1);
    float x[3], dirq[3];
    rot_sample(Q, R.o, R.d, tsum, x, dirq);       // the proposal samples are rotated like every level's (rays.py:285-286)
    float p[3];
    bool sel = contract_normalise(x, M, p);
    bool live = sel && gs != 0.f;
    if (!__any_sync(0xffffffffu, live)) continue;
    float* row = pw->stage + lane * kStageStride;
    if (live) {
      LevelCell cell[kPropLevels];
      float enc[kPropEnc], h[kPropHidden];
      {
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
          enc[2 * l] = r.x; enc[2 * l + 1] = r.y;
        }
      }
      float raw = prop_mlp_hidden(enc, W, h);
      float graw = gs * expf(fminf(fmaxf(raw, -15.f), 15.f));       // trunc_exp backward
      float genc[kPropEnc];
#pragma unroll
      for (int k = 0; k < kPropEnc; ++k) { genc[k] = 0.f; row[k] = enc[k]; }
      row[10] = 1.f;
      row[27] = graw;
#pragma unroll
      for (int j = 0; j < kPropHidden; ++j) {
        float gh = h[j] > 0.f ? W->row[j][11] * graw : 0.f;
        row[11 + j] = gh;
        row[28 + j] = fmaxf(h[j], 0.f);
#pragma unroll
        for (int k = 0; k < kPropEnc; ++k) genc[k] = fmaf(W->row[j][k], gh, genc[k]);
      }
#pragma unroll
      for (int l = 0; l < kPropLevels; ++l) {
        float cw[8];
        corner_weights(cell[l], cw);
        float gx = genc[2 * l] * loss_scale, gy = genc[2 * l + 1] * loss_scale;
#pragma unroll
        for (int c = 0; c < 8; ++c) atomic_add_f2(g_table + cell[l].idx[c], make_float2(cw[c] * gx, cw[c] * gy));
      }
    } else {
#pragma unroll 1
      for (int k = 0; k < 44; ++k) row[k] = 0.f;
    }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < kAccPerLane; ++q) {
      int e = q * 32 + lane;
      if (e < kPropParams) {
        int ca, cb;
        if (e < 160) { ca = 11 + e / 10; cb = e % 10; }
        else if (e < 176) { ca = 11 + (e - 160); cb = 10; }
        else if (e < 192) { ca = 27; cb = 28 + (e - 176); }
        else { ca = 27; cb = 10; }
        float s = 0.f;
#pragma unroll 8
        for (int t = 0; t < 32; ++t) s = fmaf(pw->stage[t * kStageStride + ca], pw->stage[t * kStageStride + cb], s);
        acc[q] += s;
      }
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(kPbWarps * 32)
k_prop_backward(const __grid_constant__ ModelDev M, const __grid_constant__ PropBwdArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PbSmem* sm = reinterpret_cast<PbSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = 0; i < 2; ++i)
    load_prop_mlp(&sm->mlp[i], A.prop_w0[i], A.prop_b0[i], A.prop_w1[i], A.prop_b1[i], tid, blockDim.x);
  for (int i = tid; i < 2 * (kPropParams + 3); i += blockDim.x) (&sm->red[0][0])[i] = 0.f;
  __syncthreads();
  PbWarp* pw = &sm->warp[wid];
  WarpScratch* sc = &pw->sc;
  float acc[2][kAccPerLane];
#pragma unroll
  for (int l = 0; l < 2; ++l)
#pragma unroll
    for (int q = 0; q < kAccPerLane; ++q) acc[l][q] = 0.f;

  const int64_t chunk = A.reg_chunk > 0 ? A.reg_chunk : A.n_total;
#pragma unroll 1
  for (int64_t r = (int64_t)blockIdx.x * kPbWarps + wid; r < A.n; r += (int64_t)gridDim.x * kPbWarps) {
    RayCtx R;
    const int64_t rg = A.ray_offset + r;
    load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, rg, ray_scale(M, A), R.o, R.d);
    const float near = A.training ? M.near_plane : 0.f;
    R.s_near = spacing_fwd(near);
    R.s_far = spacing_fwd(M.far_plane);
    const float* jit = A.jitter ? A.jitter + rg * 3 : nullptr;
    initial_edges(sc, A.lin0, jit ? jit[0] : -1.f, lane);
    for (int i = lane; i <= kS2; i += 32) pw->c2[i] = A.s_edges2_in[r * (kS2 + 1) + i];
    for (int i = lane; i < kS2; i += 32) pw->w2[i] = A.weights2[r * kS2 + i];
    __syncwarp();
    // mean over (rays of the backward chunk) x 48 intervals, times the loss multiplier (losses.py:112-131)
    const int64_t c0 = (rg / chunk) * chunk;
    const float cnt = (float)min(chunk, A.n_total - c0);
    const float coef = M.interlevel_mult / (cnt * (float)kS2);
#pragma unroll 1
    for (int lvl = 0; lvl < 2; ++lvl) {
      const int S = lvl == 0 ? kS0 : kS1;
      proposal_level_weights(sc, R, S, A.prop_table[lvl], M.prop_grid[lvl], &sm->mlp[lvl], M, lane, pw->ds, nullptr,
                             A.rot.mode ? &A.rot : nullptr);
      if (lvl == 0) {
        prop_level_backward(pw, R, S, A.prop_table[0], A.g_table[0], M.prop_grid[0], &sm->mlp[0], M, coef,
                            A.loss_scale, acc[0], lane, A.rot);
        __syncwarp();
        // level-1 edges: the ones the sampling pass of this call produced (so the proposal histogram bounded here is the
        // one the final samples were drawn from, whichever sampler arithmetic ran), not a second PDF stage
        for (int i = lane; i <= kS1; i += 32) sc->s_edge[i] = A.s_edges1_in[r * (kS1 + 1) + i];
        __syncwarp();
      } else {
        prop_level_backward(pw, R, S, A.prop_table[1], A.g_table[1], M.prop_grid[1], &sm->mlp[1], M, coef,
                            A.loss_scale, acc[1], lane, A.rot);
        __syncwarp();
      }
    }
  }
  // CTA reduction of the parameter sums, then one atomic per parameter per CTA
#pragma unroll
  for (int l = 0; l < 2; ++l)
#pragma unroll
    for (int q = 0; q < kAccPerLane; ++q) {
      int e = q * 32 + lane;
      if (e < kPropParams && acc[l][q] != 0.f) atomicAdd(&sm->red[l][e], acc[l][q]);
    }
  __syncthreads();
  for (int i = tid; i < 2 * kPropParams; i += blockDim.x) {
    int l = i / kPropParams, e = i % kPropParams;
    float g = sm->red[l][e] * A.loss_scale;
    if (g == 0.f) continue;
    if (e < 160) atomicAdd(A.g_w0[l] + e, g);                // (16,10) row-major: e = j*10 + k
    else if (e < 176) atomicAdd(A.g_b0[l] + (e - 160), g);
    else if (e < 192) atomicAdd(A.g_w1[l] + (e - 176), g);
    else atomicAdd(A.g_b1[l], g);
  }
}

cudaError_t launch_prop_backward(const ModelDev& M, const PropBwdArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(PbSmem);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_prop_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t ctas = (A.n + kPbWarps - 1) / kPbWarps;
  const int64_t max_ctas = 148 * 2 * 8;
  if (ctas > max_ctas) ctas = max_ctas;
  k_prop_backward<<<(unsigned)ctas, kPbWarps * 32, smem, stream>>>(M, A);
  return cudaGetLastError();
}

}  // namespace nemit
