// components.cu - component-level drop-ins behind the nerfstudio module API (SURVEY.md 8b "Model-side API kept
// intact"): the kernels that Field.forward / Field.density_fn / HashMLPDensityField.density_fn / SHEncoding.forward /
// MLP.forward dispatch to when they are called on their own (outside the fused emitter query).  Exact fp32
// arithmetic, same device functions as the fused kernels (common.cuh, field_common.cuh, sampling.cuh).
//
//   k_field_samples : NerfactoField.forward(ray_samples) = get_density + get_outputs
//                     (NS/fields/nerfacto_field.py:228-334, NS/fields/base_field.py:124-143) on explicit sample
//                     positions / directions; also Field.density_fn for the main field (base_field.py:49-69)
//   k_prop_density  : HashMLPDensityField.get_density / density_fn (NS/fields/density_fields.py:95-122)
//   k_sh_encode     : SHEncoding.pytorch_fwd (NS/field_components/encodings.py:708-716, NS/utils/math.py:27-90)
//   k_mlp           : MLP.pytorch_fwd (NS/field_components/mlp.py:163-183), widths <= 64
#include "kernels.h"
#include "field_common.cuh"
#include "sampling.cuh"

namespace nemit {

struct SamplesSmem {
  FieldWeightsSmem W;
  float bufA[kHidden * kTileM];
  float bufB[kHidden * kTileM];
  float app_tmp[kApp];
};

// CTA = 128 threads = one tile of 128 samples at a time (thread m owns sample m); the same register-tiled fp32
// GEMM chain as k_field_forward, without the per-ray composite.
__global__ void __launch_bounds__(128)
k_field_samples(const __grid_constant__ ModelDev M, const __grid_constant__ FieldParams P, const float* __restrict__ pos,
                const float* __restrict__ dir, int64_t n, int app_row, float* __restrict__ density,
                float* __restrict__ geo, float* __restrict__ rgb) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SamplesSmem* sm = reinterpret_cast<SamplesSmem*>(smem_raw);
  const int tid = threadIdx.x;
  load_field_weights(&sm->W, P, M.num_images, app_row, sm->app_tmp, tid, 128);
  const int64_t tiles = (n + kTileM - 1) / kTileM;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t s = tile * kTileM + tid;
    const bool live = s < n;
    float x[3] = {0.f, 0.f, 0.f}, p[3];
    if (live) { x[0] = pos[s * 3]; x[1] = pos[s * 3 + 1]; x[2] = pos[s * 3 + 2]; }
    const bool sel = contract_normalise(x, M, p) && live;
    __syncthreads();
    main_encode_to_tile(p, P.main_table, M.main_grid, sm->bufA, tid);
    __syncthreads();
    tile_gemm<kMainEnc, kHidden, true>(sm->bufA, sm->W.base_w0, sm->W.base_b0, sm->bufB, tid);
    __syncthreads();
    tile_gemm<kHidden, 16, false>(sm->bufB, sm->W.base_w1, sm->W.base_b1, sm->bufA + 15 * kTileM, tid);
    __syncthreads();
    {
      const float raw = sm->bufA[15 * kTileM + tid];
      if (live) density[s] = sel ? expf(raw) : 0.f;           // trunc_exp * selector (nerfacto_field.py:247-249)
      if (live && geo) {
#pragma unroll
        for (int k = 0; k < kGeo; ++k) geo[s * kGeo + k] = sm->bufA[(16 + k) * kTileM + tid];
      }
    }
    if (!rgb) continue;                                        // density_fn only (uniform over the CTA)
    {
      float d[3] = {0.f, 0.f, 1.f}, c[16];
      if (live) { d[0] = dir[s * 3]; d[1] = dir[s * 3 + 1]; d[2] = dir[s * 3 + 2]; }
      sh16((d[0] + 1.f) * 0.5f, (d[1] + 1.f) * 0.5f, (d[2] + 1.f) * 0.5f, c);    // get_normalized_directions
#pragma unroll
      for (int k = 0; k < 16; ++k) sm->bufA[k * kTileM + tid] = c[k];
      sm->bufA[31 * kTileM + tid] = 0.f;
    }
    __syncthreads();
    tile_gemm<32, kHidden, true>(sm->bufA, sm->W.head_w0, sm->W.head_b0_eff, sm->bufB, tid);
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
      if (live) {
        rgb[s * 3 + 0] = expf(fminf(a0 + M.rgb_bias, 88.f));   // SafeExp (nerfacto_field.py:40-57)
        rgb[s * 3 + 1] = expf(fminf(a1 + M.rgb_bias, 88.f));
        rgb[s * 3 + 2] = expf(fminf(a2 + M.rgb_bias, 88.f));
      }
    }
  }
}

cudaError_t launch_field_samples(const ModelDev& M, const FieldParams& P, const float* pos, const float* dir, int64_t n,
                                 int app_row, float* density, float* geo, float* rgb,
                                 cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  const size_t smem = sizeof(SamplesSmem);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_samples, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int64_t tiles = (n + kTileM - 1) / kTileM;
  const int64_t ctas = tiles < 148 * 2 ? tiles : 148 * 2;
  k_field_samples<<<(unsigned)ctas, 128, smem, stream>>>(M, P, pos, dir, n, app_row, density, geo, rgb);
  return cudaGetLastError();
}

// one thread per point; the 10 -> 16 -> 1 MLP sits in shared memory
__global__ void __launch_bounds__(256)
k_prop_density(const __grid_constant__ ModelDev M, const __grid_constant__ GridDev G, const float2* __restrict__ table,
               const float* __restrict__ w0, const float* __restrict__ b0, const float* __restrict__ w1,
               const float* __restrict__ b1, const float* __restrict__ pos, int64_t n, float* __restrict__ density) {
  __shared__ PropMlpSmem W;
  load_prop_mlp(&W, w0, b0, w1, b1, threadIdx.x, blockDim.x);
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    density[i] = prop_density(pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2], table, G, &W, M);
}

cudaError_t launch_prop_density(const ModelDev& M, const GridDev& G, const float2* table, const float* w0,
                                const float* b0, const float* w1, const float* b1, const float* pos, int64_t n,
                                float* density, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_prop_density<<<(unsigned)blocks, 256, 0, stream>>>(M, G, table, w0, b0, w1, b1, pos, n, density);
  return cudaGetLastError();
}

__global__ void k_sh_encode(const float* __restrict__ d, int64_t n, int levels, float* __restrict__ out) {
  const int dim = levels * levels;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float c[16];
    sh16(d[i * 3], d[i * 3 + 1], d[i * 3 + 2], c);
    for (int k = 0; k < dim; ++k) out[i * dim + k] = c[k];
  }
}

cudaError_t launch_sh_encode(const float* d, int64_t n, int levels, float* out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_sh_encode<<<(unsigned)blocks, 256, 0, stream>>>(d, n, levels, out);
  return cudaGetLastError();
}

// generic small MLP: one thread per row, activations ping-pong between two 64-wide local arrays, weights read
// through the read-only path (every lane reads the same address: one broadcast transaction per warp)
__global__ void __launch_bounds__(128)
k_mlp(const __grid_constant__ MlpDev L, const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float a[NEMIT_MLP_MAX_WIDTH], b[NEMIT_MLP_MAX_WIDTH];
    for (int k = 0; k < L.in_dim; ++k) a[k] = x[i * L.in_dim + k];
    int in_dim = L.in_dim;
    for (int l = 0; l < L.num_layers; ++l) {
      const int od = L.out_dim[l];
      const bool last = l == L.num_layers - 1;
      const float* __restrict__ w = L.weight[l];
      const float* __restrict__ bias = L.bias[l];
      float* src = (l & 1) ? b : a;
      float* dst = (l & 1) ? a : b;
      for (int o = 0; o < od; ++o) {
        float acc = bias ? __ldg(bias + o) : 0.f;
        for (int k = 0; k < in_dim; ++k) acc = fmaf(__ldg(w + o * in_dim + k), src[k], acc);
        const int act = last ? L.out_activation : L.activation;
        if (act == 1) acc = fmaxf(acc, 0.f);
        dst[o] = acc;
      }
      in_dim = od;
    }
    const float* res = (L.num_layers & 1) ? b : a;
    for (int o = 0; o < in_dim; ++o) out[i * in_dim + o] = res[o];
  }
}

// NeRFEmitter.eval_nerf up to the op call (DS/emitters/nerf.py:69-82): rays from world to emitter-local space
// (o' = A o + t as a point, v' = A v as a vector, not renormalised: mi.Transform4f @ mi.Ray3f) and the distance to skip,
// sph_t = far slab intersection with the padded unit box when the ray hits it (outside_only), else skip_distance.
// mode 1 applies only the linear part to both arrays (the transpose pass that carries o / v gradients back to world).
struct LocalRayArgs {
  float m[12];             // rows of the 3x4 world -> local matrix
  float bmin[3], bmax[3];
  float skip_distance;
  int outside_only;
  int mode;
};

__global__ void __launch_bounds__(256)
k_emitter_local_rays(const __grid_constant__ LocalRayArgs A, const float* __restrict__ o, const float* __restrict__ v,
                     int64_t n, float* __restrict__ o_out, float* __restrict__ v_out, float* __restrict__ near_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float ox = o[i], oy = o[n + i], oz = o[2 * n + i];
    const float vx = v[i], vy = v[n + i], vz = v[2 * n + i];
    float lo[3], lv[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float tr = A.mode == 1 ? 0.f : A.m[4 * r + 3];
      lo[r] = fmaf(A.m[4 * r], ox, fmaf(A.m[4 * r + 1], oy, fmaf(A.m[4 * r + 2], oz, tr)));
      lv[r] = fmaf(A.m[4 * r], vx, fmaf(A.m[4 * r + 1], vy, A.m[4 * r + 2] * vz));
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) { o_out[r * n + i] = lo[r]; v_out[r * n + i] = lv[r]; }
    if (near_out) {
      float t = A.skip_distance;
      if (A.outside_only) {
        // mi.BoundingBox3f.ray_intersect (mitsuba 3.4.1 include/mitsuba/core/bbox.h): slab test, returns the hit flag
        // and the near / far parameters; nerf.py:76 keeps the far one
        bool hit = true;
        float mint = -INFINITY, maxt = INFINITY;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          hit = hit && (lv[r] != 0.f || lo[r] > A.bmin[r] || lo[r] < A.bmax[r]);
          const float rcp = 1.f / lv[r];
          const float t1 = (A.bmin[r] - lo[r]) * rcp, t2 = (A.bmax[r] - lo[r]) * rcp;
          mint = fmaxf(mint, fminf(t1, t2));
          maxt = fminf(maxt, fmaxf(t1, t2));
        }
        hit = hit && maxt >= mint;
        if (hit) t = maxt;
      }
      near_out[i] = t;
    }
  }
}

cudaError_t launch_emitter_local_rays(const float* m34, const float* bmin, const float* bmax, float skip_distance,
                                      int outside_only, int mode, const float* o, const float* v, int64_t n,
                                      float* o_out, float* v_out, float* near_out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  LocalRayArgs A;
  for (int i = 0; i < 12; ++i) A.m[i] = m34[i];
  for (int i = 0; i < 3; ++i) { A.bmin[i] = bmin[i]; A.bmax[i] = bmax[i]; }
  A.skip_distance = skip_distance; A.outside_only = outside_only; A.mode = mode;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_emitter_local_rays<<<(unsigned)blocks, 256, 0, stream>>>(A, o, v, n, o_out, v_out, near_out);
  return cudaGetLastError();
}

cudaError_t launch_mlp(const MlpDev& L, const float* x, int64_t n, float* out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int64_t blocks = (n + 127) / 128;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_mlp<<<(unsigned)blocks, 128, 0, stream>>>(L, x, n, out);
  return cudaGetLastError();
}

}  // namespace nemit
