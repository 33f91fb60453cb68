// sampling.cu - k_sample_levels: the proposal loop of one ray per warp.
// Output: the 49 spacing / euclidean edges of the final 48 samples (+ optional per-level outputs).
#include "sampling.cuh"
#include "kernels.h"

namespace nemit {

constexpr int kWarpsPerCta = 8;

template <class MLP>
struct SampleSmem {
  MLP mlp[2];
  WarpScratch warp[kWarpsPerCta];
};

// MLP = PropMlpSmem: exact fp32 sampler (parity mode); PropMlpFastSmem: packed-half proposal networks (sampling.cuh)
template <class MLP>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 3)   // 80 registers: measured best for both variants (4 -> spills, 2 -> too few warps)
k_sample_levels(const __grid_constant__ ModelDev M, const __grid_constant__ SampleArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SampleSmem<MLP>* sm = reinterpret_cast<SampleSmem<MLP>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = 0; i < 2; ++i) {
    if constexpr (cuda::std::is_same<MLP, PropMlpSmem>::value)
      load_prop_mlp(&sm->mlp[i], A.prop_w0[i], A.prop_b0[i], A.prop_w1[i], A.prop_b1[i], tid, blockDim.x);
    else
      load_prop_mlp_fast(&sm->mlp[i], A.prop_w0[i], A.prop_b0[i], A.prop_w1[i], A.prop_b1[i], tid, blockDim.x);
  }
  __syncthreads();
  WarpScratch* sc = &sm->warp[wid];

#pragma unroll 1
  for (int64_t r = (int64_t)blockIdx.x * kWarpsPerCta + wid; r < A.n; r += (int64_t)gridDim.x * kWarpsPerCta) {
    RayCtx R;
    load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + r, ray_scale(M, A), R.o, R.d);
    const float near = A.training ? M.near_plane : 0.f;      // NearFarCollider, scene_colliders.py:217-222
    R.s_near = spacing_fwd(near);
    R.s_far = spacing_fwd(M.far_plane);
    const float* jit = A.jitter ? A.jitter + (A.ray_offset + r) * 3 : nullptr;
    initial_edges(sc, A.lin0, jit ? jit[0] : -1.f, lane);

    // levels 0 and 1 run through ONE copy of the code (instruction-cache footprint, see sampling.cuh)
#pragma unroll 1
    for (int lvl = 0; lvl < 2; ++lvl) {
      const int S = lvl == 0 ? kS0 : kS1;          // bins at this level
      const int Mn = lvl == 0 ? kS1 : kS2;         // samples of the next level
      proposal_level_weights(sc, R, S, A.prop_table[lvl], M.prop_grid[lvl], &sm->mlp[lvl], M, lane, nullptr, A.prop_table_h[lvl],
                             A.rot.mode ? &A.rot : nullptr);
      float* w_out = lvl == 0 ? A.weights0 : A.weights1;
      if (w_out) {
        for (int i = lane; i < S; i += 32) w_out[r * S + i] = sc->buf[i];
      }
      int32_t* inds = lvl == 0 ? A.inds1 : A.inds2;
      pdf_resample_warp(sc, S, Mn, lvl == 0 ? A.u1 : A.u2, jit ? jit[lvl + 1] : -1.f, A.anneal, M.cpu_sum_width,
                        lane, inds ? inds + r * (Mn + 1) : nullptr);
      if (lvl == 0) {
        for (int i = lane; i <= kS1; i += 32) {
          float e = sc->s_new[i];
          sc->s_edge[i] = e;
          if (A.s_edges1) A.s_edges1[r * (kS1 + 1) + i] = e;
        }
        __syncwarp();
      }
    }
    // ---- level 2 edges ---------------------------------------------------------------------
    for (int i = lane; i <= kS2; i += 32) {
      float s = sc->s_new[i];
      if (A.s_edges2) A.s_edges2[r * (kS2 + 1) + i] = s;
      if (A.t_edges2) A.t_edges2[r * (kS2 + 1) + i] = s_to_t(s, R.s_near, R.s_far);
    }
    __syncwarp();
  }
}

// fp32 (2 features) -> packed half2 copy of both proposal tables, one pass of 10 MB read + 5 MB written per sampler launch
// (the parameters may have changed since the last call; the library cannot know)
__global__ void k_prop_tables_half(const float2* __restrict__ t0, const float2* __restrict__ t1, uint32_t* __restrict__ h0,
                                   uint32_t* __restrict__ h1, int n0, int n1) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n0 + n1; i += stride) {
    const float2 v = i < n0 ? __ldg(t0 + i) : __ldg(t1 + (i - n0));
    const __half2 hv = __float22half2_rn(v);
    (i < n0 ? h0[i] : h1[i - n0]) = *reinterpret_cast<const uint32_t*>(&hv);
  }
}

template <class MLP>
static cudaError_t launch_sample_levels_t(const ModelDev& M, const SampleArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(SampleSmem<MLP>);
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_sample_levels<MLP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t ctas = (A.n + kWarpsPerCta - 1) / kWarpsPerCta;
  const int64_t max_ctas = 148 * 8 * 16;
  if (ctas > max_ctas) ctas = max_ctas;
  k_sample_levels<MLP><<<(unsigned)ctas, kWarpsPerCta * 32, smem, stream>>>(M, A);
  return cudaGetLastError();
}
cudaError_t launch_sample_levels(const ModelDev& M, const SampleArgs& A, cudaStream_t stream, bool fast) {
  if (fast && A.n > 0) {
    if (!A.prop_table_h[0] || !A.prop_table_h[1]) return cudaErrorInvalidValue;
    const int n0 = M.prop_grid[0].num_levels * (int)M.prop_grid[0].table_size, n1 = M.prop_grid[1].num_levels * (int)M.prop_grid[1].table_size;
    k_prop_tables_half<<<148 * 4, 256, 0, stream>>>(A.prop_table[0], A.prop_table[1], A.prop_table_h[0], A.prop_table_h[1], n0, n1);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return fast ? launch_sample_levels_t<PropMlpFastSmem>(M, A, stream) : launch_sample_levels_t<PropMlpSmem>(M, A, stream);
}

}  // namespace nemit
