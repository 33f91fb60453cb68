// gmm.cu - light point cloud -> 64 spherical Gaussians, the producer of the vMF lobe table:
//   k_lum_sum + k_compensate : compensate_pc (NS/model_components/output_light_pc.py:176-191): weight_i =
//                              max(lum_i - mean(lum), 0), zero unless lum_i > threshold.  Points are not compacted:
//                              a zero weight drops a point from every sum below.
//   k_gmm_em + k_gmm_update  : one EM iteration of cluster_light's GeneralMixtureModel([Normal(covariance_type=
//                              'sphere')] * K).fit(position, sample_weight=weight)
//                              (NS/model_components/gmm_cluster_light.py:41-85).  pomegranate is a third-party
//                              dependency that is not in the reference tree (README.md install list, unpinned):
//                              PARITY UNPINNED; restated as the textbook weighted EM for a mixture of isotropic
//                              normals (responsibilities r_ik ~ pi_k N(x_i | mu_k, s_k^2 I); mu_k = sum w r x / sum w r;
//                              s_k^2 = sum w r |x - mu_k|^2 / (3 sum w r); pi_k = sum w r / sum w; objective
//                              sum_i w_i log sum_k pi_k N(x_i | ...)).
// k_gmm_em is a stream over the points (16 B/point) with 64 x ~12 flops each: SFU/ALU-bound on exp/log; the K x 6
// sufficient statistics are reduced in shared memory per CTA and leave as fp64 atomics.
#include "kernels.h"

namespace nemit {

constexpr int kGmmMaxK = 128;

__global__ void __launch_bounds__(256)
k_lum_sum(const float* __restrict__ lum, int64_t n, double* __restrict__ out) {
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) acc += lum[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int w = 0; w < 8; ++w) v += red[w];
    atomicAdd(out, v);
  }
}

__global__ void k_compensate(const float* __restrict__ lum, int64_t n, const double* __restrict__ sum, float threshold,
                             int use_threshold, float* __restrict__ w) {
  const float mean = (float)(sum[0] / (double)n);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float l = lum[i];
    float c = fmaxf(l - mean, 0.f);                       // torch.clamp(lum - lum.mean(), 0.)
    if (use_threshold && !(l > threshold)) c = 0.f;       // mask &= lum > threshold
    w[i] = c;
  }
}

// stats layout (fp64): [k*6 + {0: sum wr, 1..3: sum wr (x - mu_k), 4: sum wr |x - mu_k|^2, 5: unused}], [K*6] = sum w,
// [K*6+1] = sum w logp   (mu_k = the iteration's input mean)
__global__ void __launch_bounds__(256)
k_gmm_em(const __grid_constant__ GmmArgs A) {
  __shared__ float4 comp[kGmmMaxK];       // mu.xyz, -1 / (2 s^2)
  __shared__ float lognorm[kGmmMaxK];     // log pi - 1.5 log(2 pi s^2)
  __shared__ float acc[kGmmMaxK][5];
  __shared__ float acc_w, acc_lp;
  const int K = A.k;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float var = fmaxf(A.var[k], 1e-20f);
    comp[k] = make_float4(A.mean[k * 3], A.mean[k * 3 + 1], A.mean[k * 3 + 2], -0.5f / var);
    lognorm[k] = (A.prior[k] > 0.f ? logf(A.prior[k]) : -INFINITY) - 1.5f * logf(6.283185307179586f * var);
    for (int j = 0; j < 5; ++j) acc[k][j] = 0.f;
  }
  if (threadIdx.x == 0) { acc_w = 0.f; acc_lp = 0.f; }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  float tw = 0.f, tlp = 0.f;
  for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ll; i0 < A.n; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = i0 + lane;
    const bool live = i < A.n;
    const float w = live ? A.w[i] : 0.f;
    const float x = live ? A.x[i * 3] : 0.f, y = live ? A.x[i * 3 + 1] : 0.f, z = live ? A.x[i * 3 + 2] : 0.f;
    if (!__any_sync(0xffffffffu, w > 0.f)) continue;
    // log-sum-exp over the components
    float m = -INFINITY;
    for (int k = 0; k < K; ++k) {
      const float4 c = comp[k];
      const float dx = x - c.x, dy = y - c.y, dz = z - c.z;
      m = fmaxf(m, fmaf(dx * dx + dy * dy + dz * dz, c.w, lognorm[k]));
    }
    float s = 0.f;
    for (int k = 0; k < K; ++k) {
      const float4 c = comp[k];
      const float dx = x - c.x, dy = y - c.y, dz = z - c.z;
      s += __expf(fmaf(dx * dx + dy * dy + dz * dz, c.w, lognorm[k]) - m);
    }
    const float inv = w > 0.f ? w / s : 0.f;
    if (w > 0.f) { tw += w; tlp += w * (m + logf(s)); }
    // M-step sums, centred on the CURRENT mean of each component (no cancellation in the variance): each component's five
    // statistics are reduced over the warp, then one shared atomic per warp
    for (int k = 0; k < K; ++k) {
      const float4 c = comp[k];
      const float dx = x - c.x, dy = y - c.y, dz = z - c.z;
      const float d2 = dx * dx + dy * dy + dz * dz;
      float r = inv * __expf(fmaf(d2, c.w, lognorm[k]) - m);
      float v[5] = {r, r * dx, r * dy, r * dz, r * d2};
#pragma unroll
      for (int j = 0; j < 5; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[j] += __shfl_xor_sync(0xffffffffu, v[j], o);
      }
      if (lane < 5 && v[0] != 0.f) {
        const float mine = lane == 0 ? v[0] : lane == 1 ? v[1] : lane == 2 ? v[2] : lane == 3 ? v[3] : v[4];
        atomicAdd(&acc[k][lane], mine);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { tw += __shfl_xor_sync(0xffffffffu, tw, o); tlp += __shfl_xor_sync(0xffffffffu, tlp, o); }
  if (lane == 0) { atomicAdd(&acc_w, tw); atomicAdd(&acc_lp, tlp); }
  __syncthreads();
  for (int t = threadIdx.x; t < K * 5; t += blockDim.x) {
    const float v = acc[t / 5][t % 5];
    if (v != 0.f) atomicAdd(A.stats + (t / 5) * 6 + (t % 5), (double)v);
  }
  if (threadIdx.x == 0) { atomicAdd(A.stats + K * 6, (double)acc_w); atomicAdd(A.stats + K * 6 + 1, (double)acc_lp); }
}

__global__ void k_gmm_update(const __grid_constant__ GmmArgs A) {
  const int K = A.k;
  const double sw = A.stats[K * 6];
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const double* s = A.stats + k * 6;
    const double nk = s[0];
    if (nk > 0.0 && sw > 0.0) {
      const double mx = s[1] / nk, my = s[2] / nk, mz = s[3] / nk;      // shift of the mean
      // sum wr |x - mu_new|^2 = sum wr |x - mu_old|^2 - nk |mu_new - mu_old|^2
      double var = (s[4] / nk - (mx * mx + my * my + mz * mz)) / 3.0;
      if (!(var > (double)A.min_var)) var = (double)A.min_var;
      A.mean_out[k * 3] = (float)(A.mean[k * 3] + mx); A.mean_out[k * 3 + 1] = (float)(A.mean[k * 3 + 1] + my);
      A.mean_out[k * 3 + 2] = (float)(A.mean[k * 3 + 2] + mz);
      A.var_out[k] = (float)var;
      A.prior_out[k] = (float)(nk / sw);
    } else {                                  // an empty component keeps its place with zero prior
      A.mean_out[k * 3] = A.mean[k * 3]; A.mean_out[k * 3 + 1] = A.mean[k * 3 + 1]; A.mean_out[k * 3 + 2] = A.mean[k * 3 + 2];
      A.var_out[k] = A.var[k];
      A.prior_out[k] = 0.f;
    }
  }
  if (threadIdx.x == 0) A.logp_out[0] = (float)A.stats[K * 6 + 1];      // sum_i w_i log p(x_i) under the INPUT parameters
}

cudaError_t launch_compensate(const float* lum, int64_t n, float threshold, int use_threshold, double* scratch, float* w,
                              cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double), stream);
  if (e != cudaSuccess) return e;
  int64_t b = (n + 255) / 256;
  const unsigned grid = (unsigned)(b < 148 * 8 ? b : 148 * 8);
  k_lum_sum<<<grid, 256, 0, stream>>>(lum, n, scratch);
  k_compensate<<<grid, 256, 0, stream>>>(lum, n, scratch, threshold, use_threshold, w);
  return cudaGetLastError();
}

cudaError_t launch_gmm_em(const GmmArgs& A, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(A.stats, 0, sizeof(double) * (A.k * 6 + 2), stream);
  if (e != cudaSuccess) return e;
  if (A.n > 0) {
    int64_t b = (A.n + 255) / 256;
    const unsigned grid = (unsigned)(b < 148 * 4 ? b : 148 * 4);
    k_gmm_em<<<grid, 256, 0, stream>>>(A);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  k_gmm_update<<<1, 128, 0, stream>>>(A);
  return cudaGetLastError();
}

}  // namespace nemit
