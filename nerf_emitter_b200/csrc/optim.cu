// optim.cu - the optimizer step on the borrowed parameter / gradient tensors, i.e. what
// Optimizers.optimizer_scaler_step_all does per parameter group right after the gradient all-reduce
// (NS/engine/optimizers.py:122-137, NS/engine/trainer.py:517-537, groups and hyper-parameters
// NS/configs/method_configs.py:331-342):
//     GradScaler.unscale_ (grad / scale, inf / nan check)  ->  clip_grad_norm_(group, max_norm)  ->
//     clip_grad_value_(group, max_value)  ->  torch.optim.Adam.step (skipped when a non-finite gradient was found)
// as TWO multi-tensor launches over all 22 tensors (19.4 M fp32):
//   k_grad_stats : per-group sum of squares of the unscaled gradients + non-finite flag         (reads 4 B / param)
//   k_adam_step  : clip, Adam moment update, parameter update, optional gradient zeroing        (28..32 B / param)
//   k_adam_finish: advances the per-group applied-step counters (one warp)
// Per group, as the reference keeps one optimizer and one GradScaler found_inf record per group: a disabled group (no
// gradients on this call) and a group with a non-finite gradient are left untouched, and only applied steps count.
// The first two are HBM streams: float4 loads / stores, grid = 148 x 8 CTAs striding over 4096-element chunks.
#include "kernels.h"

namespace nemit {

constexpr int kOptChunk = 4096;     // elements per CTA iteration (256 threads x 4 float4)

__device__ __forceinline__ bool locate_chunk(const OptimArgs& A, int64_t chunk, int& t, int64_t& first) {
  // at most 22 tensors: a linear scan over the chunk prefix is cheaper than anything clever
#pragma unroll 1
  for (int i = 0; i < A.num_tensors; ++i) {
    if (chunk < A.chunk_end[i]) {
      t = i;
      first = (chunk - (i ? A.chunk_end[i - 1] : 0)) * kOptChunk;
      return true;
    }
  }
  return false;
}

__global__ void __launch_bounds__(256)
k_grad_stats(const __grid_constant__ OptimArgs A) {
  double ss[2] = {0.0, 0.0};     // fp64 partials: the clip coefficient then agrees with torch's per-tensor norms to ~1e-7
  bool bad2[2] = {false, false};  // GradScaler keeps found_inf per optimizer, i.e. per group
  const int64_t total = A.chunk_end[A.num_tensors - 1];
  for (int64_t c = blockIdx.x; c < total; c += gridDim.x) {
    int t; int64_t first;
    if (!locate_chunk(A, c, t, first)) break;
    const float* __restrict__ g = A.grad[t];
    const int64_t n = A.numel[t];
    const int grp = A.group[t];
    if (!A.grp[grp].enabled) continue;
    float acc = 0.f;
    bool bad = false;
    if ((n & 3) == 0) {
      for (int64_t i = first + threadIdx.x * 4; i < min(first + kOptChunk, n); i += 256 * 4) {
        const float4 v = *reinterpret_cast<const float4*>(g + i);
        const float a = v.x * A.inv_scale, b = v.y * A.inv_scale, cc = v.z * A.inv_scale, d = v.w * A.inv_scale;
        bad |= !(isfinite(a) && isfinite(b) && isfinite(cc) && isfinite(d));
        acc += a * a + b * b + cc * cc + d * d;
      }
    } else {
      for (int64_t i = first + threadIdx.x; i < min(first + kOptChunk, n); i += 256) {
        const float a = g[i] * A.inv_scale;
        bad |= !isfinite(a);
        acc += a * a;
      }
    }
    ss[grp] += acc;
    bad2[grp] |= bad;
  }
  __shared__ double red[2][8];
  __shared__ int red_bad[2];
  if (threadIdx.x < 2) red_bad[threadIdx.x] = 0;
  __syncthreads();
#pragma unroll
  for (int g2 = 0; g2 < 2; ++g2) {
    double v = ss[g2];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[g2][threadIdx.x >> 5] = v;
  }
  if (bad2[0]) red_bad[0] = 1;
  if (bad2[1]) red_bad[1] = 1;
  __syncthreads();
  if (threadIdx.x < 2) {
    double v = 0.0;
    for (int w = 0; w < 8; ++w) v += red[threadIdx.x][w];
    if (v != 0.0) atomicAdd(A.stats + threadIdx.x, v);
  }
  if (threadIdx.x < 2 && red_bad[threadIdx.x]) A.stats[2 + threadIdx.x] = 1.0;
}

// one Adam element (torch/optim/adam.py _single_tensor_adam, amsgrad = False, weight_decay = 0, maximize = False):
//   exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
//   denom = (exp_avg_sq.sqrt() / sqrt(bias_correction2)).add_(eps); param.addcdiv_(exp_avg, denom, value = -lr / bias_correction1)
__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const OptimGroup& G, float coef,
                                          float inv_scale, float step_size, float bc2_sqrt) {
  g *= inv_scale;
  g *= coef;                                                   // clip_grad_norm_: grad.mul_(clip_coef_clamped)
  if (G.max_value > 0.f) g = fminf(fmaxf(g, -G.max_value), G.max_value);   // clip_grad_value_
  m = m + (g - m) * G.omb1;                           // lerp_ (weight < 0.5 form)
  v = v * G.beta2 + G.omb2 * g * g;
  const float denom = sqrtf(v) / bc2_sqrt + G.eps;
  p = p - step_size * (m / denom);
}

__global__ void __launch_bounds__(256)
k_adam_step(const __grid_constant__ OptimArgs A) {
  if (A.stats_out && blockIdx.x == 0 && threadIdx.x < 4) A.stats_out[threadIdx.x] = (float)A.stats[threadIdx.x];
  // per group: applied (enabled and every unscaled gradient finite) or left untouched.  GradScaler.step skips
  // optimizer.step() of the optimizer whose unscale_ found a non-finite gradient (torch/amp/grad_scaler.py); an optimizer
  // without gradients is not stepped at all (NS/engine/optimizers.py:136)
  bool apply[2];
  float coef[2], step_size[2], bc2_sqrt[2];
#pragma unroll
  for (int g2 = 0; g2 < 2; ++g2) {
    const OptimGroup& G = A.grp[g2];
    apply[g2] = G.enabled && A.stats[2 + g2] == 0.0;
    // clip_grad_norm_ (torch/nn/utils/clip_grad.py): clip_coef = max_norm / (total_norm + 1e-6), clamped to 1
    coef[g2] = 1.f;
    if (G.max_norm > 0.f) coef[g2] = fminf(G.max_norm / ((float)sqrt(A.stats[g2]) + 1e-6f), 1.f);
    // torch Adam: step advances only when optimizer.step() runs; bias corrections from python floats (double)
    const double step = (double)(A.step_state[g2] + 1);
    step_size[g2] = (float)(G.lr / (1.0 - pow(G.beta1_d, step)));
    bc2_sqrt[g2] = (float)sqrt(1.0 - pow(G.beta2_d, step));
  }
  if (!apply[0] && !apply[1] && !A.zero_grads) return;
  const int64_t total = A.chunk_end[A.num_tensors - 1];
  for (int64_t c = blockIdx.x; c < total; c += gridDim.x) {
    int t; int64_t first;
    if (!locate_chunk(A, c, t, first)) break;
    const int grp = A.group[t];
    const bool act = apply[grp];
    if (!act && !A.zero_grads) continue;
    float* __restrict__ p = A.param[t];
    float* __restrict__ g = A.grad[t];
    float* __restrict__ m = A.exp_avg[t];
    float* __restrict__ v = A.exp_avg_sq[t];
    const int64_t n = A.numel[t];
    const OptimGroup& G = A.grp[grp];
    const float cf = coef[grp], ss = step_size[grp], b2 = bc2_sqrt[grp];
    if ((n & 3) == 0) {
      for (int64_t i = first + threadIdx.x * 4; i < min(first + kOptChunk, n); i += 256 * 4) {
        if (act) {
          float4 pp = *reinterpret_cast<float4*>(p + i), mm = *reinterpret_cast<float4*>(m + i), vv = *reinterpret_cast<float4*>(v + i);
          const float4 gg = *reinterpret_cast<const float4*>(g + i);
          adam_elem(pp.x, gg.x, mm.x, vv.x, G, cf, A.inv_scale, ss, b2);
          adam_elem(pp.y, gg.y, mm.y, vv.y, G, cf, A.inv_scale, ss, b2);
          adam_elem(pp.z, gg.z, mm.z, vv.z, G, cf, A.inv_scale, ss, b2);
          adam_elem(pp.w, gg.w, mm.w, vv.w, G, cf, A.inv_scale, ss, b2);
          *reinterpret_cast<float4*>(p + i) = pp;
          *reinterpret_cast<float4*>(m + i) = mm;
          *reinterpret_cast<float4*>(v + i) = vv;
        }
        if (A.zero_grads) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    } else {
      for (int64_t i = first + threadIdx.x; i < min(first + kOptChunk, n); i += 256) {
        if (act) {
          float pp = p[i], mm = m[i], vv = v[i];
          adam_elem(pp, g[i], mm, vv, G, cf, A.inv_scale, ss, b2);
          p[i] = pp; m[i] = mm; v[i] = vv;
        }
        if (A.zero_grads) g[i] = 0.f;
      }
    }
  }
}

// the applied-step counters move after every CTA of k_adam_step has read them (stream order)
__global__ void k_adam_finish(const __grid_constant__ OptimArgs A) {
  if (threadIdx.x < 2 && A.grp[threadIdx.x].enabled && A.stats[2 + threadIdx.x] == 0.0) A.step_state[threadIdx.x] += 1;
}

cudaError_t launch_adam_step(OptimArgs& A, cudaStream_t stream) {
  int64_t chunks = 0;
  for (int i = 0; i < A.num_tensors; ++i) {
    chunks += (A.numel[i] + kOptChunk - 1) / kOptChunk;
    A.chunk_end[i] = chunks;
  }
  if (chunks == 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(A.stats, 0, sizeof(double) * 4, stream);
  if (e != cudaSuccess) return e;
  const unsigned grid = (unsigned)(chunks < 148 * 8 ? chunks : 148 * 8);
  k_grad_stats<<<grid, 256, 0, stream>>>(A);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  k_adam_step<<<grid, 256, 0, stream>>>(A);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  k_adam_finish<<<1, 32, 0, stream>>>(A);
  return cudaGetLastError();
}

}  // namespace nemit
