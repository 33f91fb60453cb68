// vmf.cu - vMF-mixture emitter importance sampling and pdf (DS/emitters/vMF.py:24-81).
//
// The reference composes four Mitsuba 3.4.1 primitives that are NOT in its tree (C++ in the mitsuba
// wheel): DiscreteDistribution.sample_reuse / eval_pmf_normalized, Frame3f (coordinate_system, Duff et
// al. 2017), warp::square_to_von_mises_fisher (concentric-disk warp) and its pdf.  They are restated
// here from the published Mitsuba 3 sources; parity is pinned only by properties (tests/test_vmf*.py).
//
// Without a VmfFrame everything is in emitter-local coordinates; with one the kernels do what the emitter does: it.p and
// ds.d are world-space, p = W^-1 it.p, d = normalize(W d_local), pdf evaluated at normalize(W^-1 d), and the rest of the
// DirectionSample (dist = 2 max(r_bsphere, |it.p - c|), ds.p = it.p + d dist) is written too (vMF.py:25,37-50,65-67).
// One thread per sample; the K lobes (position, 1/std^2, pmf, cdf) live in shared memory.
#include <cstring>

#include "vmf.h"
#include "common.cuh"

namespace nemit {

constexpr int kMaxLobes = 1024;
constexpr float kInvTwoPi = 0.15915494309189535f;
constexpr float kInvFourPi = 0.07957747154594767f;
constexpr float kPi = 3.14159265358979323846f;

struct LobeSmem {
  float px[kMaxLobes], py[kMaxLobes], pz[kMaxLobes];
  float inv_var[kMaxLobes];
  float pmf[kMaxLobes];     // normalised
  float cdf[kMaxLobes];     // unnormalised running sum (float64 accumulate, like mitsuba's DiscreteDistribution)
  float sum;
  int first, last;          // first / last lobe with non-zero weight (m_valid)
};

__device__ void load_lobes(LobeSmem* L, const float* __restrict__ pos, const float* __restrict__ weight,
                           const float* __restrict__ std, int K) {
  for (int i = threadIdx.x; i < K; i += blockDim.x) {
    L->px[i] = pos[3 * i]; L->py[i] = pos[3 * i + 1]; L->pz[i] = pos[3 * i + 2];
    float s = std[i];
    L->inv_var[i] = 1.0f / (s * s);                 // vMF.py:15,21: inv_var = 1/sqr(std)
  }
  if (threadIdx.x == 0) {
    double acc = 0.0;
    int first = -1, last = -1;
    for (int i = 0; i < K; ++i) {
      float w = weight[i];
      acc += (double)w;
      L->cdf[i] = (float)acc;
      if (w > 0.f) { if (first < 0) first = i; last = i; }
    }
    L->sum = (float)acc;
    L->first = first < 0 ? 0 : first;
    L->last = last < 0 ? 0 : last;
  }
  __syncthreads();
  float norm = 1.0f / L->sum;
  for (int i = threadIdx.x; i < K; i += blockDim.x) L->pmf[i] = weight[i] * norm;
  __syncthreads();
}

// warp::square_to_von_mises_fisher_pdf: exp(k (z-1)) k / (2 pi (1 - exp(-2k))), 1/(4 pi) at k = 0
__device__ __forceinline__ float vmf_pdf(float cos_theta, float kappa) {
  float res = expf(kappa * (cos_theta - 1.f)) * (kappa * kInvTwoPi) / (1.f - expf(-2.f * kappa));
  return kappa > 0.f ? res : kInvFourPi;
}

__device__ __forceinline__ float mixture_pdf(const LobeSmem* L, int K, float px, float py, float pz, float dx,
                                             float dy, float dz) {
  float sum = 0.f;
  for (int i = 0; i < K; ++i) {
    float lx = L->px[i] - px, ly = L->py[i] - py, lz = L->pz[i] - pz;
    float len = sqrtf(lx * lx + ly * ly + lz * lz);
    float inv = 1.f / len;                        // vMF.py:75: |lobe| = 0 is not handled by the reference either
    float cos_theta = (dx * lx + dy * ly + dz * lz) * inv;    // z of Frame(mu).to_local(d)
    sum += vmf_pdf(cos_theta, len * L->inv_var[i]) * L->pmf[i];
  }
  return sum;
}

// Transform4f::transform_affine of a point / of a vector (3 x 4 row-major)
__device__ __forceinline__ void xf_point(const float* m, float& x, float& y, float& z) {
  const float a = m[0] * x + m[1] * y + m[2] * z + m[3], b = m[4] * x + m[5] * y + m[6] * z + m[7],
              c = m[8] * x + m[9] * y + m[10] * z + m[11];
  x = a; y = b; z = c;
}
__device__ __forceinline__ void xf_vector(const float* m, float& x, float& y, float& z) {
  const float a = m[0] * x + m[1] * y + m[2] * z, b = m[4] * x + m[5] * y + m[6] * z, c = m[8] * x + m[9] * y + m[10] * z;
  x = a; y = b; z = c;
}

__global__ void __launch_bounds__(256)
k_vmf_pdf(const float* __restrict__ pos, const float* __restrict__ weight, const float* __restrict__ std, int K,
          const float* __restrict__ p, const float* __restrict__ d, int64_t n, float* __restrict__ pdf,
          const __grid_constant__ VmfFrame F) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  LobeSmem* L = reinterpret_cast<LobeSmem*>(smem_raw);
  load_lobes(L, pos, weight, std, K);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float dx = d[3 * i], dy = d[3 * i + 1], dz = d[3 * i + 2];
    float px = p[3 * i], py = p[3 * i + 1], pz = p[3 * i + 2];
    if (F.on) { xf_vector(F.w2l, dx, dy, dz); xf_point(F.w2l, px, py, pz); }      // vMF.py:65-67
    float inv = rsqrtf(dx * dx + dy * dy + dz * dz);          // vMF.py:66 normalize(d)
    pdf[i] = mixture_pdf(L, K, px, py, pz, dx * inv, dy * inv, dz * inv);
  }
}

__global__ void __launch_bounds__(256)
k_vmf_sample(const float* __restrict__ pos, const float* __restrict__ weight, const float* __restrict__ std, int K,
             const float* __restrict__ p, const float* __restrict__ u, int64_t n, float* __restrict__ d_out,
             float* __restrict__ pdf, const __grid_constant__ VmfFrame F, float* __restrict__ dist_out,
             float* __restrict__ ds_p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  LobeSmem* L = reinterpret_cast<LobeSmem*>(smem_raw);
  load_lobes(L, pos, weight, std, K);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float px = p[3 * i], py = p[3 * i + 1], pz = p[3 * i + 2];
    const float wpx = px, wpy = py, wpz = pz;                 // it.p (world)
    if (F.on) xf_point(F.w2l, px, py, pz);                    // vMF.py:25
    float ux = u[2 * i], uy = u[2 * i + 1];
    // DiscreteDistribution::sample_reuse: first index in [first,last] with cdf[index] >= u*sum
    float value = ux * L->sum;
    int lo = L->first, hi = L->last;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (L->cdf[mid] < value) lo = mid + 1; else hi = mid;
    }
    const int idx = lo;
    float norm = 1.f / L->sum;
    float cdf_prev = idx > 0 ? L->cdf[idx - 1] * norm : 0.f;
    float ure = (ux - cdf_prev) / L->pmf[idx];
    // lobe seen from the shading point (vMF.py:28-34)
    float lx = L->px[idx] - px, ly = L->py[idx] - py, lz = L->pz[idx] - pz;
    float len = sqrtf(lx * lx + ly * ly + lz * lz);
    float mx = lx / len, my = ly / len, mz = lz / len;
    float kappa = len * L->inv_var[idx];
    // warp::square_to_uniform_disk_concentric
    float x = 2.f * ure - 1.f, y = 2.f * uy - 1.f;
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (q13) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float sphi, cphi;
    sincosf(phi, &sphi, &cphi);
    float dxk = r * cphi, dyk = r * sphi;
    // warp::square_to_von_mises_fisher (low-distortion concentric variant)
    float r2 = dxk * dxk + dyk * dyk;
    float sy = fmaxf(1.f - r2, 1e-6f);
    float cos_theta = 1.f + logf(sy + (1.f - sy) * expf(-2.f * kappa)) / kappa;
    float scale = sqrtf(fmaxf(1.f - cos_theta * cos_theta, 0.f) / r2);
    float vx = dxk * scale, vy = dyk * scale, vz = cos_theta;
    if (!(r2 > 0.f)) { vx = 0.f; vy = 0.f; }
    if (!(kappa > 0.f)) {                           // uniform sphere fallback
      float z = 1.f - 2.f * uy, rr = sqrtf(fmaxf(1.f - z * z, 0.f));
      float s2, c2;
      sincosf(2.f * kPi * ure, &s2, &c2);
      vx = rr * c2; vy = rr * s2; vz = z;
    }
    // Frame3f(mu): coordinate_system (Duff et al.), then to_world
    float sign = copysignf(1.f, mz), a = -1.f / (sign + mz), b = mx * my * a;
    float sx = 1.f + sign * mx * mx * a, sy_ = sign * b, sz = -sign * mx;
    float tx = b, ty = sign + my * my * a, tz = -my;
    float wx = sx * vx + tx * vy + mx * vz;
    float wy = sy_ * vx + ty * vy + my * vz;
    float wz = sz * vx + tz * vy + mz * vz;
    float inv = rsqrtf(wx * wx + wy * wy + wz * wz);          // vMF.py:38 normalize
    wx *= inv; wy *= inv; wz *= inv;
    float lx2 = wx, ly2 = wy, lz2 = wz;                       // the direction pdf_direction will see in local space
    if (F.on) {
      xf_vector(F.l2w, wx, wy, wz);                           // vMF.py:37-38: to world, normalize
      inv = rsqrtf(wx * wx + wy * wy + wz * wz);
      wx *= inv; wy *= inv; wz *= inv;
      lx2 = wx; ly2 = wy; lz2 = wz;
      xf_vector(F.w2l, lx2, ly2, lz2);                        // vMF.py:65-66: back, normalize
      inv = rsqrtf(lx2 * lx2 + ly2 * ly2 + lz2 * lz2);
      lx2 *= inv; ly2 *= inv; lz2 *= inv;
    }
    d_out[3 * i] = wx; d_out[3 * i + 1] = wy; d_out[3 * i + 2] = wz;
    pdf[i] = mixture_pdf(L, K, px, py, pz, lx2, ly2, lz2);    // vMF.py:54 ds.pdf = pdf_direction(...)
    if (dist_out || ds_p) {
      // vMF.py:40-46: the reference point may lie outside the bounding sphere (a sensor): dist = 2 max(r, |it.p - c|)
      const float ax = wpx - F.centre[0], ay = wpy - F.centre[1], az = wpz - F.centre[2];
      const float dist = 2.f * fmaxf(F.radius, sqrtf(ax * ax + ay * ay + az * az));
      if (dist_out) dist_out[i] = dist;
      if (ds_p) { ds_p[3 * i] = wpx + wx * dist; ds_p[3 * i + 1] = wpy + wy * dist; ds_p[3 * i + 2] = wpz + wz * dist; }
    }
  }
}

static cudaError_t prep(const void* fn, int K, size_t* smem) {
  if (K < 1 || K > kMaxLobes) return cudaErrorInvalidValue;
  *smem = sizeof(LobeSmem);
  return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem);
}

cudaError_t launch_vmf_sample(const float* pos, const float* weight, const float* std, int K, const float* p,
                              const float* u, int64_t n, float* d, float* pdf, cudaStream_t stream, const VmfFrame* frame,
                              float* dist, float* ds_p) {
  if (n <= 0) return cudaSuccess;
  VmfFrame F;
  memset(&F, 0, sizeof(F));
  if (frame) F = *frame;
  size_t smem;
  cudaError_t e = prep((const void*)k_vmf_sample, K, &smem);
  if (e != cudaSuccess) return e;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_vmf_sample<<<(unsigned)blocks, 256, smem, stream>>>(pos, weight, std, K, p, u, n, d, pdf, F, dist, ds_p);
  return cudaGetLastError();
}

cudaError_t launch_vmf_pdf(const float* pos, const float* weight, const float* std, int K, const float* p,
                           const float* d, int64_t n, float* pdf, cudaStream_t stream, const VmfFrame* frame) {
  if (n <= 0) return cudaSuccess;
  VmfFrame F;
  memset(&F, 0, sizeof(F));
  if (frame) F = *frame;
  size_t smem;
  cudaError_t e = prep((const void*)k_vmf_pdf, K, &smem);
  if (e != cudaSuccess) return e;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_vmf_pdf<<<(unsigned)blocks, 256, smem, stream>>>(pos, weight, std, K, p, d, n, pdf, F);
  return cudaGetLastError();
}

}  // namespace nemit
