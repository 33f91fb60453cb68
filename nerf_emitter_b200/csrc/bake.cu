// bake.cu - headless spherical-sensor bake of the NeRF radiance (BASELINE config 5):
//   k_spherical_rays  : SphericalCamera.sample_ray (DS/sensors/spherical_sensor.py:19-44, use_sphere_uv = False,
//                       to_world = translation as in assets/sensors/spherical_sensor.xml) for a range of film samples
//   k_film_accumulate : the film's ImageBlock::put - box filter, or the gaussian reconstruction filter of the sensor
//                       XML (stddev 0.25, radius 4 stddev, value exp(-x^2 / 2 s^2) - exp(-r^2 / 2 s^2), separable),
//                       accumulating (w * rgb, w) per pixel           [restated from the published Mitsuba 3.4.1
//                       algorithm; mitsuba is not in the reference tree: parity unpinned, DESIGN.md section 4]
//   k_film_develop    : hdrfilm develop: rgb / weight, alpha = 1 where the pixel received weight
// Everything between the two is the emitter forward (nemit_forward); the image never leaves the device.
#include "kernels.h"

namespace nemit {

// counter-based RNG for the sub-pixel positions (the independent sampler's role): two uniforms in [0,1) per sample
__device__ __forceinline__ uint32_t pcg_hash(uint32_t x) {
  x = x * 747796405u + 2891336453u;
  uint32_t w = ((x >> ((x >> 28u) + 4u)) ^ x) * 277803737u;
  return (w >> 22u) ^ w;
}

__global__ void k_spherical_rays(const __grid_constant__ SphericalArgs A) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < A.n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = A.first + i;                       // global sample index, pixel-major
    const int64_t pix = s / A.spp;
    const int px = (int)(pix % A.width), py = (int)(pix / A.width);
    float jx, jy;
    if (A.jitter) { jx = A.jitter[i * 2]; jy = A.jitter[i * 2 + 1]; }
    else {
      const uint32_t h0 = pcg_hash((uint32_t)s * 2u + A.seed * 0x9E3779B9u), h1 = pcg_hash((uint32_t)s * 2u + 1u + A.seed * 0x9E3779B9u);
      jx = (float)(h0 >> 8) * (1.f / 16777216.f);
      jy = (float)(h1 >> 8) * (1.f / 16777216.f);
    }
    const float fx = (float)px + jx, fy = (float)py + jy;            // film position in pixel units
    const float u = fx / (float)A.width, v = fy / (float)A.height;   // position_sample
    float sp, cp, st, ct;
    sincospif(2.f * u, &sp, &cp);                                    // phi = 2 pi u
    sincospif(v, &st, &ct);                                          // theta = pi v
    const float d[3] = {sp * st, ct, -cp * st};                      // spherical_sensor.py:29-31
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      A.v[k * A.n + i] = d[k];
      A.o[k * A.n + i] = A.origin[k] + d[k] * A.near_clip;           // ray.o += ray.d * near_t
    }
    A.pos[i * 2] = fx;
    A.pos[i * 2 + 1] = fy;
  }
}

__device__ __forceinline__ float gauss_filter(float x, float alpha, float bias, float radius) {
  return fabsf(x) < radius ? fmaxf(0.f, __expf(alpha * x * x) - bias) : 0.f;
}

__global__ void k_film_accumulate(const __grid_constant__ FilmArgs A) {
  const float radius = 4.f * A.stddev;
  const float alpha = -1.f / (2.f * A.stddev * A.stddev);
  const float bias = __expf(alpha * radius * radius);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < A.n; i += (int64_t)gridDim.x * blockDim.x) {
    const float x = A.pos[i * 2], y = A.pos[i * 2 + 1];
    const float r = A.rgb[i * 3], g = A.rgb[i * 3 + 1], b = A.rgb[i * 3 + 2];
    if (A.filter == 0) {
      const int px = min(max((int)floorf(x), 0), A.width - 1), py = min(max((int)floorf(y), 0), A.height - 1);
      atomicAdd(A.accum + (int64_t)py * A.width + px, make_float4(r, g, b, 1.f));
      continue;
    }
    // pixel centres (p + 0.5) within the filter radius of the sample
    const int x0 = max((int)ceilf(x - radius - 0.5f), 0), x1 = min((int)floorf(x + radius - 0.5f), A.width - 1);
    const int y0 = max((int)ceilf(y - radius - 0.5f), 0), y1 = min((int)floorf(y + radius - 0.5f), A.height - 1);
    for (int py = y0; py <= y1; ++py) {
      const float wy = gauss_filter((float)py + 0.5f - y, alpha, bias, radius);
      if (wy == 0.f) continue;
      for (int px = x0; px <= x1; ++px) {
        const float w = wy * gauss_filter((float)px + 0.5f - x, alpha, bias, radius);
        if (w != 0.f) atomicAdd(A.accum + (int64_t)py * A.width + px, make_float4(w * r, w * g, w * b, w));
      }
    }
  }
}

__global__ void k_film_develop(int64_t n, const float4* __restrict__ accum, float4* __restrict__ rgba) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = accum[i];
    const float inv = a.w > 0.f ? 1.f / a.w : 0.f;
    rgba[i] = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w > 0.f ? 1.f : 0.f);
  }
}

static unsigned grid_for(int64_t n, int threads) {
  int64_t b = (n + threads - 1) / threads;
  return (unsigned)(b < 148 * 16 ? (b < 1 ? 1 : b) : 148 * 16);
}

cudaError_t launch_spherical_rays(const SphericalArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  k_spherical_rays<<<grid_for(A.n, 256), 256, 0, stream>>>(A);
  return cudaGetLastError();
}
cudaError_t launch_film_accumulate(const FilmArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  k_film_accumulate<<<grid_for(A.n, 256), 256, 0, stream>>>(A);
  return cudaGetLastError();
}
cudaError_t launch_film_develop(int64_t n, const float* accum, float* rgba, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  k_film_develop<<<grid_for(n, 256), 256, 0, stream>>>(n, reinterpret_cast<const float4*>(accum), reinterpret_cast<float4*>(rgba));
  return cudaGetLastError();
}

}  // namespace nemit
