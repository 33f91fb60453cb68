// microbench.cu - nemit_microbench: the measured ceilings the gather / scatter kernels are reported against
// (bench.py roofline), re-measured on the handle's device in the same run: random L1-missing / L1-hitting / warp-coherent
// 8-byte lane-loads, random red.global.add.v2.f32 lane-ops, and streaming reads of an L2-resident buffer.
// Same kernels as tools/gather_bench.cu and tools/red_bench.cu (round 1), reduced to the variants the report uses.
#include "handle.h"

namespace {

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s ^ (s >> 15); }

// PAT 0: every lane its own random entry; PAT 1: the 32 lanes of a warp read 32 consecutive entries
template <int PAT>
__global__ void k_mb_gather(const float2* __restrict__ table, uint32_t mask, int iters, float* out) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  uint32_t ws = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u)) * 2654435761u + 999u;
  float acc = 0.f;
  for (int i = 0; i < iters; ++i) {
    uint32_t e[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) e[j] = PAT == 0 ? (rng(s) & mask) : (((rng(ws) & mask) & ~31u) | (threadIdx.x & 31));
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float2 v = __ldg(table + e[j]); acc += v.x + v.y; }
  }
  if (acc == 123.456f) out[0] = acc;
}

__global__ void k_mb_red(float2* table, uint32_t mask, int ops) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  for (int i = 0; i < ops; ++i) atomicAdd(table + (rng(s) & mask), make_float2(1.f, 2.f));
}

__global__ void k_mb_l2_read(const float4* __restrict__ buf, int64_t n4, int reps, float* out) {
  float acc = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      const float4 v = __ldcg(buf + i);
      acc += v.x + v.y + v.z + v.w;
    }
  if (acc == 123.456f) out[0] = acc;
}

}  // namespace

extern "C" int nemit_microbench(NemitHandle* h, int32_t kind, double* result_out, void* stream) {
  if (!h || !result_out) return nemit_fail(NEMIT_E_INVALID, "nemit_microbench: null argument%s");
  if (kind < NEMIT_MB_GATHER_L1MISS || kind > NEMIT_MB_L2_READ) return nemit_fail(NEMIT_E_INVALID, "nemit_microbench: unknown kind%s");
  DeviceGuard guard(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t bytes = 64u << 20;
  void* buf = nullptr;
  float* out = nullptr;
  cudaEvent_t a = nullptr, b = nullptr;
  cudaError_t e = cudaMalloc(&buf, bytes);
  if (e == cudaSuccess) e = cudaMalloc(&out, 16);
  if (e == cudaSuccess) e = cudaMemsetAsync(buf, 0, bytes, st);
  if (e == cudaSuccess) e = cudaEventCreate(&a);
  if (e == cudaSuccess) e = cudaEventCreate(&b);
  double ops = 0.0;
  float ms = 0.f;
  if (e == cudaSuccess) {
    const int threads = 256, blocks = 148 * 6, iters = 256;     // 48 warps per SM
    for (int rep = 0; rep < 2; ++rep) {                        // first pass warms the caches, second one is timed
      if (rep == 1) cudaEventRecord(a, st);
      switch (kind) {
        case NEMIT_MB_GATHER_L1MISS:
          k_mb_gather<0><<<blocks, threads, 0, st>>>(static_cast<const float2*>(buf), (uint32_t)(bytes / 8) - 1, iters, out);
          ops = (double)blocks * threads * iters * 8;
          break;
        case NEMIT_MB_GATHER_L1HIT:
          k_mb_gather<0><<<blocks, threads, 0, st>>>(static_cast<const float2*>(buf), (uint32_t)((64u << 10) / 8) - 1, iters, out);
          ops = (double)blocks * threads * iters * 8;
          break;
        case NEMIT_MB_GATHER_COHERENT:
          k_mb_gather<1><<<blocks, threads, 0, st>>>(static_cast<const float2*>(buf), (uint32_t)(bytes / 8) - 1, iters, out);
          ops = (double)blocks * threads * iters * 8;
          break;
        case NEMIT_MB_RED:
          k_mb_red<<<148 * 8, threads, 0, st>>>(static_cast<float2*>(buf), (uint32_t)(bytes / 8) - 1, 512);
          ops = (double)148 * 8 * threads * 512;
          break;
        default:
          k_mb_l2_read<<<148 * 8, threads, 0, st>>>(static_cast<const float4*>(buf), (int64_t)(32u << 20) / 16, 16, out);
          ops = (double)(32u << 20) * 16;
          break;
      }
      if (rep == 1) cudaEventRecord(b, st);
    }
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaEventSynchronize(b);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, a, b);
  }
  if (a) cudaEventDestroy(a);
  if (b) cudaEventDestroy(b);
  cudaFree(buf);
  cudaFree(out);
  if (e != cudaSuccess) return nemit_cuda_fail(e, "nemit_microbench");
  result_out[0] = ops / (ms * 1e-3);
  result_out[1] = ms;
  h->launches += 2;
  return NEMIT_OK;
}
