// red_bench.cu - throughput of fire-and-forget reductions (red.global.add) on a 64 MiB fp32 table with random
// addresses: scalar f32, v2.f32 (one table entry) and v4.f32 (an aligned pair of entries), per lane-op.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s ^ (s >> 15); }

template <int W>   // 1, 2, 4 floats per op
__global__ void k_red(float* table, uint32_t mask_entries, int ops_per_thread, int spread) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  for (int i = 0; i < ops_per_thread; ++i) {
    uint32_t e = rng(s) & mask_entries;                 // entry index (8-byte entries)
    if (spread == 0) e = (e & ~31u) | (threadIdx.x & 31);   // warp-coalesced variant: 32 consecutive entries
    if (W == 1) atomicAdd(table + 2 * e, 1.f);
    if (W == 2) atomicAdd(reinterpret_cast<float2*>(table) + e, make_float2(1.f, 2.f));
    if (W == 4) atomicAdd(reinterpret_cast<float4*>(table) + (e >> 1), make_float4(1.f, 2.f, 3.f, 4.f));
  }
}

// partially filled warps: only lanes < active issue the reduction (cost per lane-op or per instruction?)
__global__ void k_red_partial(float* table, uint32_t mask_entries, int ops_per_thread, int active) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  const bool on = (threadIdx.x & 31) < active;
  for (int i = 0; i < ops_per_thread; ++i) {
    uint32_t e = rng(s) & mask_entries;
    if (on) atomicAdd(reinterpret_cast<float2*>(table) + e, make_float2(1.f, 2.f));
  }
}

// do random L1-missing gathers and random reductions share one path?  Each thread issues `nl` float2 gathers and `nr`
// float2 reductions per iteration on two different 64 MiB tables.
__global__ void k_mix(const float2* __restrict__ src, float* dst, uint32_t mask_entries, int iters, int nl, int nr, float* sink) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  float acc = 0.f;
  for (int i = 0; i < iters; ++i) {
    for (int j = 0; j < nl; ++j) { float2 v = __ldg(src + (rng(s) & mask_entries)); acc += v.x + v.y; }
    for (int j = 0; j < nr; ++j) atomicAdd(reinterpret_cast<float2*>(dst) + (rng(s) & mask_entries), make_float2(1.f, 2.f));
  }
  if (acc == 123.456f) *sink = acc;
}

int main() {
  const size_t entries = 1u << 23;                      // 8 Mi entries x 8 B = 64 MiB
  float* t; CK(cudaMalloc(&t, entries * 8)); CK(cudaMemset(t, 0, entries * 8));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int spread = 1; spread >= 0; --spread)
    for (int threads : {128, 256, 1024})
      for (int W : {1, 2, 4}) {
        const int blocks = 148 * (2048 / threads), ops = 512;
        auto launch = [&] {
          if (W == 1) k_red<1><<<blocks, threads>>>(t, entries - 1, ops, spread);
          if (W == 2) k_red<2><<<blocks, threads>>>(t, entries - 1, ops, spread);
          if (W == 4) k_red<4><<<blocks, threads>>>(t, entries - 1, ops, spread);
        };
        launch(); CK(cudaDeviceSynchronize());
        cudaEventRecord(a); launch(); cudaEventRecord(b); CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, a, b);
        double lane_ops = (double)blocks * threads * ops;
        printf("%s W=%d threads/CTA=%4d (full occupancy): %.3f ms, %.1f G lane-ops/s, %.2f SM-cycles per lane-op @1.9GHz\n",
               spread ? "random " : "coalesc", W, threads, ms, lane_ops / ms * 1e-6, ms * 1e-3 * 1.9e9 * 148 / lane_ops);
      }
  // is the 1.5 cycles per lane-op an SM-side or an L2-side (chip-wide) limit?  Same kernel on fewer SMs: an SM-side limit
  // keeps the per-SM rate, a chip-wide one raises it
  for (int nsm : {148, 111, 74, 37, 18}) {
    const int blocks = nsm * 8, threads = 256, ops = 512;
    auto launch = [&] { k_red<2><<<blocks, threads>>>(t, entries - 1, ops, 1); };
    launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(a); launch(); cudaEventRecord(b); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, a, b);
    double lane_ops = (double)blocks * threads * ops;
    printf("SM sweep: %3d SMs (8 CTAs of 256 each): %.3f ms, %.1f G lane-ops/s chip-wide, %.2f SM-cycles per lane-op per SM @1.965GHz\n",
           nsm, ms, lane_ops / ms * 1e-6, ms * 1e-3 * 1.965e9 * nsm / lane_ops);
  }
  // low occupancy: 8 warps per SM
  for (int W : {2, 4}) {
    const int blocks = 148, threads = 256, ops = 4096;
    auto launch = [&] { if (W == 2) k_red<2><<<blocks, threads>>>(t, entries - 1, ops, 1); else k_red<4><<<blocks, threads>>>(t, entries - 1, ops, 1); };
    launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(a); launch(); cudaEventRecord(b); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, a, b);
    double lane_ops = (double)blocks * threads * ops;
    printf("random  W=%d 8 warps/SM: %.3f ms, %.1f G lane-ops/s, %.2f SM-cycles per lane-op\n", W, ms, lane_ops / ms * 1e-6, ms * 1e-3 * 1.9e9 * 148 / lane_ops);
  }
  for (int active : {32, 16, 8, 4, 1}) {
    const int blocks = 148 * 8, threads = 256, ops = 512;
    auto launch = [&] { k_red_partial<<<blocks, threads>>>(t, entries - 1, ops, active); };
    launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(a); launch(); cudaEventRecord(b); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double instr = (double)blocks * threads / 32 * ops;
    printf("partial warps: %2d active lanes per red.v2 instruction: %.3f ms, %.1f SM-cycles per warp instruction, %.2f per active lane\n",
           active, ms, ms * 1e-3 * 1.9e9 * 148 / instr, ms * 1e-3 * 1.9e9 * 148 / instr / active);
  }
  {
    float2* src; CK(cudaMalloc(&src, entries * 8)); CK(cudaMemset(src, 0, entries * 8));
    float* sink; CK(cudaMalloc(&sink, 4));
    const int blocks = 148 * 8, threads = 256, iters = 128;
    const int cfg[5][2] = {{4, 0}, {0, 4}, {4, 4}, {4, 2}, {2, 4}};
    for (uint32_t mask : {(uint32_t)entries - 1, (uint32_t)entries / 4 - 1})      // 2 x 64 MiB (exceeds the 126 MB L2) and 2 x 16 MiB
    for (auto& c : cfg) {
      auto launch = [&] { k_mix<<<blocks, threads>>>(src, t, mask, iters, c[0], c[1], sink); };
      launch(); CK(cudaDeviceSynchronize());
      cudaEventRecord(a); launch(); cudaEventRecord(b); CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, a, b);
      const double cyc = ms * 1e-3 * 1.9e9 * 148 / ((double)blocks * threads * iters);
      printf("mixed (%u MiB per table): %d gathers + %d reds per iteration: %.3f ms, %.2f SM-cycles per iteration (sum of the separate costs: %.2f)\n",
             (mask + 1) >> 17, c[0], c[1], ms, cyc, c[0] * 1.0 + c[1] * 1.5);
    }
  }
  return 0;
}
