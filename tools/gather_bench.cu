// gather_bench.cu - ceiling of random table gathers on one B200: lane-loads/s for 4 / 8 / 16-byte loads from an
// L2-resident table with the cache operators the field kernels can use.  Patterns: "random" (every lane-load its own
// line, = one corner of a hashed level), "pair" (two 8-byte loads from one aligned 16-byte pair, = x-neighbour corners),
// "coher" (the 32 lanes of a warp read 32 consecutive entries, = a coarse level with ray-coherent samples).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_bench tools/gather_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s ^ (s >> 15); }

enum { OP_NC = 0, OP_CG = 1, OP_NOALLOC = 2 };
enum { PAT_RANDOM = 0, PAT_PAIR = 1, PAT_COHER = 2 };

template <int OP>
__device__ __forceinline__ float ld4(const uint32_t* p) {
  uint32_t v;
  if (OP == OP_NC) v = __ldg(p);
  if (OP == OP_CG) v = __ldcg(p);
  if (OP == OP_NOALLOC) asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
  return __uint_as_float(v);
}
template <int OP>
__device__ __forceinline__ float ld8(const float2* p) {
  float2 v;
  if (OP == OP_NC) v = __ldg(p);
  if (OP == OP_CG) v = __ldcg(p);
  if (OP == OP_NOALLOC) asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
  return v.x + v.y;
}
template <int OP>
__device__ __forceinline__ float ld16(const float4* p) {
  float4 v;
  if (OP == OP_NC) v = __ldg(p);
  if (OP == OP_CG) v = __ldcg(p);
  if (OP == OP_NOALLOC) asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v.x + v.y + v.z + v.w;
}

template <int OP>
__device__ __forceinline__ float ld32(const float4* p) {   // one whole 32-byte sector per lane (LDG.256, sm_100+)
  float a, b, c, d, e, f, g, h;
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d), "=f"(e), "=f"(f), "=f"(g), "=f"(h) : "l"(p));
  return a + b + c + d + e + f + g + h;
}

// W = bytes per lane-load; table has `mask + 1` entries of W bytes
template <int W, int OP, int PAT>
__global__ void k_gather(const void* table, uint32_t mask, int iters, float* out) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  uint32_t ws = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u)) * 2654435761u + 999u;   // per-warp stream
  float acc = 0.f;
  for (int i = 0; i < iters; ++i) {
    uint32_t e[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (PAT == PAT_RANDOM) e[j] = rng(s) & mask;
      if (PAT == PAT_PAIR) { if ((j & 1) == 0) e[j] = rng(s) & mask; else e[j] = e[j - 1] ^ 1u; }
      if (PAT == PAT_COHER) e[j] = ((rng(ws) & mask) & ~31u) | (threadIdx.x & 31);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (W == 4) acc += ld4<OP>(reinterpret_cast<const uint32_t*>(table) + e[j]);
      if (W == 8) acc += ld8<OP>(reinterpret_cast<const float2*>(table) + e[j]);
      if (W == 16) acc += ld16<OP>(reinterpret_cast<const float4*>(table) + e[j]);
      if (W == 32) acc += ld32<OP>(reinterpret_cast<const float4*>(table) + 2 * e[j]);
    }
  }
  if (acc == 123.456f) out[0] = acc;
}

template <int W, int OP, int PAT>
void run(const void* t, size_t table_bytes, int threads, int ctas_per_sm, float* out, const char* name) {
  const int blocks = 148 * ctas_per_sm, iters = 256;
  uint32_t mask = (uint32_t)(table_bytes / W) - 1;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k_gather<W, OP, PAT><<<blocks, threads>>>(t, mask, iters, out); CK(cudaDeviceSynchronize());
  cudaEventRecord(a); k_gather<W, OP, PAT><<<blocks, threads>>>(t, mask, iters, out); cudaEventRecord(b); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b);
  double loads = (double)blocks * threads * iters * 8;
  printf("%-34s table %4zu MiB warps/SM %2d: %7.3f ms  %7.1f G lane-loads/s  %5.2f SM-cycles/lane-load  %6.0f GB/s useful\n",
         name, table_bytes >> 20, threads * ctas_per_sm / 32, ms, loads / ms * 1e-6, ms * 1e-3 * 1.965e9 * 148 / loads,
         loads * W / ms * 1e-6);
}

int main() {
  const size_t bytes = 64u << 20;
  void* t; CK(cudaMalloc(&t, bytes)); CK(cudaMemset(t, 0, bytes));
  float* out; CK(cudaMalloc(&out, 4));
  for (int occ : {16, 48}) {
    int threads = 256, c = occ * 32 / threads;
    run<8, OP_NC, PAT_RANDOM>(t, bytes, threads, c, out, "8B  nc       random");
    run<8, OP_CG, PAT_RANDOM>(t, bytes, threads, c, out, "8B  cg       random");
    run<8, OP_NOALLOC, PAT_RANDOM>(t, bytes, threads, c, out, "8B  noalloc  random");
    run<8, OP_NC, PAT_PAIR>(t, bytes, threads, c, out, "8B  nc       pair (2 loads/16B)");
    run<8, OP_NOALLOC, PAT_PAIR>(t, bytes, threads, c, out, "8B  noalloc  pair (2 loads/16B)");
    run<16, OP_NC, PAT_RANDOM>(t, bytes, threads, c, out, "16B nc       random");
    run<16, OP_NOALLOC, PAT_RANDOM>(t, bytes, threads, c, out, "16B noalloc  random");
    run<4, OP_NC, PAT_RANDOM>(t, bytes / 2, threads, c, out, "4B  nc       random (half2 table)");
    run<4, OP_NOALLOC, PAT_RANDOM>(t, bytes / 2, threads, c, out, "4B  noalloc  random (half2 table)");
    run<8, OP_NC, PAT_RANDOM>(t, bytes / 2, threads, c, out, "8B  nc       random (2 x half2)");
    run<8, OP_NC, PAT_COHER>(t, bytes, threads, c, out, "8B  nc       coherent");
    run<8, OP_NC, PAT_RANDOM>(t, 64u << 10, threads, c, out, "8B  nc       random, 64 KiB (L1)");
    run<8, OP_NC, PAT_RANDOM>(t, 8u << 20, threads, c, out, "8B  nc       random, 8 MiB");
    run<32, OP_NC, PAT_RANDOM>(t, bytes, threads, c, out, "32B nc       random");
    run<16, OP_NC, PAT_RANDOM>(t, 64u << 10, threads, c, out, "16B nc       random, 64 KiB (L1)");
    run<32, OP_NC, PAT_RANDOM>(t, 64u << 10, threads, c, out, "32B nc       random, 64 KiB (L1)");
    run<16, OP_NC, PAT_COHER>(t, bytes, threads, c, out, "16B nc       coherent");
    run<32, OP_NC, PAT_COHER>(t, bytes, threads, c, out, "32B nc       coherent");
    printf("\n");
  }
  return 0;
}
