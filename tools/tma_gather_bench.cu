// tma_gather_bench.cu - is there a second path for L1-missing table gathers beside the LSU one?  
// The 64 MiB table is viewed as a 2-D tensor of (rows = entries / 2) x (4 floats): one row = one aligned 16-byte pair of
// hash-table entries.  Rows are fetched by the TMA engine with cp.async.bulk.tensor.2d ... tile::gather4 (four arbitrary
// rows per instruction, landing in shared memory, completion on an mbarrier) and, for comparison / concurrently, by
// ordinary 16-byte LSU gathers.  Prints rows/s for: LSU only, TMA only, and both at once (each warp issues both).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_gather_bench tools/tma_gather_bench.cu
//   tools/tma_gather_bench [box_rows=1|4]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s ^ (s >> 15); }
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

constexpr int kWarps = 16;            // per CTA
constexpr int kSlot = 128;            // bytes of shared memory per lane (one gather4 = 4 x 16 B, 128-byte aligned slot)

// MODE bit 0: TMA gather4 (4 rows per lane per iteration); bit 1: LSU gathers (4 x 16-byte loads per lane per iteration)
template <int MODE>
__global__ void __launch_bounds__(kWarps * 32)
k_gather(const __grid_constant__ CUtensorMap tmap, const float4* __restrict__ table, uint32_t row_mask, int iters, float* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned char* buf = smem + (size_t)w * 32 * kSlot;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + (size_t)kWarps * 32 * kSlot) + w;
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  uint32_t phase = 0;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    uint32_t r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = rng(s) & row_mask;
    if (MODE & 1) {
      if (lane == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(32 * 64) : "memory");
      __syncwarp();
      asm volatile(
          "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
          ::"r"(smem_u32(buf + lane * kSlot)), "l"(&tmap), "r"(0), "r"((int)r[0]), "r"((int)r[1]), "r"((int)r[2]), "r"((int)r[3]),
            "r"(smem_u32(bar))
          : "memory");
    }
    if (MODE & 2) {
#pragma unroll
      for (int j = 4; j < 8; ++j) {      // all four components used: one LDG.128 per row (two of them would turn into two LDG.32)
        const float4 v = __ldg(table + r[j]);
        acc += (v.x + v.y) + (v.z + v.w);
      }
    }
    if (MODE & 1) {
      uint32_t spins = 0;
      while (!mbar_try_wait(bar, phase)) { if (++spins > (1u << 22)) __trap(); }
      phase ^= 1u;
      acc += *reinterpret_cast<const float*>(buf + lane * kSlot);
      __syncwarp();
    }
  }
  if (acc == 123.456f) out[0] = acc;
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int MODE>
static void run(const CUtensorMap& tm, const float4* t, uint32_t rows, int ctas_per_sm, float* out, const char* name) {
  const int blocks = 148 * ctas_per_sm, iters = 256;
  const size_t smem = (size_t)kWarps * 32 * kSlot + kWarps * 8;
  CK(cudaFuncSetAttribute(k_gather<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k_gather<MODE><<<blocks, kWarps * 32, smem>>>(tm, t, rows - 1, iters, out); CK(cudaDeviceSynchronize());
  cudaEventRecord(a); k_gather<MODE><<<blocks, kWarps * 32, smem>>>(tm, t, rows - 1, iters, out); cudaEventRecord(b);
  CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double lanes = (double)blocks * kWarps * 32 * iters;
  const double tma_rows = (MODE & 1) ? lanes * 4 : 0.0, lsu_rows = (MODE & 2) ? lanes * 4 : 0.0;
  printf("%-28s %2d warps/SM: %8.3f ms   TMA %7.1f G rows/s   LSU %7.1f G rows/s   total %7.1f G rows/s (%.2f SM-cycles per row at 1.965 GHz)\n",
         name, kWarps * ctas_per_sm, ms, tma_rows / ms * 1e-6, lsu_rows / ms * 1e-6, (tma_rows + lsu_rows) / ms * 1e-6,
         ms * 1e-3 * 1.965e9 * 148 / (tma_rows + lsu_rows));
}

int main(int argc, char** argv) {
  const int box_rows = argc > 1 ? atoi(argv[1]) : 1;
  const size_t bytes = 64u << 20;
  const uint32_t rows = (uint32_t)(bytes / 16);
  float4* t; CK(cudaMalloc(&t, bytes)); CK(cudaMemset(t, 0, bytes));
  float* out; CK(cudaMalloc(&out, 4));
  EncodeTiled encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &q));
  if (!encode || q != cudaDriverEntryPointSuccess) { printf("cuTensorMapEncodeTiled not available\n"); return 1; }
  CUtensorMap tm;
  const cuuint64_t gdim[2] = {4, rows};
  const cuuint64_t gstride[1] = {16};
  const cuuint32_t box[2] = {4, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d (box rows %d)\n", (int)r, box_rows); return 1; }
  printf("table 64 MiB = %u rows of 16 B, tensor map box = 4 x %d\n", rows, box_rows);
  for (int c : {1, 2, 3}) {
    run<2>(tm, t, rows, c, out, "LSU 16-byte gathers only");
    run<1>(tm, t, rows, c, out, "TMA tile::gather4 only");
    run<3>(tm, t, rows, c, out, "TMA gather4 + LSU together");
  }
  return 0;
}
