// tc_common.cuh - sm_100a tensor-core plumbing for the fused MLP kernels: tcgen05.mma (UMMA) issue, TMEM
// allocation / loads, mbarrier waits and the shared-memory matrix descriptors, written as inline PTX.
//
// Operand layout used everywhere here: K-major, no swizzle ("interleave") canonical UMMA layout of fp16
// elements.  A core matrix is 8 rows x 16 bytes (8 halves along K), stored as 8 consecutive 16-byte rows.
// Element (row r, k) of an R x K operand lives at byte
//      (k / 8) * LBO + (r / 8) * SBO + (r % 8) * 16 + (k % 8) * 2
// with SBO = 128 (row groups contiguous) and LBO = R * 16 (one K-chunk of all rows after the other).
// One tcgen05.mma of kind::f16 consumes K = 16, i.e. two K-chunks; k-step j starts at byte 2 * j * LBO.
// The accumulator D (M = 128 rows, N fp32 columns) sits in TMEM: row m = lane m, column n = column n.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nemit {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// bounded spin: a lost tensor-core completion traps (launch error) instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t it = 0; it < (1u << 24); ++it)
    if (mbar_try_wait(bar, parity)) return;
  __trap();
}

// transaction-count arrival + 1D bulk copy global -> shared (TMA engine; 16-byte aligned, size % 16 == 0)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// the source is read once: L2 evict-first policy, so the stream does not push the gradient table out of L2
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                         uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// 1D bulk copy shared -> global (TMA engine, no LSU traffic), as one bulk group; wait_read returns once the engine has
// finished READING shared memory, i.e. the source buffer may be overwritten
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes, uint64_t policy) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes), "l"(policy)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMEM ----------------------------------------------------------------------------------------
// one full warp allocates NCOLS (power of two >= 32) columns; the base address lands in *slot (shared memory)
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "n"(NCOLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
// several allocations by one CTA: allocate each, then give the permit back once
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc_only(uint32_t* slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "n"(NCOLS)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}

// 32 lanes x 32-bit, 16 consecutive columns: thread t of the warp gets lane (warp % 4) * 32 + t
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// same shape, 8 consecutive columns
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- descriptors ---------------------------------------------------------------------------------
// shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor bit layout):
// [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version = 1, [61,64) layout type = 0
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor, kind::f16, fp16 A/B (K-major), fp32 accumulate (cute::UMMA::InstrDescriptor):
// [4,6) c_format = 1 (F32), [7,10) a_format = 0 (F16), [10,13) b_format = 0, bit 15/16 majors = 0 (K),
// [17,23) N >> 3, [24,29) M >> 4
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// general form: operand majors selectable (1 = MN-major: the operand's M/N index is the contiguous one)
__host__ __device__ constexpr uint32_t idesc_f16_mj(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]^T, issued by ONE thread
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// A operand from TENSOR MEMORY (the .ts form): row m of A = TMEM lane m, 32-bit column j of the operand = its K elements
// (2j, 2j + 1) as a packed 16-bit pair, so one k-step of 16 reads 8 columns; `a_tmem` = column base, lane 0.
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32-bit, 4 consecutive columns: thread t of the warp writes lane (warp % 4) * 32 + t
__device__ __forceinline__ void tmem_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// One GEMM of the MLP chain: D[128 x N] = A[128 x K] * B[N x K]^T, fp16 operands in the layout above
// (A: LBO = 128*16, B: LBO = N*16), issued by the calling thread; completion arrives on `bar`.
template <int N, int K>
__device__ __forceinline__ void issue_layer(uint32_t d_tmem, const void* a_smem, const void* b_smem, uint64_t* bar) {
  constexpr uint32_t kALbo = 128 * 16, kBLbo = N * 16;
  const uint32_t a0 = smem_u32(a_smem), b0 = smem_u32(b_smem);
#pragma unroll
  for (int j = 0; j < K / 16; ++j) {
    mma_f16_ss(d_tmem, smem_desc(a0 + 2 * j * kALbo, kALbo, 128), smem_desc(b0 + 2 * j * kBLbo, kBLbo, 128),
               idesc_f16(128, N), j > 0 ? 1u : 0u);
  }
  mma_commit(bar);
}

// A view of one fp16 buffer in the layout above as an MMA operand.  The SAME buffer of R rows x C columns,
// element (r, c) at (c / 8) * R * 16 + (r / 8) * 128 + (r % 8) * 16 + (c % 8) * 2, can be read
//   * K-major  (rows = the operand's M/N index, columns = K):  LBO = R * 16, SBO = 128, 16 columns per k-step
//   * MN-major (columns = the operand's M/N index, rows = K):  LBO = 128 (next 8 rows), SBO = R * 16 (next 8
//     columns), 16 rows per k-step (verified on B200 with tools/umma_probe.cu)
struct Operand {
  uint32_t addr, lbo, sbo, step;
};
__device__ __forceinline__ Operand op_kmajor(const void* buf, int rows) {
  return Operand{smem_u32(buf), (uint32_t)rows * 16u, 128u, 2u * (uint32_t)rows * 16u};
}
__device__ __forceinline__ Operand op_mnmajor(const void* buf, int rows, int first_col = 0) {
  return Operand{smem_u32(buf) + (uint32_t)(first_col >> 3) * (uint32_t)rows * 16u, 128u, (uint32_t)rows * 16u, 256u};
}
// D (+)= A * B^T over `ksteps` k-steps of 16; the first k-step overwrites D unless `accumulate`.  ROLL keeps the k-step
// loop rolled (instruction footprint) - for chains whose issue rate is not on the critical path
template <bool ROLL = false>
__device__ __forceinline__ void mma_chain(uint32_t d_tmem, const Operand& a, const Operand& b, uint32_t idesc, int ksteps,
                                          bool accumulate) {
  if (ROLL) {
#pragma unroll 1
    for (int j = 0; j < ksteps; ++j)
      mma_f16_ss(d_tmem, smem_desc(a.addr + j * a.step, a.lbo, a.sbo), smem_desc(b.addr + j * b.step, b.lbo, b.sbo), idesc,
                 (accumulate || j > 0) ? 1u : 0u);
  } else {
    for (int j = 0; j < ksteps; ++j)
      mma_f16_ss(d_tmem, smem_desc(a.addr + j * a.step, a.lbo, a.sbo), smem_desc(b.addr + j * b.step, b.lbo, b.sbo), idesc,
                 (accumulate || j > 0) ? 1u : 0u);
  }
}

// byte offset of the 16-byte chunk (row r, K-chunk kc) of an operand with R rows
template <int R>
__device__ __forceinline__ uint32_t chunk_off(int r, int kc) { return (uint32_t)(kc * R * 16 + r * 16); }

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// max(x, 0) and the conversion in one instruction (F2FP.RELU.F16.F32.PACK_AB)
__device__ __forceinline__ uint32_t pack_h2_relu(float a, float b) {
  uint32_t r;
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// saturating variant for gradient operands: +-inf would poison a whole accumulator column.  One instruction
// (F2FP.SATFINITE.F16.F32.PACK_AB): out-of-range values go to +-65504, in-range values round to nearest as in pack_h2
__device__ __forceinline__ uint32_t pack_h2_sat(float a, float b) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));      // first source operand = upper half
  return r;
}

// ---- split precision (NP = 2, 3) ---------------------------------------------------------------------
// "fp32 on the tensor cores": every operand element is carried as NP bf16 pieces x0 + x1 (+ x2) with x0 = rn_bf16(x),
// x1 = rn_bf16(x - x0), x2 = rn_bf16(x - x0 - x1): 16 (NP = 2) or 24 (NP = 3) significant bits and fp32's exponent range,
// so no operand scaling is needed.  Each piece has its own buffer in the SAME layout as the fp16 operand, and a product
// A * B is the sum of the piece products A_i B_j with i + j < NP (3 or 6 MMAs into one fp32 accumulator, smallest terms
// first; the dropped terms are below 2^-8NP of the product).  NP = 1 is the fp16 path.
template <int NP>
struct OpFmt { static constexpr uint32_t bits = (1u << 7) | (1u << 10); };         // a_format = b_format = BF16
template <>
struct OpFmt<1> { static constexpr uint32_t bits = 0u; };                          // a_format = b_format = F16

__device__ __forceinline__ uint32_t bf16x2_rn(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// one pair of operand elements -> w[0] (the fp16 pair, or the leading bf16 pair), w[1], w[2] (the lower bf16 pieces);
// SAT clamps to the finite range first (gradient operands: +-inf would poison a whole accumulator column)
// RELU: max(x, 0) folded into the conversion (F2FP.RELU), for the hidden-layer epilogues
template <int NP, bool SAT = false, bool RELU = false>
__device__ __forceinline__ void pack_pair(float a, float b, uint32_t (&w)[3]) {
  if (NP == 1) {
    w[0] = SAT ? pack_h2_sat(a, b) : (RELU ? pack_h2_relu(a, b) : pack_h2(a, b));
  } else {
    if (RELU) { a = fmaxf(a, 0.f); b = fmaxf(b, 0.f); }
    if (SAT) { a = fminf(fmaxf(a, -3.0e38f), 3.0e38f); b = fminf(fmaxf(b, -3.0e38f), 3.0e38f); }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      w[p] = bf16x2_rn(a, b);
      if (p + 1 < NP) { a -= __uint_as_float(w[p] << 16); b -= __uint_as_float(w[p] & 0xffff0000u); }
    }
  }
}
// eight consecutive K elements of row `row` (= the 16-byte chunk kc) of an R-row operand; piece p at buf + p * piece_bytes
template <int NP, int R, bool SAT = false, bool RELU = false>
__device__ __forceinline__ void store_chunk(uint8_t* buf, uint32_t piece_bytes, int row, int kc, const float (&x)[8]) {
  uint32_t w[4][3];
#pragma unroll
  for (int q = 0; q < 4; ++q) pack_pair<NP, SAT, RELU>(x[2 * q], x[2 * q + 1], w[q]);
#pragma unroll
  for (int p = 0; p < NP; ++p)
    *reinterpret_cast<uint4*>(buf + p * piece_bytes + chunk_off<R>(row, kc)) = make_uint4(w[0][p], w[1][p], w[2][p], w[3][p]);
}
// D (+)= A * B^T with NP-piece operands: piece p of an operand sits p * (a_piece | b_piece) bytes after piece 0; a piece
// stride of 0 says the operand has one piece only (exactly representable values: the ones operand)
template <int NP, bool ROLL = false>
__device__ __forceinline__ void mma_chain_np(uint32_t d_tmem, const Operand& a, uint32_t a_piece, const Operand& b,
                                             uint32_t b_piece, uint32_t idesc, int ksteps, bool accumulate) {
  const uint32_t id = idesc | OpFmt<NP>::bits;
  bool acc = accumulate;
#pragma unroll
  for (int order = NP - 1; order >= 0; --order) {          // order = i + j: the small products first
#pragma unroll
    for (int i = 0; i <= order; ++i) {
      const int j = order - i;
      if ((i > 0 && a_piece == 0) || (j > 0 && b_piece == 0)) continue;
      Operand ai = a, bj = b;
      ai.addr += i * a_piece;
      bj.addr += j * b_piece;
      mma_chain<ROLL>(d_tmem, ai, bj, id, ksteps, acc);
      acc = true;
    }
  }
}

// the same with the A operand in tensor memory: piece i of A at column a_tmem + i * a_piece_cols, 8 columns per k-step
template <int NP>
__device__ __forceinline__ void mma_chain_np_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t a_piece_cols, const Operand& b,
                                                uint32_t b_piece, uint32_t idesc, int ksteps, bool accumulate) {
  const uint32_t id = idesc | OpFmt<NP>::bits;
  bool acc = accumulate;
#pragma unroll
  for (int order = NP - 1; order >= 0; --order) {
#pragma unroll
    for (int i = 0; i <= order; ++i) {
      const int j = order - i;
      for (int k = 0; k < ksteps; ++k) {
        mma_f16_ts(d_tmem, a_tmem + i * a_piece_cols + 8 * k, smem_desc(b.addr + j * b_piece + k * b.step, b.lbo, b.sbo), id,
                   (acc || k > 0) ? 1u : 0u);
      }
      acc = true;
    }
  }
}

// W (rows x cols_src, torch (out,in) layout, fp32) -> K-major operand buffer(s) with R rows and K columns (fp16, or
// NP bf16 pieces R * K * 2 bytes apart); entries outside [rows, cols_used) are zero.
template <int R, int K, int NP>
__device__ __forceinline__ void stage_weight_np(uint8_t* dst, const float* __restrict__ W, int rows, int cols_src,
                                                int cols_used, int tid, int nthreads) {
  for (int i = tid; i < R * K; i += nthreads) {
    int n = i / K, k = i % K;
    float v = (n < rows && k < cols_used) ? W[n * cols_src + k] : 0.f;
    uint8_t* at = dst + chunk_off<R>(n, k >> 3) + (k & 7) * 2;
    if (NP == 1) {
      *reinterpret_cast<__half*>(at) = __float2half_rn(v);
    } else {
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        const __nv_bfloat16 piece = __float2bfloat16_rn(v);
        *reinterpret_cast<__nv_bfloat16*>(at + p * (R * K * 2)) = piece;
        v -= __bfloat162float(piece);
      }
    }
  }
}

// W (rows x cols_src, torch (out,in) layout, fp32) -> fp16 K-major operand buffer with R rows and K columns;
// entries outside [rows, cols_used) are zero.
template <int R, int K>
__device__ __forceinline__ void stage_weight(uint8_t* dst, const float* __restrict__ W, int rows, int cols_src,
                                             int cols_used, int tid, int nthreads) {
  for (int i = tid; i < R * K; i += nthreads) {
    int n = i / K, k = i % K;
    float v = (n < rows && k < cols_used) ? W[n * cols_src + k] : 0.f;
    *reinterpret_cast<__half*>(dst + chunk_off<R>(n, k >> 3) + (k & 7) * 2) = __float2half_rn(v);
  }
}

}  // namespace tc
}  // namespace nemit
