// field_tc.cu - k_field_forward_tc: the main field with the 64-wide MLP chain on the 5th-generation tensor
// cores (tcgen05.mma, fp16 or split-bf16 operands, fp32 accumulation in TMEM).
//
// One group = 128 threads = one 128-sample tile at a time (thread m owns sample row m: it gathers the row's hash
// encoding, and after every layer reads ITS accumulator row back from TMEM lane m, applies bias / ReLU in fp32,
// and writes the row of the next layer's A operand).  The five GEMMs of a tile
//     enc[128x32] W_b0^T -> h1[128x64] W_b1^T -> [raw | geo15] ;  [SH16 | geo15 | 0] W_h0'^T -> h2 W_h1^T -> h3 W_h2^T
// are issued by one thread each (M = 128, N = 64 / 16, K = 32 / 64), the accumulator reuses the same 64 TMEM
// columns, the weights stay resident in shared memory in the UMMA K-major layout for the CTA's lifetime.
// Several groups per SM overlap one group's gather phase with another's MMA / epilogue phase.
// Density activation, SafeExp, weights scan and composite stay in fp32 (same code as the fp32 kernel).
//
// Operand precision (template NP):
//   NP = 1  fp16 operands: the reference's own mixed-precision mode (torch.autocast / tiny-cuda-nn FullyFusedMLP,
//           NS/models/base_model.py:305, NS/field_components/mlp.py:117-144), BASELINE.json configs[1]; one group per
//           CTA, four CTAs per SM.
//   NP = 3  fp32 arithmetic on the tensor cores: every operand as three bf16 pieces (24 significant bits), six MMAs per
//           layer (tc_common.cuh).  The weights take 60 KB of shared memory, so one CTA per SM holds them once for THREE
//           groups of 128 threads (named barriers), and the A operand does not go through shared memory at all: a thread
//           writes its row of the three pieces straight into TENSOR MEMORY (tcgen05.st; lane = row, 32 columns per
//           piece) and the MMAs take A from there (the .ts form) - 64 accumulator + 96 operand columns per group, 480
//           of the SM's 512.  The seed pass of this mode keeps the two leading pieces of E / H1 / IN (written by the
//           threads in the tile layout), the Jacobian in fp32 and the ReLU masks of the two head layers for
//           k_field_backward_tc.
//
// Reference: NerfactoField.get_density / get_outputs (NS/fields/nerfacto_field.py:228-334), MLP
// (NS/field_components/mlp.py:163-183).
#include "kernels.h"
#include "field_common.cuh"
#include "tc_common.cuh"

namespace nemit {

constexpr uint32_t kAPiece = 128 * 64 * 2;
constexpr uint32_t kAPieceCols = 32;            // A in tensor memory: 64 K-elements of 16 bits per row and piece
#ifndef NEMIT_SPLIT_ATM
#define NEMIT_SPLIT_ATM 1                       // split-operand mode: A operand in tensor memory (0: in shared memory, two groups)
#endif
constexpr bool kSplitAtm = NEMIT_SPLIT_ATM != 0;
#ifndef NEMIT_FP16_FWD_ATM
#define NEMIT_FP16_FWD_ATM 1                    // fp16 forward (MODE 0): A operand in tensor memory too, five CTAs per SM instead of four
#endif
constexpr bool kFp16FwdAtm = NEMIT_FP16_FWD_ATM != 0;
#ifndef NEMIT_FP16_SEED_ATM
#define NEMIT_FP16_SEED_ATM 1                   // the same for the fp16 seed pass (MODE 1; stays at four CTAs per SM: 128 registers)
#endif
constexpr bool kFp16SeedAtm = NEMIT_FP16_SEED_ATM != 0;
template <int MODE, int NP>
struct FwdCfg {
  static constexpr bool kAtm = (NP > 1 && kSplitAtm) || (NP == 1 && (MODE == 0 ? kFp16FwdAtm : kFp16SeedAtm));
  // fp16 operands: four CTAs of one group per SM (shared memory: weights + A operand = 47 KB each); with A in tensor
  // memory 31 KB and 64 + 32 TMEM columns each: five
#ifndef NEMIT_SEED_CTAS
#define NEMIT_SEED_CTAS 4
#endif
  static constexpr int kCtasPerSm = NP == 1 ? (kAtm ? (MODE == 0 ? 5 : NEMIT_SEED_CTAS) : 4) : 1;
};
constexpr int kSplitGroups = kSplitAtm ? 3 : 2;
template <int NP, int NG, bool ATM>
struct TcSmem {
  // B operands (UMMA K-major layout, rows = output features), piece p at p * (one-piece size)
  alignas(128) uint8_t b1[NP * 64 * 32 * 2];   // base layer 0   (64 x 32)
  alignas(128) uint8_t b2[NP * 16 * 64 * 2];   // base layer 1   (16 x 64)
  alignas(128) uint8_t b3[NP * 64 * 32 * 2];   // head layer 0   (64 x [SH16 | geo15 | 0])
  alignas(128) uint8_t b4[NP * 64 * 64 * 2];   // head layer 1   (64 x 64)
  alignas(128) uint8_t b5[NP * 16 * 64 * 2];   // head layer 2   (3 rows used of 16)
  struct Group {
    alignas(128) uint8_t a[ATM ? 128 : NP * kAPiece];   // A operand of the current layer (128 x K<=64) unless it lives in TMEM
    float sigma[kGroupSamples];
    float rgb[kGroupSamples * 3];
    float t_edge[kRaysPerGroup * (kS2 + 1)];
    float ray_o[kRaysPerGroup * 3];
    float ray_d[kRaysPerGroup * 3];
    float s_edge[kRaysPerGroup * (kS2 + 1)];   // adjoint-seed mode only
    alignas(8) uint64_t mbar;
  } g[NG];
  float bias_b0[64], bias_b1[16], bias_h0[64], bias_h1[64], bias_h2[4];
  float app_tmp[kApp];
  uint32_t tmem_slot;
  uint32_t tmem_slot_a;                        // fp16 operands with A in tensor memory: the A operand's own allocation (32 columns)
  int seed_max_bits;                           // adjoint-seed mode: max seed magnitude of this CTA (float bits)
};

// all threads of one group (the whole CTA when it has one group)
template <int NG>
__device__ __forceinline__ void group_sync(int grp) {
  if (NG == 1) __syncthreads();
  else asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
}

// where a thread's row of the next A operand goes: the shared-memory operand buffer (the stash copy is then made from
// there by the TMA engine, run_layer), or this thread's TMEM lane (ATM; the kept pieces are then stored from registers)
template <int NP, bool ATM>
struct AOut {
  uint8_t* smem;         // !ATM
  uint32_t taddr;        // ATM: column base of piece 0, this warp's lanes
  int row;
  uint8_t* keep;         // ATM: tile of the kept operand in the workspace, or null
  uint32_t keep_bytes;   // ATM: bytes of one kept piece
  template <bool RELU = false>
  __device__ __forceinline__ void put(int kc, const float (&x)[8]) const {
    if (!ATM) {
      tc::store_chunk<NP, 128, false, RELU>(smem, kAPiece, row, kc, x);
    } else {
      uint32_t w[4][3];
#pragma unroll
      for (int q = 0; q < 4; ++q) tc::pack_pair<NP, false, RELU>(x[2 * q], x[2 * q + 1], w[q]);
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        tc::tmem_st4(taddr + p * kAPieceCols + kc * 4, w[0][p], w[1][p], w[2][p], w[3][p]);
        if (keep && p < 2) __stcs(reinterpret_cast<uint4*>(keep + p * keep_bytes) + kc * 128 + row, make_uint4(w[0][p], w[1][p], w[2][p], w[3][p]));
      }
    }
  }
};

// accumulator row (64 columns) -> +bias, ReLU -> operand row of the next layer (8 K-chunks); returns the ReLU mask.
// The columns are walked in groups of kFwdEpiCols x 16; with fewer than 4 the loop stays rolled (code size: the tile loop
// is several times the 32 KB instruction cache, and the groups / CTAs of an SM sit in different phases of it)
#ifndef NEMIT_FWD_EPI_COLS
#define NEMIT_FWD_EPI_COLS 2
#endif
constexpr int kFwdEpiCols = NEMIT_FWD_EPI_COLS;
#if NEMIT_FWD_EPI_COLS == 4
#define NEMIT_FWD_EPI_LOOP _Pragma("unroll")
#else
#define NEMIT_FWD_EPI_LOOP _Pragma("unroll 1")
#endif
template <int NP, bool ATM>
__device__ __forceinline__ uint2 epilogue_hidden(uint32_t taddr, const float* __restrict__ bias, const AOut<NP, ATM>& out) {
  uint32_t m[2] = {0u, 0u};
  NEMIT_FWD_EPI_LOOP
  for (int c0 = 0; c0 < 4; c0 += kFwdEpiCols) {
    float v[kFwdEpiCols][16];
#pragma unroll
    for (int cc = 0; cc < kFwdEpiCols; ++cc) tc::tmem_ld16(taddr + (c0 + cc) * 16, v[cc]);
    tc::tmem_ld_wait();
#pragma unroll
    for (int cc = 0; cc < kFwdEpiCols; ++cc) {
      const int c = c0 + cc;
      uint32_t mc = 0u;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float x[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          x[q] = v[cc][hh * 8 + q] + bias[c * 16 + hh * 8 + q];          // the ReLU sits in the conversion (F2FP.RELU)
          if (NP > 1) mc |= (x[q] > 0.f ? 1u : 0u) << (hh * 8 + q);
        }
        out.template put<true>(c * 2 + hh, x);
      }
      if (NP > 1) {
        if (c >> 1) m[1] |= mc << ((c & 1) * 16);
        else m[0] |= mc << ((c & 1) * 16);
      }
    }
  }
  return make_uint2(m[0], m[1]);
}

// make this thread's A-operand writes visible to the tensor core, then let one thread of the group issue the layer
template <int N, int K, int NP, int NG, bool ATM>
__device__ __forceinline__ void run_layer(typename TcSmem<NP, NG, ATM>::Group* G, const uint8_t* b, uint32_t tmem, uint32_t a_tmem,
                                          uint32_t& phase, int gt, int grp, uint8_t* keep = nullptr, uint32_t keep_bytes = 0,
                                          uint64_t policy = 0) {
  if (ATM) {
    // the rows were written with tcgen05.st: complete them, order them before the barrier; D = columns [0, 64), A pieces
    // from column 64 (the kept pieces went out from the registers, AOut::put)
    tc::tmem_st_wait();
    tc::tc_fence_before();
    group_sync<NG>(grp);
    if (gt == 0) {
      tc::tc_fence_after();
      tc::mma_chain_np_ts<NP>(tmem, a_tmem, kAPieceCols, tc::op_kmajor(b, N), N * K * 2, tc::idesc_f16(128, N), K / 16, false);
      tc::mma_commit(&G->mbar);
    }
    tc::mbar_wait(&G->mbar, phase);
    phase ^= 1u;
    tc::tc_fence_after();
    return;
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  group_sync<NG>(grp);
  if (gt == 0) {
    tc::tc_fence_after();
    tc::mma_chain_np<NP>(tmem, tc::op_kmajor(G->a, 128), kAPiece, tc::op_kmajor(b, N), N * K * 2, tc::idesc_f16(128, N), K / 16, false);
    tc::mma_commit(&G->mbar);
    if (keep) {
      // adjoint-seed mode: the A operand (= this layer's input rows, already in the tile layout) goes to the workspace
      // through the TMA engine - the fp16 operand, or the two leading bf16 pieces; the operand buffer is reused once the
      // engine has read it
#pragma unroll
      for (int p = 0; p < (NP > 1 ? 2 : 1); ++p) tc::bulk_s2g(keep + p * keep_bytes, G->a + p * kAPiece, keep_bytes, policy);
      tc::bulk_wait_read();
    }
  }
  tc::mbar_wait(&G->mbar, phase);
  phase ^= 1u;
  tc::tc_fence_after();
  if (keep) group_sync<NG>(grp);
}

// MODE 0: forward.  MODE 1: first pass of the backward (same forward, then the per-ray composite adjoint).
template <int MODE, int NP, int NG>
__global__ void __launch_bounds__(128 * NG, FwdCfg<MODE, NP>::kCtasPerSm)
k_field_forward_tc(const __grid_constant__ ModelDev M, const __grid_constant__ FieldArgs A) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr bool ATM = FwdCfg<MODE, NP>::kAtm;  // A operand in tensor memory
  constexpr bool kTwoAllocs = ATM && NP == 1;  // accumulator (64 columns) and A operand (32) as two power-of-two allocations
  constexpr int kGroupCols = ATM ? 64 + NP * (int)kAPieceCols : 64;
  constexpr int kAllocCols = kTwoAllocs ? 64 : (NG * kGroupCols <= 64 ? 64 : (NG * kGroupCols <= 128 ? 128 : (NG * kGroupCols <= 256 ? 256 : 512)));
  static_assert(NG * kGroupCols <= 512, "tensor memory of one SM");
  using Smem = TcSmem<NP, NG, ATM>;
  Smem* sm = reinterpret_cast<Smem*>(smem_raw);
  constexpr int kThreads = 128 * NG;
  constexpr int SP = NP > 1 ? 2 : 1;           // operand pieces kept for the backward
  const int tid = threadIdx.x, lane = tid & 31, grp = tid >> 7, gt = tid & 127, wq = gt >> 5;
  typename Smem::Group* G = &sm->g[grp];

  // ---- one-time setup: weights -> operands, biases, effective head bias, TMEM, mbarriers ------------
  tc::stage_weight_np<64, 32, NP>(sm->b1, A.P.base_w0, 64, 32, 32, tid, kThreads);
  tc::stage_weight_np<16, 64, NP>(sm->b2, A.P.base_w1, 16, 64, 64, tid, kThreads);
  tc::stage_weight_np<64, 32, NP>(sm->b3, A.P.head_w0, 64, kHeadIn, kHeadK, tid, kThreads);
  tc::stage_weight_np<64, 64, NP>(sm->b4, A.P.head_w1, 64, 64, 64, tid, kThreads);
  tc::stage_weight_np<16, 64, NP>(sm->b5, A.P.head_w2, 3, 64, 64, tid, kThreads);
  for (int i = tid; i < 64; i += kThreads) { sm->bias_b0[i] = A.P.base_b0[i]; sm->bias_h1[i] = A.P.head_b1[i]; }
  for (int i = tid; i < 16; i += kThreads) sm->bias_b1[i] = A.P.base_b1[i];
  for (int i = tid; i < 4; i += kThreads) sm->bias_h2[i] = i < 3 ? A.P.head_b2[i] : 0.f;
  for (int i = tid; i < kApp; i += kThreads) {
    sm->app_tmp[i] = appearance_component(A.P.appearance, M.num_images, A.app_row, i);
  }
  if (gt == 0) tc::mbar_init(&G->mbar, 1);
  if (tid == 0) sm->seed_max_bits = 0;
  if (gt == 0) tc::fence_mbar_init();
  float seed_max = 0.f;
  if (tid < 32) {
    if (kTwoAllocs) {
      tc::tmem_alloc_only<64>(&sm->tmem_slot);
      tc::tmem_alloc_only<32>(&sm->tmem_slot_a);
      tc::tmem_relinquish();
    } else {
      tc::tmem_alloc<kAllocCols>(&sm->tmem_slot);
    }
  }
  __syncthreads();
  for (int n = tid; n < 64; n += kThreads) {
    float b = A.P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + k], sm->app_tmp[k], b);
    sm->bias_h0[n] = b;
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = sm->tmem_slot + (kTwoAllocs ? 0u : (uint32_t)grp * kGroupCols);   // this group's columns: accumulator, then A pieces
  const uint32_t a_tmem = kTwoAllocs ? sm->tmem_slot_a : tmem + 64;
  const uint32_t taddr = tmem + ((uint32_t)(wq * 32) << 16);     // this warp's 32 TMEM lanes
  uint32_t phase = 0;
  const uint64_t policy = MODE == 1 ? tc::l2_evict_first_policy() : 0;

  const int64_t n_groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  for (int64_t g = (int64_t)blockIdx.x * NG + grp; g < n_groups; g += (int64_t)gridDim.x * NG) {
    const int64_t r0 = g * kRaysPerGroup;
    const int rays_here = (int)min((int64_t)kRaysPerGroup, A.n - r0);
    group_sync<NG>(grp);
    if (gt < kRaysPerGroup) {
      float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
      if (gt < rays_here) load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + r0 + gt,
                                   ray_scale(M, A), o, d);
#pragma unroll
      for (int k = 0; k < 3; ++k) { G->ray_o[gt * 3 + k] = o[k]; G->ray_d[gt * 3 + k] = d[k]; }
    }
    for (int i = gt; i < kRaysPerGroup * (kS2 + 1); i += 128) {
      int rr = i / (kS2 + 1);
      G->t_edge[i] = rr < rays_here ? A.t_edges2[r0 * (kS2 + 1) + i] : 0.f;
      if (MODE == 1) G->s_edge[i] = rr < rays_here ? A.s_edges2[r0 * (kS2 + 1) + i] : 0.f;
    }
    group_sync<NG>(grp);

    for (int tile = 0; tile < kGroupSamples / kTileM; ++tile) {
      const int s = tile * kTileM + gt;
      const int rr = s / kS2, si = s % kS2;
      const float t0 = G->t_edge[rr * (kS2 + 1) + si], t1 = G->t_edge[rr * (kS2 + 1) + si + 1];
      const float tsum = add_rn(t0, t1);
      float x[3], p[3], dirq[3];
      rot_sample(A.rot, &G->ray_o[rr * 3], &G->ray_d[rr * 3], tsum, x, dirq);      // Rotater.apply_frustums when set for this stage
      const bool sel = contract_normalise(x, M, p) && rr < rays_here;
      const int64_t tile_id = g * (kGroupSamples / kTileM) + tile;
      AOut<NP, ATM> aout{G->a, a_tmem + ((uint32_t)(wq * 32) << 16), gt, nullptr, 0};
      // ---- hash encoding: 4 levels = 8 values = one 16-byte K-chunk of this thread's A row ------------
#pragma unroll 1
      for (int l0 = 0; l0 < kMainLevels; l0 += 4) {
        LevelCell cell[4];
        float2 v[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          level_cell<true>(p, M.main_grid.scal[l0 + j], M.main_grid.table_mask, (uint32_t)(l0 + j) * M.main_grid.table_size, cell[j]);
          gather_cell(A.P.main_table, cell[j], v[j]);
        }
        float e8[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 r = blend_cell(v[j], cell[j]);
          e8[2 * j] = r.x; e8[2 * j + 1] = r.y;
        }
        aout.keep = (ATM && MODE == 1 && A.enc_tiles) ? A.enc_tiles + tile_id * (SP * kEncTileBytes) : nullptr;
        aout.keep_bytes = 8192;
        aout.put(l0 >> 2, e8);
        if (MODE == 1 && A.enc_tiles) {
          // keep the row's Jacobian for the backward kernel (coalesced 16-byte stores, chunk-major tiles): per level
          // dx.xy dy.xy dz.xy, as fp16 pairs (NP = 1: 12 chunks per row) or in fp32 (split operands: 24 chunks per row)
          if (NP == 1) {
            uint32_t wj[12];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float2 dx, dy, dz;
              blend_cell_grad(v[j], cell[j], dx, dy, dz);
              wj[3 * j + 0] = tc::pack_h2_sat(dx.x, dx.y);
              wj[3 * j + 1] = tc::pack_h2_sat(dy.x, dy.y);
              wj[3 * j + 2] = tc::pack_h2_sat(dz.x, dz.y);
            }
            uint4* jp = reinterpret_cast<uint4*>(A.jac_tiles + tile_id * kJacTileBytes) + (l0 >> 2) * 3 * 128 + gt;
            __stcs(jp, make_uint4(wj[0], wj[1], wj[2], wj[3]));
            __stcs(jp + 128, make_uint4(wj[4], wj[5], wj[6], wj[7]));
            __stcs(jp + 256, make_uint4(wj[8], wj[9], wj[10], wj[11]));
          } else {
            float4* jp = reinterpret_cast<float4*>(A.jac_tiles + tile_id * (2 * kJacTileBytes)) + (l0 >> 2) * 6 * 128 + gt;
#pragma unroll
            for (int j = 0; j < 4; j += 2) {
              float2 dx0, dy0, dz0, dx1, dy1, dz1;
              blend_cell_grad(v[j], cell[j], dx0, dy0, dz0);
              blend_cell_grad(v[j + 1], cell[j + 1], dx1, dy1, dz1);
              __stcs(jp + (3 * (j >> 1) + 0) * 128, make_float4(dx0.x, dx0.y, dy0.x, dy0.y));
              __stcs(jp + (3 * (j >> 1) + 1) * 128, make_float4(dz0.x, dz0.y, dx1.x, dx1.y));
              __stcs(jp + (3 * (j >> 1) + 2) * 128, make_float4(dy1.x, dy1.y, dz1.x, dz1.y));
            }
          }
        }
      }
      // adjoint-seed mode keeps the activations (operand rows) for k_field_backward_tc: [H1 pieces | IN pieces]
      uint8_t* act = nullptr;
      if (MODE == 1 && A.act_tiles) act = A.act_tiles + tile_id * (SP * kActTileBytes);
      // ---- base MLP ------------------------------------------------------------------------------------
      run_layer<64, 32, NP, NG, ATM>(G, sm->b1, tmem, a_tmem, phase, gt, grp, act ? A.enc_tiles + tile_id * (SP * kEncTileBytes) : nullptr, 8192, policy);
      aout.keep = ATM ? act : nullptr;
      aout.keep_bytes = 16384;
      epilogue_hidden<NP, ATM>(taddr, sm->bias_b0, aout);
      run_layer<16, 64, NP, NG, ATM>(G, sm->b2, tmem, a_tmem, phase, gt, grp, act, 16384, policy);
      {
        float v[16];
        tc::tmem_ld16(taddr, v);
        tc::tmem_ld_wait();
        const float raw = v[0] + sm->bias_b1[0];
        G->sigma[s] = sel ? expf(raw) : 0.f;                          // trunc_exp * selector, fp32
        float c[16];
        sh16((dirq[0] + 1.f) * 0.5f, (dirq[1] + 1.f) * 0.5f, (dirq[2] + 1.f) * 0.5f, c);
        aout.keep = (ATM && act) ? act + SP * 16384 : nullptr;
        aout.keep_bytes = 8192;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float x8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) x8[q] = c[hh * 8 + q];
          aout.put(hh, x8);
        }
        float gfe[16];
#pragma unroll
        for (int k = 0; k < 15; ++k) gfe[k] = v[k + 1] + sm->bias_b1[k + 1];
        gfe[15] = MODE == 1 ? 1.f : 0.f;   // pad column (zero weights); the backward's wgrad reads it as a ones column
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float x8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) x8[q] = gfe[hh * 8 + q];
          aout.put(2 + hh, x8);
        }
        aout.keep = nullptr;
      }
      // ---- head MLP ------------------------------------------------------------------------------------
      run_layer<64, 32, NP, NG, ATM>(G, sm->b3, tmem, a_tmem, phase, gt, grp, act ? act + SP * 16384 : nullptr, 8192, policy);
      const uint2 m2 = epilogue_hidden<NP, ATM>(taddr, sm->bias_h0, aout);
      // H2 / H3 are not kept: the backward recomputes them from IN (kernels.h); with split operands its recompute has
      // 16 significant bits where this pass has 24, so the ReLU masks decided HERE are handed over
      run_layer<64, 64, NP, NG, ATM>(G, sm->b4, tmem, a_tmem, phase, gt, grp);
      const uint2 m3 = epilogue_hidden<NP, ATM>(taddr, sm->bias_h1, aout);
      if (MODE == 1 && NP > 1 && A.masks4) __stcs(A.masks4 + tile_id * 128 + gt, make_uint4(m2.x, m2.y, m3.x, m3.y));
      run_layer<16, 64, NP, NG, ATM>(G, sm->b5, tmem, a_tmem, phase, gt, grp);
      {
        float v[16];
        tc::tmem_ld16(taddr, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 3; ++j)
          G->rgb[s * 3 + j] = expf(fminf(v[j] + sm->bias_h2[j] + M.rgb_bias, 88.f));   // SafeExp
      }
      // the next tile's gather overwrites A: the last MMA has completed (waited above); TMEM loads are done
      tc::tc_fence_before();
      group_sync<NG>(grp);
    }

    // composite: warp w handles rays w and w+4; lane handles samples lane and lane+32 (fp32, as field.cu)
    for (int rr = wq; rr < rays_here; rr += 4) {
      if (MODE == 1) {
        seed_max = fmaxf(seed_max, composite_adjoint_ray<true>(M, A, &G->t_edge[rr * (kS2 + 1)], &G->s_edge[rr * (kS2 + 1)],
                                                         &G->sigma[rr * kS2], &G->rgb[rr * kS2 * 3], r0 + rr, lane));
        continue;
      }
      float ds[2], w[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        ds[k] = 0.f;
        if (si < kS2) {
          float t0 = G->t_edge[rr * (kS2 + 1) + si], t1 = G->t_edge[rr * (kS2 + 1) + si + 1];
          ds[k] = mul_rn(sub_rn(t1, t0), G->sigma[rr * kS2 + si]);
        }
      }
      weights_from_ds<2>(ds, w, lane);
      float cr = 0.f, cg = 0.f, cb = 0.f, acc = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        if (si < kS2) {
          const float* c = &G->rgb[(rr * kS2 + si) * 3];
          float r = c[0], gg = c[1], b = c[2];
          if (!A.training) { r = nan_to_num(r); gg = nan_to_num(gg); b = nan_to_num(b); }
          cr = fmaf(w[k], r, cr); cg = fmaf(w[k], gg, cg); cb = fmaf(w[k], b, cb);
          acc += w[k];
          if (A.weights2) A.weights2[(r0 + rr) * kS2 + si] = w[k];
        }
      }
      cr = warp_sum(cr); cg = warp_sum(cg); cb = warp_sum(cb); acc = warp_sum(acc);
      if (lane == 0) {
        const float* last = &G->rgb[(rr * kS2 + kS2 - 1) * 3];
        float lr = last[0], lg = last[1], lb = last[2];
        if (!A.training) { lr = nan_to_num(lr); lg = nan_to_num(lg); lb = nan_to_num(lb); }
        float bg = 1.f - acc;
        cr += lr * bg; cg += lg * bg; cb += lb * bg;
        if (!A.training) { cr = fmaxf(cr, 0.f); cg = fmaxf(cg, 0.f); cb = fmaxf(cb, 0.f); }
        A.rgb[(r0 + rr) * 3 + 0] = cr; A.rgb[(r0 + rr) * 3 + 1] = cg; A.rgb[(r0 + rr) * 3 + 2] = cb;
        if (A.accumulation) A.accumulation[r0 + rr] = acc;
      }
    }
  }
  if (MODE == 1 && A.g_absmax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) seed_max = fmaxf(seed_max, __shfl_xor_sync(0xffffffffu, seed_max, o));
    if (lane == 0) atomicMax(&sm->seed_max_bits, __float_as_int(seed_max));   // non-negative floats order as ints
  }
  tc::tc_fence_before();
  __syncthreads();
  if (MODE == 1 && A.g_absmax && tid == 0) atomicMax(reinterpret_cast<int*>(A.g_absmax), sm->seed_max_bits);
  if (tid < 32) {
    tc::tmem_dealloc<kAllocCols>(sm->tmem_slot);
    if (kTwoAllocs) tc::tmem_dealloc<32>(sm->tmem_slot_a);
  }
}

template <int MODE, int NP, int NG>
static cudaError_t launch_tc_mode(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(TcSmem<NP, NG, FwdCfg<MODE, NP>::kAtm>) + 128;
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_forward_tc<MODE, NP, NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int64_t groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  // persistent: the groups of an SM (4 CTAs of one with fp16 operands, one CTA of two with split operands) share its
  // tensor core
  const int64_t want = (groups + NG - 1) / NG, cap = 148 * FwdCfg<MODE, NP>::kCtasPerSm;
  k_field_forward_tc<MODE, NP, NG><<<(unsigned)(want < cap ? want : cap), 128 * NG, smem, stream>>>(M, A);
  return cudaGetLastError();
}
cudaError_t launch_field_forward_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream, int pieces) {
  return pieces > 1 ? launch_tc_mode<0, 3, kSplitGroups>(M, A, stream) : launch_tc_mode<0, 1, 1>(M, A, stream);
}
cudaError_t launch_field_adjoint_seed_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream, int pieces) {
  return pieces > 1 ? launch_tc_mode<1, 3, kSplitGroups>(M, A, stream) : launch_tc_mode<1, 1, 1>(M, A, stream);
}

}  // namespace nemit
