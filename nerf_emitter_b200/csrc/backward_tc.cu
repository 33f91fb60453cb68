// backward_tc.cu - k_field_backward_tc: reverse mode of the main field (48 final samples per ray) with every
// dense contraction on the 5th-generation tensor cores (tcgen05.mma, fp16 operands, fp32 accumulation in TMEM).
//
// The seed pass (k_field_forward_tc<1>, field_tc.cu) has already run the forward of the chunk on the tensor cores and
// kept, per 128-sample tile and in the UMMA operand layout: the hash encoding E, its Jacobian d enc / d offset, the
// operand rows H1 and IN = [SH16 | geo15 | 1], and per sample the seed d loss / d pre-activation (rgb xyz with the
// SafeExp derivative, density w with the trunc_exp derivative and the selector): 464 B per sample.  So this kernel
// never gathers from the parameter table (its L2 working set is the gradient table alone).  The two head hidden layers
// H2, H3 (256 B per sample in round 1) are recomputed here from IN with two forward MMAs + epilogues per tile: the
// compute halves wait for the scatter warps anyway, and the same fp16 operands through the same instruction shapes
// give the same bits, hence the same ReLU masks, as the seed pass had.
//
// One CTA = 256 threads = two independent halves of 128 threads; a half owns one tile at a time (thread t = sample
// row t).  Per tile:
//   load     : E, H1, IN arrive by three 1-D bulk copies (cp.async.bulk -> mbarrier complete_tx) straight into
//              the operand buffers; the copies of the NEXT tile are issued as soon as a buffer's last reader (an MMA
//              commit group) has completed, so they fly under the remaining layers and the scatter
//   backward : for every layer, in ONE commit group,
//                 dgrad  G_in [128 x K]  = G_out [128 x N] W        (A = G_out K-major, B = the FORWARD weight
//                                                                    buffer read MN-major: no transposed copy)
//                 wgrad  dW  [N x K]    += G_out^T X                (A = G_out buffer read MN-major, M = 64,
//                                                                    B = activation buffer read MN-major, K = 128
//                                                                    samples; accumulator lives in TMEM for the
//                                                                    whole kernel)
//                 bias   db  [N]        += G_out^T 1                (a constant ones operand; the head input
//                                                                    carries a ones column instead)
//              the epilogue masks the accumulator row with (activation > 0) and writes the fp16 gradient row over
//              the activation row
//   scatter  : d enc (TMEM) -> red.add on the gradient table (x-neighbour corners that share an aligned 16-byte pair
//              go out as one red.v4.f32), d enc . Jacobian -> d/d(position) -> per-ray sums.  This phase is bound by
//              the SM's reduction issue rate (1.5 cycles per lane, tools/red_bench.cu).
// Warp specialisation (round 2): the table scatter does not run on the compute halves any more.  A half parks the tile's
// d-encoding rows and sample positions in shared memory (one column per sample) and signals an mbarrier; a separate
// set of scatter warps (two warpgroups: a thread per parked sample in each, group g taking the levels l with l % 2 == g;
// 64 registers each after setmaxnreg.dec) drains them into red.global.add while the MMA chains run, serving the two halves
// alternately so that a full slot is always waiting, and hands the slot back through a second mbarrier.  Measured per
// 2^16 rays: round-1 kernel 1.93 ms; one scatter warpgroup 1.86; + run aggregation 1.61; + early exit of the scan 1.54;
// two scatter warpgroups 1.42 (the scatter warps are instruction-bound, ncu: profiles/r2_ncu_backward_ws_agg_summary.txt).  The compute halves (setmaxnreg.inc) only ever wait for their own
// commit groups and for a free slot.
// What bounds the scatter (measured, tools/red_bench.cu SM sweep, profiles/r2_red_bench.txt): red.global.add is limited
// CHIP-WIDE - 184 G lane-ops/s whether 18 or 148 SMs issue them, whatever the width up to 16 bytes - i.e. by the L2
// atomic units, not by the SM's issue rate as round 1 assumed.  So the only lever is the NUMBER of lane-ops:
//   * x-neighbour corners of an aligned pair go out as one red.v4.f32 (round 1),
//   * consecutive samples of a ray that fall into the same grid cell (coarse levels everywhere, all levels where the
//     sampler has concentrated the samples) are summed across the lanes of a scatter warp with a segmented shuffle scan
//     and issued once by the last lane of the run,
//   * the six per-ray sums d/d(origin, direction) are reduced over the samples of a ray inside the warp before the
//     atomics (32 x 6 lane-ops per warp -> 6 or 12).
// Gradient operands are fp16 as in the reference's autocast path (torch.cuda.amp, NS/pipelines/mitsuba_sdf.py:
// 1177,1193): the seed pass records max |seed| of the launch, the seeds are multiplied by the power of two that
// brings that maximum to 16, and every result is divided by it again (exact), so the fp16 range is used without
// an externally tuned loss scale.
//
// Reference: SdfNerfactoModel.get_backward_outputs (NS/models/sdf_nerfacto.py:115-168),
// MitsubaSdfPipeline.backward_for_camera_ray_bundle (NS/pipelines/mitsuba_sdf.py:1144-1219),
// NerfactoField.get_density / get_outputs (NS/fields/nerfacto_field.py:228-334), _TruncExp.backward
// (NS/field_components/activations.py:44-48).
#include "kernels.h"
#include "field_common.cuh"
#include "tc_common.cuh"

namespace nemit {
namespace {

// TMEM columns of one half (256 per half, 512 per CTA = the whole SM)
constexpr uint32_t kColD0 = 0;      // 128 x 64 work accumulator (forward layers and dgrads)
constexpr uint32_t kColH1 = 64;     // d head_w1   [n][k]   64 x 64   (M = 64: row r in lane (r/16)*32 + r%16)
constexpr uint32_t kColH0 = 128;    // d head_w0   [n][k]   64 x 32   (column 31 = d head_b0_eff)
constexpr uint32_t kColB0 = 160;    // d base_w0   [n][k]   64 x 32
constexpr uint32_t kColB1T = 192;   // d base_w1^T [k][n]   64 x 16
constexpr uint32_t kColH2T = 208;   // d head_w2^T [k][n]   64 x 8
constexpr uint32_t kColBH1 = 216;   // d head_b1   [n][*]   64 x 8 (8 identical columns)
constexpr uint32_t kColBB0 = 224;   // d base_b0   [n][*]   64 x 8
constexpr uint32_t kColBB1 = 232;   // d base_b1   [*][n]   64 x 16 (64 identical rows: A = ones)
constexpr uint32_t kColBH2 = 248;   // d head_b2   [*][n]   64 x 8
constexpr uint32_t kHalfCols = 256;
#ifndef NEMIT_SCATTER_UNROLL
#define NEMIT_SCATTER_UNROLL 2
#endif
constexpr int kScatterUnroll = NEMIT_SCATTER_UNROLL;
// scatter warps: 128 threads = one sample of the parked tile each; they serve the two compute halves alternately
#ifndef NEMIT_SCATTER_GROUPS
#define NEMIT_SCATTER_GROUPS 2         // scatter warpgroups: group g takes the levels l with l % groups == g
#endif
constexpr int kScatterGroups = NEMIT_SCATTER_GROUPS;
constexpr int kScatterThreads = 128 * kScatterGroups;
// NP = 1: fp16 operands, two compute halves.  NP = 2: the fp32-accurate mode (bf16 hi | lo operand pieces, tc_common.cuh):
// the operand buffers are twice as large, so ONE compute half per CTA (its chain is not overlapped with a second one; the
// scatter warps still run beside it).
// Register split: the CTA is launched with kLaunchRegs per thread (65536 / threads rounded down to a multiple of 8), i.e.
// a pool of threads x kLaunchRegs; the scatter warpgroups shrink and the compute warpgroups grow out of what they released
// (setmaxnreg).  The split must fit the POOL, not the register file: a setmaxnreg.inc that cannot be served blocks for ever.
template <int NP>
struct BwdCfg {
  static constexpr int kHalves = NP == 1 ? 2 : 1;
  static constexpr int kComputeThreads = 128 * kHalves;
  static constexpr int kThreads = kComputeThreads + kScatterThreads;
  static constexpr int kLaunchRegs = kThreads == 640 ? 96 : (kThreads == 512 ? 128 : (kThreads == 384 ? 168 : 255));
  static constexpr int kComputeRegs = NP == 1 ? (kScatterGroups == 1 ? 200 : (kScatterGroups == 2 ? 192 : 144)) : 232;
  static constexpr int kScatterRegs = NP == 1 ? (kScatterGroups == 1 ? 104 : 64) : 64;
  static_assert(kThreads == 640 || kThreads == 512 || kThreads == 384, "launch register count is derived for 384 / 512 / 640 threads");
  static_assert(kComputeThreads * kComputeRegs + kScatterThreads * kScatterRegs <= kThreads * kLaunchRegs, "register pool of the CTA");
};
#ifndef NEMIT_BWD_AGGREGATE
#define NEMIT_BWD_AGGREGATE 1          // sum the contributions of runs of lanes that fall into the same grid cell before the red
#endif

// one-piece sizes of the operand buffers (piece p of a buffer sits p * size after piece 0)
constexpr uint32_t kPcB1 = 64 * 32 * 2, kPcB2 = 16 * 64 * 2, kPcB3 = 64 * 32 * 2, kPcB4 = 64 * 64 * 2, kPcB5 = 16 * 64 * 2;
constexpr uint32_t kPcE = 128 * 32 * 2, kPcH = 128 * 64 * 2, kPcIn = 128 * 32 * 2, kPcGs = 128 * 16 * 2;
template <int NP>
struct BwdTcCommon {
  alignas(128) uint8_t b1[NP * kPcB1];    // base layer 0   (64 x 32)   forward K-major layout, rows = outputs
  alignas(128) uint8_t b2[NP * kPcB2];    // base layer 1   (16 x 64)
  alignas(128) uint8_t b3[NP * kPcB3];    // head layer 0   (64 x [SH16 | geo15 | 0])
  alignas(128) uint8_t b4[NP * kPcB4];    // head layer 1   (64 x 64)
  alignas(128) uint8_t b5[NP * kPcB5];    // head layer 2   (3 rows used of 16)
  alignas(128) uint8_t ones[8 * 128 * 2]; // MN-major 8 x 128 operand of ones (fp16 or bf16: one piece, exact)
  float app[kApp];
  float bias_h0[64];                      // head layer 0 bias with the appearance columns folded in (as the forward)
  float bias_h1[64];
  uint32_t tmem_slot;
};
template <int NP>
struct BwdTcHalf {
  alignas(128) uint8_t e[NP * kPcE];      // hash encoding
  alignas(128) uint8_t h1[NP * kPcH];     // base hidden         -> its gradient
  alignas(128) uint8_t in[NP * kPcIn];    // [SH16 | geo15 | 1]
  alignas(128) uint8_t h2[NP * kPcH];     // head hidden 1       -> its gradient
  alignas(128) uint8_t h3[NP * kPcH];     // head hidden 2       -> its gradient
  alignas(128) uint8_t gs[NP * kPcGs];    // d rgb pre-activation (3 of 16), later d [raw | geo15]
  float dvec[64];
  float denc[32 * 128];                   // d encoding of the tile handed to the scatter warps, [column][sample]
  float pbuf[3 * 128];                    // its normalised sample positions, [axis][sample]
  alignas(8) uint64_t mbar;               // MMA commit groups
  alignas(8) uint64_t ld_bar;             // bulk copies of one tile (65536 bytes per phase)
  alignas(8) uint64_t full_bar;           // 128 compute threads arrive: denc / pbuf of a tile are parked
  alignas(8) uint64_t empty_bar;          // the half's scatter threads arrive: the slot may be overwritten
};
template <int NP>
struct BwdTcSmem {
  BwdTcCommon<NP> c;
  BwdTcHalf<NP> h[BwdCfg<NP>::kHalves];
};

__device__ __forceinline__ void half_sync(int half) {
  asm volatile("bar.sync %0, 128;" ::"r"(half + 1) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}

// make the half's operand writes visible to the tensor core, let the issuers put the step's MMAs in flight, wait for all.
// The chains of one step (weight gradient, bias gradient, data gradient) write DIFFERENT accumulators, so they are issued
// by lane 0 of kIssuers different warps at the same time - one thread issues ~10 instructions per MMA (descriptors into
// uniform registers), which at up to 52 MMAs per step was a visible part of the chain; every issuer commits to the same
// mbarrier (count kIssuers).  A given accumulator is always fed by the same thread, so its MMAs stay ordered.
#ifndef NEMIT_BWD_ISSUERS
#define NEMIT_BWD_ISSUERS 3
#endif
constexpr int kIssuers = NEMIT_BWD_ISSUERS;
// which chains keep their k-step loops rolled (weight-gradient / bias-gradient chains: issued beside the data-gradient
// chain by their own threads, so their issue rate is off the critical path; the instruction footprint is not)
#ifndef NEMIT_BWD_ROLL_W
#define NEMIT_BWD_ROLL_W 0
#endif
#ifndef NEMIT_BWD_ROLL_B
#define NEMIT_BWD_ROLL_B 0
#endif
constexpr bool kRollW = NEMIT_BWD_ROLL_W != 0, kRollB = NEMIT_BWD_ROLL_B != 0;
static_assert(kIssuers == 1 || kIssuers == 3, "one issuing thread, or one per chain");
template <class F>
__device__ __forceinline__ void mma_step_on(uint64_t* mbar, int group, int t, uint32_t& phase, F&& issue) {
  tc::fence_proxy_async();
  tc::tc_fence_before();
  half_sync(group);
  if ((t & 31) == 0 && t < 32 * kIssuers) {
    tc::tc_fence_after();
    if (kIssuers == 1) { issue(0); issue(1); issue(2); }
    else issue(t >> 5);
    tc::mma_commit(mbar);
  }
  tc::mbar_wait(mbar, phase);
  phase ^= 1u;
  tc::tc_fence_after();
}
template <class HalfT, class F>
__device__ __forceinline__ void mma_step(HalfT* H, int half, int t, uint32_t& phase, F&& issue) {
  mma_step_on(&H->mbar, half, t, phase, issue);
}

// Table scatter of one parked tile, all 16 levels: run by the four scatter warps, thread = sample `smp` of the tile
// (so a warp holds 32 consecutive samples, i.e. runs of consecutive samples of a ray).  Per level:
//   * an all-zero gradient pair (also: invalid rows) contributes nothing;
//   * lanes whose samples fall into the SAME grid cell (equal integer cell coordinates => equal 8 table indices) form a
//     run; the 16 corner contributions are summed over the run with a segmented shuffle scan and the last lane issues them;
//   * corners (2q, 2q+1) differ in x only (ceil, floor); for an even floor coordinate the two hashed indices differ in
//     bit 0 only, i.e. they are the two halves of one aligned 16-byte pair: one red.v4.f32.
template <class HalfT>
__device__ __forceinline__ void scatter_tile(const HalfT* H, int smp, int lane, int group, int groups, const ModelDev& M,
                                             float2* table, float gscale, uint32_t& nred) {
  const float p[3] = {H->pbuf[smp], H->pbuf[128 + smp], H->pbuf[256 + smp]};
#ifndef NEMIT_BWD_SCATTER_FROM
#define NEMIT_BWD_SCATTER_FROM 0       // experiments only: first level that is scattered
#endif
#ifndef NEMIT_BWD_SCATTER_TO
#define NEMIT_BWD_SCATTER_TO 16        // experiments only: one past the last level that is scattered
#endif
#pragma unroll 1
  for (int l = NEMIT_BWD_SCATTER_FROM + group; l < NEMIT_BWD_SCATTER_TO; l += groups) {
    const float gx = H->denc[(2 * l) * 128 + smp], gy = H->denc[(2 * l + 1) * 128 + smp];
    const bool act = gx != 0.f || gy != 0.f;
    if (!__any_sync(0xffffffffu, act)) continue;
    LevelCell cell;
    const float sc = M.main_grid.scal[l];
    uint32_t c0[3];
    level_cell<true>(p, sc, M.main_grid.table_mask, (uint32_t)l * M.main_grid.table_size, cell, c0);
    float cw[8];
    corner_weights(cell, cw);
    const float gxs = gx * gscale, gys = gy * gscale;
    const uint32_t x0 = c0[0];
    const bool pair = ((x0 & 1) == 0) && (cell.idx[0] != cell.idx[1]);
    float w[16];                         // (corner c, feature f) -> w[2c + f]
#pragma unroll
    for (int c = 0; c < 8; ++c) { w[2 * c] = cw[c] * gxs; w[2 * c + 1] = cw[c] * gys; }
    bool issue = act;
#if NEMIT_BWD_AGGREGATE
    {
      // cell key: the integer cell coordinates, 11 + 11 + 10 bits in one word; on the levels whose resolution exceeds 1024
      // the eleventh bit of z travels in a second word.  Lanes that must stay alone get a key of their own: inactive ones,
      // and samples that sit EXACTLY on a grid plane (a zero fractional coordinate): their ceil corner coincides with the
      // floor corner, so their 8 indices are not those of the cell's other samples although the integer coordinates agree -
      // as the last lane of a run they would send the run's ceil-corner sums to the floor-corner entries (seen on 1 of
      // 49 152 sample-levels of a 64-ray batch, tools/diag_table_grad.py).
      const uint32_t y0 = c0[1], z0 = c0[2];
      const bool alone = !act || cell.fx == 0.f || cell.fy == 0.f || cell.fz == 0.f;
      const uint32_t key = alone ? (0xffffffe0u | (uint32_t)lane) : (x0 | (y0 << 11) | (z0 << 22));
      const uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);        // (every lane takes part: no short-circuit around it)
      bool head = lane == 0 || key != prev;
      if (sc > 1023.f) {                 // warp-uniform
        const uint32_t zhi = alone ? 0xffu : (z0 >> 10);
        const uint32_t prev_zhi = __shfl_up_sync(0xffffffffu, zhi, 1);
        head = head || zhi != prev_zhi;
      }
      const uint32_t heads = __ballot_sync(0xffffffffu, head);
      if (heads != 0xffffffffu) {        // at least one run of two or more lanes in this warp
        const int start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));      // first lane of my run
        const bool tail = lane == 31 || ((heads >> (lane + 1)) & 1u);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const bool take = lane - d >= start;
          if (!__any_sync(0xffffffffu, take)) break;       // no run of this warp is longer than d
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float up = __shfl_up_sync(0xffffffffu, w[i], d);
            if (take) w[i] += up;
          }
        }
        issue = act && tail;             // every lane of a multi-lane run is active; the last one holds the run's sums
      }
    }
#endif
    if (!issue) continue;
    nred += pair ? 4u : 8u;
    if (pair) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t ia = cell.idx[2 * q], ib = cell.idx[2 * q + 1];
        const bool odd = ia & 1u;        // the ceil-x corner sits at the odd entry of the pair
        const float ax = odd ? w[4 * q + 2] : w[4 * q], ay = odd ? w[4 * q + 3] : w[4 * q + 1];
        const float bx = odd ? w[4 * q] : w[4 * q + 2], by = odd ? w[4 * q + 1] : w[4 * q + 3];
        atomicAdd(reinterpret_cast<float4*>(table + (ia < ib ? ia : ib)), make_float4(ax, ay, bx, by));
      }
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) atomic_add_f2(table + cell.idx[c], make_float2(w[2 * c], w[2 * c + 1]));
    }
  }
}

// The epilogues walk the 64 accumulator columns in groups of kEpiCols x 16; with fewer than 4 the loop stays rolled
// (code size: the kernel's instruction footprint is far beyond the instruction cache, ncu: no_inst stalls)
#ifndef NEMIT_EPI_COLS
#define NEMIT_EPI_COLS 1
#endif
constexpr int kEpiCols = NEMIT_EPI_COLS;
#if NEMIT_EPI_COLS == 4
#define NEMIT_EPI_LOOP _Pragma("unroll")
#else
#define NEMIT_EPI_LOOP _Pragma("unroll 1")
#endif

// forward recompute of a head hidden layer: accumulator row + bias, ReLU -> operand row (k_field_forward_tc's
// epilogue_hidden, same arithmetic in the same order)
// With split operands (NP = 2) the ReLU decision is the seed pass's (24 significant bits there, 16 here): unit n is active
// iff bit n of `mask` is set, whatever the sign of this pass's own pre-activation (they differ only within ~1e-5 of zero).
template <int NP>
__device__ __forceinline__ void epi_fwd_hidden(uint32_t taddr, const float* __restrict__ bias, uint8_t* buf, int row, uint2 mask) {
  NEMIT_EPI_LOOP
  for (int c0 = 0; c0 < 4; c0 += kEpiCols) {
    float v[kEpiCols][16];
#pragma unroll
    for (int cc = 0; cc < kEpiCols; ++cc) tc::tmem_ld16(taddr + (c0 + cc) * 16, v[cc]);
    tc::tmem_ld_wait();
#pragma unroll
    for (int cc = 0; cc < kEpiCols; ++cc) {
      const int c = c0 + cc;
      const uint32_t mw = ((c >> 1) ? mask.y : mask.x) >> ((c & 1) * 16);
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float x[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          x[q] = v[cc][hh * 8 + q] + bias[c * 16 + hh * 8 + q];           // ReLU: in the conversion (NP = 1), by the mask (NP = 2)
          if (NP > 1) x[q] = ((mw >> (hh * 8 + q)) & 1u) ? fmaxf(x[q], 0.f) : 0.f;
        }
        tc::store_chunk<NP, 128, false, (NP == 1)>(buf, kPcH, row, c * 2 + hh, x);
      }
    }
  }
}

// hidden layer: accumulator row x (activation row > 0) -> gradient row written over the activation row
// (USE_MASK: the activity comes from the seed pass's mask bits instead, see epi_fwd_hidden)
template <int NP, bool USE_MASK = false>
__device__ __forceinline__ void epi_bwd_hidden(uint32_t taddr, uint8_t* buf, int row, uint2 mask = make_uint2(0u, 0u)) {
  NEMIT_EPI_LOOP
  for (int c0 = 0; c0 < 4; c0 += kEpiCols) {
    float v[kEpiCols][16];
#pragma unroll
    for (int cc = 0; cc < kEpiCols; ++cc) tc::tmem_ld16(taddr + (c0 + cc) * 16, v[cc]);
    tc::tmem_ld_wait();
#pragma unroll
    for (int cc = 0; cc < kEpiCols; ++cc) {
      const int c = c0 + cc;
      const uint32_t mc = ((c >> 1) ? mask.y : mask.x) >> ((c & 1) * 16);
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        uint32_t aw[4] = {0u, 0u, 0u, 0u};
        if (!USE_MASK) {
          const uint4 act = *reinterpret_cast<const uint4*>(buf + tc::chunk_off<128>(row, c * 2 + hh));
          aw[0] = act.x; aw[1] = act.y; aw[2] = act.z; aw[3] = act.w;
        }
        float x[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int j = hh * 8 + q * 2;
          if (USE_MASK) {
            const uint32_t mw = mc >> j;
            x[2 * q] = (mw & 1u) ? v[cc][j] : 0.f;
            x[2 * q + 1] = (mw & 2u) ? v[cc][j + 1] : 0.f;
          } else {
            // post-ReLU activations are >= 0: a non-zero fp16 / bf16-hi pattern (sign bit aside) means the unit was active
            x[2 * q] = (aw[q] & 0x00007fffu) ? v[cc][j] : 0.f;
            x[2 * q + 1] = (aw[q] & 0x7fff0000u) ? v[cc][j + 1] : 0.f;
          }
        }
        tc::store_chunk<NP, 128, true>(buf, kPcH, row, c * 2 + hh, x);
      }
    }
  }
}

}  // namespace

template <int NP>
__global__ void __launch_bounds__(BwdCfg<NP>::kThreads, 1)
k_field_backward_tc(const __grid_constant__ ModelDev M, const __grid_constant__ FieldBwdArgs A) {
  using Cfg = BwdCfg<NP>;
  constexpr int kHalves = Cfg::kHalves, kComputeThreads = Cfg::kComputeThreads, kBwdThreads = Cfg::kThreads;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwdTcSmem<NP>* sm = reinterpret_cast<BwdTcSmem<NP>*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int tid = threadIdx.x, lane = tid & 31, wq = (tid >> 5) & 3;
  // roles: threads [0, kComputeThreads) = compute halves of 128 (thread t = sample row t of the half's tile);
  //        the rest = scatter warpgroups (thread t = sample t of whichever half's parked tile is being drained)
  const bool is_scatter = tid >= kComputeThreads;
  const int half = is_scatter ? 0 : tid >> 7;
  const int t = tid & 127;
  BwdTcCommon<NP>* C = &sm->c;
  BwdTcHalf<NP>* H = &sm->h[half];

  // ---- one-time setup ------------------------------------------------------------------------------
  tc::stage_weight_np<64, 32, NP>(C->b1, A.P.base_w0, 64, 32, 32, tid, kBwdThreads);
  tc::stage_weight_np<16, 64, NP>(C->b2, A.P.base_w1, 16, 64, 64, tid, kBwdThreads);
  tc::stage_weight_np<64, 32, NP>(C->b3, A.P.head_w0, 64, kHeadIn, kHeadK, tid, kBwdThreads);
  tc::stage_weight_np<64, 64, NP>(C->b4, A.P.head_w1, 64, 64, 64, tid, kBwdThreads);
  tc::stage_weight_np<16, 64, NP>(C->b5, A.P.head_w2, 3, 64, 64, tid, kBwdThreads);
  for (int i = tid; i < 8 * 128; i += kBwdThreads) reinterpret_cast<uint16_t*>(C->ones)[i] = NP == 1 ? 0x3C00u : 0x3F80u;
  for (int i = tid; i < kApp; i += kBwdThreads) {
    C->app[i] = appearance_component(A.P.appearance, M.num_images, A.app_row, i);
  }
  if (!is_scatter && t == 0) {
    tc::mbar_init(&H->mbar, kIssuers);
    tc::mbar_init(&H->ld_bar, 1);
    tc::mbar_init(&H->full_bar, 128);
    tc::mbar_init(&H->empty_bar, kScatterThreads);
    tc::fence_mbar_init();
  }
  if (tid < 32) tc::tmem_alloc<512>(&C->tmem_slot);
  for (int i = tid; i < 64; i += kBwdThreads) C->bias_h1[i] = A.P.head_b1[i];
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  for (int n = tid; n < 64; n += kBwdThreads) {          // as k_field_forward_tc: b0 + W0[:, 31:63] . appearance, same order
    float b = A.P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + k], C->app[k], b);
    C->bias_h0[n] = b;
  }
  __syncthreads();
  const uint32_t tm = C->tmem_slot + (uint32_t)half * kHalfCols;   // this half's columns, lane 0
  const uint32_t taddr = tm + ((uint32_t)(wq * 32) << 16);         // this warp's 32 TMEM lanes
  uint32_t phase = 0;

  // gradient operand scale: power of two that maps the launch's largest seed to 16 (fp16 operands; bf16 pieces have
  // fp32's range and need none)
  float S = 1.f, invS = 1.f;
  if (NP == 1) {
    const float gmax = A.g_absmax ? *A.g_absmax : 0.f;
    if (gmax > 0.f && gmax < 3.0e38f) {
      int e;
      frexpf(gmax, &e);                       // gmax = m * 2^e, m in [0.5, 1)
      e = 4 - e;
      e = e < -60 ? -60 : (e > 60 ? 60 : e);
      S = ldexpf(1.f, e);
      invS = ldexpf(1.f, -e);
    }
  }
  const tc::Operand op_ones{tc::smem_u32(C->ones), 128u, 2048u, 256u};
  // the same ones as an A operand with M = 64: every 8-row group re-reads the one group that exists (SBO = 0)
  const tc::Operand op_ones_a{tc::smem_u32(C->ones), 128u, 0u, 256u};

  const int64_t n_samples = A.n * kS2;
  const int64_t n_tiles = (n_samples + kTileM - 1) / kTileM;
  const int64_t tile_stride = (int64_t)gridDim.x * kHalves;
  constexpr uint32_t kTileLoadBytes = NP * (kEncTileBytes + kActTileBytes);
  const float gscale = invS * A.loss_scale;

  if (is_scatter) {
    // ---- scatter warps: drain the half's parked tiles into the gradient table ------------------------------------
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(Cfg::kScatterRegs));
    uint32_t nred = 0;               // reduction lane-ops this thread issued (reported when profiling)
    uint32_t j = 0;
    // half 0 owns tiles 2b, 2b + stride, ...; half 1 owns 2b + 1, ...: drained alternately, so that while one half's tile
    // is being scattered the other half parks its next one
    for (int64_t tile = (int64_t)blockIdx.x * kHalves; tile < n_tiles; tile += tile_stride, ++j) {
#pragma unroll 1
      for (int hh = 0; hh < kHalves; ++hh) {
        if (tile + hh >= n_tiles) break;
        BwdTcHalf<NP>* Hh = &sm->h[hh];
        tc::mbar_wait(&Hh->full_bar, j & 1u);
        if (A.G.main_table)            // null: ray-gradient-only pass (nemit_point_lights), nothing to scatter
          scatter_tile(Hh, t, lane, (tid - kComputeThreads) >> 7, kScatterGroups, M, A.G.main_table, gscale, nred);
        mbar_arrive(&Hh->empty_bar);   // release: the reads of denc / pbuf above are done
      }
    }
    if (A.red_ops) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) nred += __shfl_xor_sync(0xffffffffu, nred, o);
      if (lane == 0) atomicAdd(A.red_ops, (unsigned long long)nred);
    }
  } else {
  // ---- compute halves -----------------------------------------------------------------------------------------------
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(Cfg::kComputeRegs));
  bool first = true;
  uint32_t ld_phase = 0;
  uint32_t slot_phase = 1;           // empty_bar parity to wait for: a fresh barrier passes parity 1 (slot free)
  // thread 0 of the half: one bulk copy per operand buffer of tile `tl`
  const uint64_t pol = tc::l2_evict_first_policy();
  // (the pieces of a buffer are adjacent in the kept tile exactly as in shared memory: one copy per buffer)
  auto load_e = [&](int64_t tl) { tc::bulk_g2s(H->e, A.enc_tiles + tl * (NP * kEncTileBytes), NP * 8192, &H->ld_bar, pol); };
  auto load_h1 = [&](int64_t tl) { tc::bulk_g2s(H->h1, A.act_tiles + tl * (NP * kActTileBytes), NP * 16384, &H->ld_bar, pol); };
  auto load_in = [&](int64_t tl) { tc::bulk_g2s(H->in, A.act_tiles + tl * (NP * kActTileBytes) + NP * 16384, NP * 8192, &H->ld_bar, pol); };
  {
    const int64_t tile0 = (int64_t)blockIdx.x * kHalves + half;
    if (t == 0 && tile0 < n_tiles) {
      tc::mbar_expect_tx(&H->ld_bar, kTileLoadBytes);
      load_in(tile0); load_h1(tile0); load_e(tile0);
    }
  }
  for (int64_t tile = (int64_t)blockIdx.x * kHalves + half; tile < n_tiles; tile += tile_stride) {
    const int64_t next_tile = tile + tile_stride;
    const bool has_next = next_tile < n_tiles;
    const int64_t s = tile * kTileM + t;
    const bool valid = s < n_samples;
    const int64_t rr = valid ? s / kS2 : 0;
    const int si = valid ? (int)(s % kS2) : 0;
    float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f}, t0 = 0.f, t1 = 0.f;
    float4 seed = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + rr, ray_scale(M, A), o, d);
      t0 = A.t_edges2[rr * (kS2 + 1) + si];
      t1 = A.t_edges2[rr * (kS2 + 1) + si + 1];
      seed = __ldcs(A.seeds4 + s);
    }
    uint4 mk = make_uint4(0u, 0u, 0u, 0u);
    if (NP > 1) mk = __ldcs(A.masks4 + tile * kTileM + t);      // ReLU masks of H2 | H3 as the seed pass decided them
    const float tsum = add_rn(t0, t1);
    float x[3], p[3], dirq[3];
    // Rotater.apply_frustums where the reverse pass sees rotated final samples (the light-point-cloud pass only, see backward.cu)
    const bool rotated = rot_sample(A.rot, o, d, tsum, x, dirq);
    const bool sel = contract_normalise(x, M, p);

    // d loss / d (rgb pre-activation): 3 of 16 columns
    {
      const float g8[8] = {seed.x * S, seed.y * S, seed.z * S, 0.f, 0.f, 0.f, 0.f, 0.f};
      const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      tc::store_chunk<NP, 128, true>(H->gs, kPcGs, t, 0, g8);
      tc::store_chunk<NP, 128>(H->gs, kPcGs, t, 1, z8);
    }
    // this tile's operand buffers have landed
    tc::mbar_wait(&H->ld_bar, ld_phase);
    ld_phase ^= 1u;

    // ---- head hidden layers, forward again (H2, H3 are not kept) ------------------------------------------
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->in, 128), kPcIn, tc::op_kmajor(C->b3, 64), kPcB3, tc::idesc_f16(128, 64), 2, false);
    });
    epi_fwd_hidden<NP>(taddr + kColD0, C->bias_h0, H->h2, t, make_uint2(mk.x, mk.y));
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h2, 128), kPcH, tc::op_kmajor(C->b4, 64), kPcB4, tc::idesc_f16(128, 64), 4, false);
    });
    epi_fwd_hidden<NP>(taddr + kColD0, C->bias_h1, H->h3, t, make_uint2(mk.z, mk.w));

    // ---- head MLP backward ---------------------------------------------------------------------------
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 1) tc::mma_chain_np<NP, kRollW>(tm + kColH2T, tc::op_mnmajor(H->h3, 128), kPcH, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
      if (c == 2) tc::mma_chain_np<NP, kRollB>(tm + kColBH2, op_ones_a, 0, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->gs, 128), kPcGs, tc::op_mnmajor(C->b5, 16), kPcB5, tc::idesc_f16_mj(128, 64, 0, 1), 1, false);
    });
    epi_bwd_hidden<NP, (NP > 1)>(taddr + kColD0, H->h3, t, make_uint2(mk.z, mk.w));
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 1) tc::mma_chain_np<NP, kRollW>(tm + kColH1, tc::op_mnmajor(H->h3, 128), kPcH, tc::op_mnmajor(H->h2, 128), kPcH, tc::idesc_f16_mj(64, 64, 1, 1), 8, !first);
      if (c == 2) tc::mma_chain_np<NP, kRollB>(tm + kColBH1, tc::op_mnmajor(H->h3, 128), kPcH, op_ones, 0, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h3, 128), kPcH, tc::op_mnmajor(C->b4, 64), kPcB4, tc::idesc_f16_mj(128, 64, 0, 1), 4, false);
    });
    epi_bwd_hidden<NP, (NP > 1)>(taddr + kColD0, H->h2, t, make_uint2(mk.x, mk.y));
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 1) tc::mma_chain_np<NP, kRollW>(tm + kColH0, tc::op_mnmajor(H->h2, 128), kPcH, tc::op_mnmajor(H->in, 128), kPcIn, tc::idesc_f16_mj(64, 32, 1, 1), 8, !first);
      // geo-feature columns 16..31 only; SH carries no reverse-mode gradient (encodings.py:708)
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h2, 128), kPcH, tc::op_mnmajor(C->b3, 64, 16), kPcB3, tc::idesc_f16_mj(128, 16, 0, 1), 4, false);
    });
    if (t == 0 && has_next) {       // IN has no reader left: the next tile's copies start here
      tc::mbar_expect_tx(&H->ld_bar, kTileLoadBytes);
      load_in(next_tile);
    }
    {
      float v[16];
      tc::tmem_ld16(taddr + kColD0, v);
      tc::tmem_ld_wait();
      float g16[16];
      g16[0] = seed.w * S;
#pragma unroll
      for (int k = 0; k < 15; ++k) g16[k + 1] = v[k];
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float x8[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) x8[q] = g16[hh * 8 + q];
        tc::store_chunk<NP, 128, true>(H->gs, kPcGs, t, hh, x8);
      }
    }

    // ---- base MLP backward ---------------------------------------------------------------------------
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 1) tc::mma_chain_np<NP, kRollW>(tm + kColB1T, tc::op_mnmajor(H->h1, 128), kPcH, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 16, 1, 1), 8, !first);
      if (c == 2) tc::mma_chain_np<NP, kRollB>(tm + kColBB1, op_ones_a, 0, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 16, 1, 1), 8, !first);
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->gs, 128), kPcGs, tc::op_mnmajor(C->b2, 16), kPcB2, tc::idesc_f16_mj(128, 64, 0, 1), 1, false);
    });
    epi_bwd_hidden<NP>(taddr + kColD0, H->h1, t);
    // this tile's Jacobian rows: the loads fly under the last commit group
    uint4 jn[12 * NP];               // NP = 1: fp16 pairs; NP = 2: fp32
    {
      const uint4* jp = reinterpret_cast<const uint4*>(A.jac_tiles + tile * (NP * kJacTileBytes)) + t;
#pragma unroll
      for (int c = 0; c < 12 * NP; ++c) jn[c] = __ldcs(jp + c * 128);
    }
    mma_step(H, half, t, phase, [&](int c) {
      if (c == 1) tc::mma_chain_np<NP, kRollW>(tm + kColB0, tc::op_mnmajor(H->h1, 128), kPcH, tc::op_mnmajor(H->e, 128), kPcE, tc::idesc_f16_mj(64, 32, 1, 1), 8, !first);
      if (c == 2) tc::mma_chain_np<NP, kRollB>(tm + kColBB0, tc::op_mnmajor(H->h1, 128), kPcH, op_ones, 0, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
      if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h1, 128), kPcH, tc::op_mnmajor(C->b1, 64), kPcB1, tc::idesc_f16_mj(128, 32, 0, 1), 4, false);
    });
    if (t == 0 && has_next) { load_h1(next_tile); load_e(next_tile); }

    // ---- hash grid backward: d/d(position) here; the table scatter of this tile is handed to the scatter warps: the rows
    // are parked in shared memory once they have given the slot back (previous tile drained) --------------------------
    {
      float gp[3] = {0.f, 0.f, 0.f};
      tc::mbar_wait(&H->empty_bar, slot_phase);
      slot_phase ^= 1u;
      H->pbuf[t] = p[0]; H->pbuf[128 + t] = p[1]; H->pbuf[256 + t] = p[2];
#pragma unroll
      for (int l0 = 0; l0 < kMainLevels; l0 += 4) {
        float g[8];
        tc::tmem_ld8(taddr + kColD0 + 2 * l0, g);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 8; ++k) H->denc[(2 * l0 + k) * 128 + t] = valid ? g[k] : 0.f;
        uint32_t wj[12 * NP];
#pragma unroll
        for (int c = 0; c < 3 * NP; ++c) {
          const uint4 q = jn[(l0 >> 2) * 3 * NP + c];
          wj[4 * c] = q.x; wj[4 * c + 1] = q.y; wj[4 * c + 2] = q.z; wj[4 * c + 3] = q.w;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gx = g[2 * j], gy = g[2 * j + 1];
          float2 dx, dy, dz;
          if (NP == 1) {
            dx = __half22float2(*reinterpret_cast<const __half2*>(&wj[3 * j + 0]));
            dy = __half22float2(*reinterpret_cast<const __half2*>(&wj[3 * j + 1]));
            dz = __half22float2(*reinterpret_cast<const __half2*>(&wj[3 * j + 2]));
          } else {
            dx = make_float2(__uint_as_float(wj[6 * j + 0]), __uint_as_float(wj[6 * j + 1]));
            dy = make_float2(__uint_as_float(wj[6 * j + 2]), __uint_as_float(wj[6 * j + 3]));
            dz = make_float2(__uint_as_float(wj[6 * j + 4]), __uint_as_float(wj[6 * j + 5]));
          }
          const float scs = M.main_grid.scal[l0 + j] * invS;
          gp[0] = fmaf(scs, gx * dx.x + gy * dx.y, gp[0]);
          gp[1] = fmaf(scs, gx * dy.x + gy * dy.y, gp[1]);
          gp[2] = fmaf(scs, gx * dz.x + gy * dz.y, gp[2]);
        }
      }
      // per-ray sums d/d(origin), d/d(direction): reduced over the samples of a ray inside the warp first (a warp's 32
      // consecutive samples belong to at most two rays), then one lane per ray issues the six atomics
      float v6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (valid && sel) {
        float gx[3];
        contract_vjp(x, sel, gp, gx);
        if (rotated) {
          float gr[3];
          rot_transpose(A.rot, gx, gr);
          gx[0] = gr[0]; gx[1] = gr[1]; gx[2] = gr[2];
        }
        const float tmid = tsum * 0.5f;
#pragma unroll
        for (int k = 0; k < 3; ++k) { v6[k] = gx[k]; v6[3 + k] = gx[k] * tmid; }
      }
      {
        const long long r_any = s / kS2;                  // also defined for rows past the end (their values are zero)
        const long long r_first = __shfl_sync(0xffffffffu, r_any, 0), r_last = __shfl_sync(0xffffffffu, r_any, 31);
        const bool lo = r_any == r_first;
        float a6[6], b6[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) { a6[k] = lo ? v6[k] : 0.f; b6[k] = lo ? 0.f : v6[k]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            a6[k] += __shfl_xor_sync(0xffffffffu, a6[k], o);
            b6[k] += __shfl_xor_sync(0xffffffffu, b6[k], o);
          }
        }
        if (lane == 0 && r_first < A.n) {
#pragma unroll
          for (int k = 0; k < 6; ++k) atomicAdd(A.ray_grad + r_first * 6 + k, a6[k]);
        }
        if (lane == 31 && r_last != r_first && r_last < A.n) {
#pragma unroll
          for (int k = 0; k < 6; ++k) atomicAdd(A.ray_grad + r_last * 6 + k, b6[k]);
        }
      }
      mbar_arrive(&H->full_bar);     // release: this thread's column of denc / pbuf is written
    }
    first = false;
  }

  // ---- flush this half's weight-gradient accumulators (torch layouts, loss scale applied) ---------------
  if (!first && A.G.main_table) {
    tc::tc_fence_before();
    half_sync(half);
    tc::tc_fence_after();
    const bool has = lane < 16;             // M = 64 accumulators: row r lives in lane (r / 16) * 32 + r % 16
    const int r = wq * 16 + (lane & 15);
    const float fs = A.loss_scale * invS;
    float v[16];
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      tc::tmem_ld16(taddr + kColH1 + c * 16, v);
      tc::tmem_ld_wait();
      if (has)
        for (int i = 0; i < 16; ++i) atomicAdd(A.G.head_w1 + r * kHidden + c * 16 + i, fs * v[i]);
    }
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      tc::tmem_ld16(taddr + kColH0 + c * 16, v);
      tc::tmem_ld_wait();
      if (has) {
        for (int i = 0; i < 16; ++i) {
          const int k = c * 16 + i;
          if (k < kHeadK) {
            atomicAdd(A.G.head_w0 + r * kHeadIn + k, fs * v[i]);
          } else {
            // ones column = d head_b0_eff; the appearance columns are rank-1 with it (same 32-vector every sample)
            const float db = fs * v[i];
            H->dvec[r] = db;
            atomicAdd(A.G.head_b0 + r, db);
            for (int j = 0; j < kApp; ++j) atomicAdd(A.G.head_w0 + r * kHeadIn + kHeadK + j, db * C->app[j]);
          }
        }
      }
    }
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      tc::tmem_ld16(taddr + kColB0 + c * 16, v);
      tc::tmem_ld_wait();
      if (has)
        for (int i = 0; i < 16; ++i) atomicAdd(A.G.base_w0 + r * kMainEnc + c * 16 + i, fs * v[i]);
    }
    tc::tmem_ld16(taddr + kColB1T, v);
    tc::tmem_ld_wait();
    if (has)
      for (int n = 0; n < 16; ++n) atomicAdd(A.G.base_w1 + n * kHidden + r, fs * v[n]);
    {
      float u[8];
      tc::tmem_ld8(taddr + kColH2T, u);
      tc::tmem_ld_wait();
      if (has)
        for (int n = 0; n < 3; ++n) atomicAdd(A.G.head_w2 + n * kHidden + r, fs * u[n]);
      tc::tmem_ld8(taddr + kColBH1, u);
      tc::tmem_ld_wait();
      if (has) atomicAdd(A.G.head_b1 + r, fs * u[0]);
      tc::tmem_ld8(taddr + kColBB0, u);
      tc::tmem_ld_wait();
      if (has) atomicAdd(A.G.base_b0 + r, fs * u[0]);
    }
    tc::tmem_ld16(taddr + kColBB1, v);
    tc::tmem_ld_wait();
    if (t == 0)
      for (int n = 0; n < 16; ++n) atomicAdd(A.G.base_b1 + n, fs * v[n]);
    {
      float u[8];
      tc::tmem_ld8(taddr + kColBH2, u);
      tc::tmem_ld_wait();
      if (t == 0)
        for (int n = 0; n < 3; ++n) atomicAdd(A.G.head_b2 + n, fs * u[n]);
    }
    half_sync(half);
    // appearance embedding (NS/fields/nerfacto_field.py:272-285): row / mean / zeros as resolved by the host (appearance_component)
    if (t < kApp) {
      float g = 0.f;
      for (int n = 0; n < kHidden; ++n) g = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + t], H->dvec[n], g);
      appearance_component_grad(A.G.appearance, M.num_images, A.app_row, t, g);
    }
  }
  }  // compute halves
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_dealloc<512>(C->tmem_slot);
}

// ---------------------------------------------------------------------------------------------------------------------
// Split-operand backward, pipelined (NP = 2).  The doubled operand buffers leave room for ONE tile in flight per SM, so
// instead of two halves that each run the whole chain, the chain of a tile is cut in two and run by two warpgroups as a
// pipeline:
//   HEAD warpgroup (threads   0..127): H2 / H3 recompute, head layers 2, 1, 0 backward  ->  d [raw | geo15] rows (gs2)
//   BASE warpgroup (threads 128..255): base layers 1, 0 backward, d/d(position), per-ray sums, parks d enc for the scatter
//   scatter warpgroups (threads 256..511) as before.
// While BASE finishes tile i, HEAD is already on tile i + 1.  The two stages own disjoint operand buffers (HEAD: IN, H2, H3,
// gs; BASE: H1, E, gs2), disjoint accumulators (BASE's work accumulator sits in columns 256..319) and their own barriers;
// the hand-over is the 8 KB gs2 buffer behind an mbarrier pair.
constexpr uint32_t kColD0B = 256;    // BASE stage work accumulator (128 x 64)
struct BwdPipeSmem {
  BwdTcCommon<2> c;
  BwdTcHalf<2> h;                           // gs = HEAD's gradient operand, mbar / ld_bar = HEAD's
  alignas(128) uint8_t gs2[2 * kPcGs];      // d [raw | geo15] rows handed from HEAD to BASE
  alignas(8) uint64_t mbar_b;               // BASE's MMA commit groups
  alignas(8) uint64_t ld_bar_b;             // BASE's bulk copies (H1, E)
  alignas(8) uint64_t gs_full;              // 128 HEAD threads arrive: gs2 holds the rows of a tile
  alignas(8) uint64_t gs_empty;             // 128 BASE threads arrive: gs2 has been consumed
};
constexpr int kPipeScatterThreads = 256;         // two scatter warpgroups
constexpr int kPipeThreads = 256 + kPipeScatterThreads;
constexpr int kPipeLaunchRegs = 128, kPipeComputeRegs = 192, kPipeScatterRegs = 64;
static_assert(256 * kPipeComputeRegs + kPipeScatterThreads * kPipeScatterRegs <= kPipeThreads * kPipeLaunchRegs, "register pool of the CTA");

__global__ void __launch_bounds__(kPipeThreads, 1)
k_field_backward_tc_pipe(const __grid_constant__ ModelDev M, const __grid_constant__ FieldBwdArgs A) {
  constexpr int NP = 2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwdPipeSmem* sm = reinterpret_cast<BwdPipeSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int tid = threadIdx.x, lane = tid & 31, wq = (tid >> 5) & 3;
  const int role = tid >> 7;               // 0 = HEAD, 1 = BASE, 2 / 3 = scatter groups
  const int t = tid & 127;
  BwdTcCommon<NP>* C = &sm->c;
  BwdTcHalf<NP>* H = &sm->h;

  // ---- one-time setup ------------------------------------------------------------------------------
  tc::stage_weight_np<64, 32, NP>(C->b1, A.P.base_w0, 64, 32, 32, tid, kPipeThreads);
  tc::stage_weight_np<16, 64, NP>(C->b2, A.P.base_w1, 16, 64, 64, tid, kPipeThreads);
  tc::stage_weight_np<64, 32, NP>(C->b3, A.P.head_w0, 64, kHeadIn, kHeadK, tid, kPipeThreads);
  tc::stage_weight_np<64, 64, NP>(C->b4, A.P.head_w1, 64, 64, 64, tid, kPipeThreads);
  tc::stage_weight_np<16, 64, NP>(C->b5, A.P.head_w2, 3, 64, 64, tid, kPipeThreads);
  for (int i = tid; i < 8 * 128; i += kPipeThreads) reinterpret_cast<uint16_t*>(C->ones)[i] = 0x3F80u;      // bf16 1.0
  for (int i = tid; i < kApp; i += kPipeThreads) C->app[i] = appearance_component(A.P.appearance, M.num_images, A.app_row, i);
  if (tid == 0) {
    tc::mbar_init(&H->mbar, kIssuers);
    tc::mbar_init(&sm->mbar_b, kIssuers);
    tc::mbar_init(&H->ld_bar, 1);
    tc::mbar_init(&sm->ld_bar_b, 1);
    tc::mbar_init(&H->full_bar, 128);
    tc::mbar_init(&H->empty_bar, kPipeScatterThreads);
    tc::mbar_init(&sm->gs_full, 128);
    tc::mbar_init(&sm->gs_empty, 128);
    tc::fence_mbar_init();
  }
  if (tid < 32) tc::tmem_alloc<512>(&C->tmem_slot);
  for (int i = tid; i < 64; i += kPipeThreads) C->bias_h1[i] = A.P.head_b1[i];
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  for (int n = tid; n < 64; n += kPipeThreads) {         // as k_field_forward_tc: b0 + W0[:, 31:63] . appearance, same order
    float b = A.P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + k], C->app[k], b);
    C->bias_h0[n] = b;
  }
  __syncthreads();
  const uint32_t tm = C->tmem_slot;                                // lane 0
  const uint32_t taddr = tm + ((uint32_t)(wq * 32) << 16);         // this warp's 32 TMEM lanes
  uint32_t phase = 0;
  const tc::Operand op_ones{tc::smem_u32(C->ones), 128u, 2048u, 256u};
  const tc::Operand op_ones_a{tc::smem_u32(C->ones), 128u, 0u, 256u};
  const int64_t n_samples = A.n * kS2;
  const int64_t n_tiles = (n_samples + kTileM - 1) / kTileM;
  const int64_t tile_stride = gridDim.x;
  const float gscale = A.loss_scale;
  const uint64_t pol = tc::l2_evict_first_policy();
  bool first = true;

  if (role >= 2) {
    // ---- scatter warps ---------------------------------------------------------------------------------------------
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kPipeScatterRegs));
    uint32_t nred = 0, j = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += tile_stride, ++j) {
      tc::mbar_wait(&H->full_bar, j & 1u);
      if (A.G.main_table) scatter_tile(H, t, lane, role - 2, 2, M, A.G.main_table, gscale, nred);
      mbar_arrive(&H->empty_bar);
    }
    if (A.red_ops) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) nred += __shfl_xor_sync(0xffffffffu, nred, o);
      if (lane == 0) atomicAdd(A.red_ops, (unsigned long long)nred);
    }
  } else if (role == 0) {
    // ---- HEAD warpgroup ----------------------------------------------------------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kPipeComputeRegs));
    uint32_t ld_phase = 0, gs_phase = 1;      // gs_empty parity to wait for: a fresh barrier passes parity 1 (gs2 free)
    auto load_in = [&](int64_t tl) { tc::bulk_g2s(H->in, A.act_tiles + tl * (NP * kActTileBytes) + NP * 16384, NP * 8192, &H->ld_bar, pol); };
    if (t == 0 && (int64_t)blockIdx.x < n_tiles) {
      tc::mbar_expect_tx(&H->ld_bar, NP * 8192);
      load_in(blockIdx.x);
    }
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += tile_stride) {
      const int64_t next_tile = tile + tile_stride;
      const int64_t s = tile * kTileM + t;
      const bool valid = s < n_samples;
      float4 seed = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) seed = __ldcs(A.seeds4 + s);
      const uint4 mk = __ldcs(A.masks4 + tile * kTileM + t);      // ReLU masks of H2 | H3 as the seed pass decided them
      {
        const float g8[8] = {seed.x, seed.y, seed.z, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        tc::store_chunk<NP, 128, true>(H->gs, kPcGs, t, 0, g8);
        tc::store_chunk<NP, 128>(H->gs, kPcGs, t, 1, z8);
      }
      tc::mbar_wait(&H->ld_bar, ld_phase);
      ld_phase ^= 1u;
      // head hidden layers, forward again (H2, H3 are not kept)
      mma_step_on(&H->mbar, 0, t, phase, [&](int c) {
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->in, 128), kPcIn, tc::op_kmajor(C->b3, 64), kPcB3, tc::idesc_f16(128, 64), 2, false);
      });
      epi_fwd_hidden<NP>(taddr + kColD0, C->bias_h0, H->h2, t, make_uint2(mk.x, mk.y));
      mma_step_on(&H->mbar, 0, t, phase, [&](int c) {
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h2, 128), kPcH, tc::op_kmajor(C->b4, 64), kPcB4, tc::idesc_f16(128, 64), 4, false);
      });
      epi_fwd_hidden<NP>(taddr + kColD0, C->bias_h1, H->h3, t, make_uint2(mk.z, mk.w));
      // head MLP backward
      mma_step_on(&H->mbar, 0, t, phase, [&](int c) {
        if (c == 1) tc::mma_chain_np<NP>(tm + kColH2T, tc::op_mnmajor(H->h3, 128), kPcH, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
        if (c == 2) tc::mma_chain_np<NP>(tm + kColBH2, op_ones_a, 0, tc::op_mnmajor(H->gs, 128), kPcGs, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->gs, 128), kPcGs, tc::op_mnmajor(C->b5, 16), kPcB5, tc::idesc_f16_mj(128, 64, 0, 1), 1, false);
      });
      epi_bwd_hidden<NP, true>(taddr + kColD0, H->h3, t, make_uint2(mk.z, mk.w));
      mma_step_on(&H->mbar, 0, t, phase, [&](int c) {
        if (c == 1) tc::mma_chain_np<NP>(tm + kColH1, tc::op_mnmajor(H->h3, 128), kPcH, tc::op_mnmajor(H->h2, 128), kPcH, tc::idesc_f16_mj(64, 64, 1, 1), 8, !first);
        if (c == 2) tc::mma_chain_np<NP>(tm + kColBH1, tc::op_mnmajor(H->h3, 128), kPcH, op_ones, 0, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h3, 128), kPcH, tc::op_mnmajor(C->b4, 64), kPcB4, tc::idesc_f16_mj(128, 64, 0, 1), 4, false);
      });
      epi_bwd_hidden<NP, true>(taddr + kColD0, H->h2, t, make_uint2(mk.x, mk.y));
      mma_step_on(&H->mbar, 0, t, phase, [&](int c) {
        if (c == 1) tc::mma_chain_np<NP>(tm + kColH0, tc::op_mnmajor(H->h2, 128), kPcH, tc::op_mnmajor(H->in, 128), kPcIn, tc::idesc_f16_mj(64, 32, 1, 1), 8, !first);
        // geo-feature columns 16..31 only; SH carries no reverse-mode gradient (encodings.py:708)
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0, tc::op_kmajor(H->h2, 128), kPcH, tc::op_mnmajor(C->b3, 64, 16), kPcB3, tc::idesc_f16_mj(128, 16, 0, 1), 4, false);
      });
      if (t == 0 && next_tile < n_tiles) {        // IN has no reader left: the next tile's copy starts here
        tc::mbar_expect_tx(&H->ld_bar, NP * 8192);
        load_in(next_tile);
      }
      {
        float v[16];
        tc::tmem_ld16(taddr + kColD0, v);
        tc::tmem_ld_wait();
        float g16[16];
        g16[0] = seed.w;
#pragma unroll
        for (int k = 0; k < 15; ++k) g16[k + 1] = v[k];
        tc::mbar_wait(&sm->gs_empty, gs_phase);      // BASE has consumed the previous tile's rows
        gs_phase ^= 1u;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float x8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) x8[q] = g16[hh * 8 + q];
          tc::store_chunk<NP, 128, true>(sm->gs2, kPcGs, t, hh, x8);
        }
        tc::fence_proxy_async();                     // the reader is BASE's tensor-core work
        mbar_arrive(&sm->gs_full);
      }
      first = false;
    }
    // flush HEAD's weight-gradient accumulators
    if (!first && A.G.main_table) {
      tc::tc_fence_before();
      half_sync(0);
      tc::tc_fence_after();
      const bool has = lane < 16;             // M = 64 accumulators: row r lives in lane (r / 16) * 32 + r % 16
      const int r = wq * 16 + (lane & 15);
      const float fs = A.loss_scale;
      float v[16];
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        tc::tmem_ld16(taddr + kColH1 + c * 16, v);
        tc::tmem_ld_wait();
        if (has)
          for (int i = 0; i < 16; ++i) atomicAdd(A.G.head_w1 + r * kHidden + c * 16 + i, fs * v[i]);
      }
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        tc::tmem_ld16(taddr + kColH0 + c * 16, v);
        tc::tmem_ld_wait();
        if (has) {
          for (int i = 0; i < 16; ++i) {
            const int k = c * 16 + i;
            if (k < kHeadK) {
              atomicAdd(A.G.head_w0 + r * kHeadIn + k, fs * v[i]);
            } else {
              const float db = fs * v[i];     // ones column = d head_b0_eff; the appearance columns are rank-1 with it
              H->dvec[r] = db;
              atomicAdd(A.G.head_b0 + r, db);
              for (int j = 0; j < kApp; ++j) atomicAdd(A.G.head_w0 + r * kHeadIn + kHeadK + j, db * C->app[j]);
            }
          }
        }
      }
      {
        float u[8];
        tc::tmem_ld8(taddr + kColH2T, u);
        tc::tmem_ld_wait();
        if (has)
          for (int n = 0; n < 3; ++n) atomicAdd(A.G.head_w2 + n * kHidden + r, fs * u[n]);
        tc::tmem_ld8(taddr + kColBH1, u);
        tc::tmem_ld_wait();
        if (has) atomicAdd(A.G.head_b1 + r, fs * u[0]);
        tc::tmem_ld8(taddr + kColBH2, u);
        tc::tmem_ld_wait();
        if (t == 0)
          for (int n = 0; n < 3; ++n) atomicAdd(A.G.head_b2 + n, fs * u[n]);
      }
      half_sync(0);
      if (t < kApp) {                         // appearance embedding (NS/fields/nerfacto_field.py:272-285)
        float g = 0.f;
        for (int n = 0; n < kHidden; ++n) g = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + t], H->dvec[n], g);
        appearance_component_grad(A.G.appearance, M.num_images, A.app_row, t, g);
      }
    }
  } else {
    // ---- BASE warpgroup ----------------------------------------------------------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kPipeComputeRegs));
    uint32_t ld_phase = 0, gs_phase = 0, slot_phase = 1;
    auto load_e = [&](int64_t tl) { tc::bulk_g2s(H->e, A.enc_tiles + tl * (NP * kEncTileBytes), NP * 8192, &sm->ld_bar_b, pol); };
    auto load_h1 = [&](int64_t tl) { tc::bulk_g2s(H->h1, A.act_tiles + tl * (NP * kActTileBytes), NP * 16384, &sm->ld_bar_b, pol); };
    if (t == 0 && (int64_t)blockIdx.x < n_tiles) {
      tc::mbar_expect_tx(&sm->ld_bar_b, NP * (8192 + 16384));
      load_h1(blockIdx.x); load_e(blockIdx.x);
    }
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += tile_stride) {
      const int64_t next_tile = tile + tile_stride;
      const int64_t s = tile * kTileM + t;
      const bool valid = s < n_samples;
      const int64_t rr = valid ? s / kS2 : 0;
      const int si = valid ? (int)(s % kS2) : 0;
      float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f}, t0 = 0.f, t1 = 0.f;
      if (valid) {
        load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + rr, ray_scale(M, A), o, d);
        t0 = A.t_edges2[rr * (kS2 + 1) + si];
        t1 = A.t_edges2[rr * (kS2 + 1) + si + 1];
      }
      const float tsum = add_rn(t0, t1);
      float x[3], p[3], dirq[3];
      const bool rotated = rot_sample(A.rot, o, d, tsum, x, dirq);
      const bool sel = contract_normalise(x, M, p);
      tc::mbar_wait(&sm->ld_bar_b, ld_phase);
      ld_phase ^= 1u;
      tc::mbar_wait(&sm->gs_full, gs_phase);       // HEAD's d [raw | geo15] rows of this tile
      gs_phase ^= 1u;
      mma_step_on(&sm->mbar_b, 1, t, phase, [&](int c) {
        if (c == 1) tc::mma_chain_np<NP>(tm + kColB1T, tc::op_mnmajor(H->h1, 128), kPcH, tc::op_mnmajor(sm->gs2, 128), kPcGs, tc::idesc_f16_mj(64, 16, 1, 1), 8, !first);
        if (c == 2) tc::mma_chain_np<NP>(tm + kColBB1, op_ones_a, 0, tc::op_mnmajor(sm->gs2, 128), kPcGs, tc::idesc_f16_mj(64, 16, 1, 1), 8, !first);
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0B, tc::op_kmajor(sm->gs2, 128), kPcGs, tc::op_mnmajor(C->b2, 16), kPcB2, tc::idesc_f16_mj(128, 64, 0, 1), 1, false);
      });
      mbar_arrive(&sm->gs_empty);                  // every reader of gs2 (the commit group above) has completed
      epi_bwd_hidden<NP>(taddr + kColD0B, H->h1, t);
      // this tile's Jacobian rows (fp32): the loads fly under the last commit group
      uint4 jn[12 * NP];
      {
        const uint4* jp = reinterpret_cast<const uint4*>(A.jac_tiles + tile * (NP * kJacTileBytes)) + t;
#pragma unroll
        for (int c = 0; c < 12 * NP; ++c) jn[c] = __ldcs(jp + c * 128);
      }
      mma_step_on(&sm->mbar_b, 1, t, phase, [&](int c) {
        if (c == 1) tc::mma_chain_np<NP>(tm + kColB0, tc::op_mnmajor(H->h1, 128), kPcH, tc::op_mnmajor(H->e, 128), kPcE, tc::idesc_f16_mj(64, 32, 1, 1), 8, !first);
        if (c == 2) tc::mma_chain_np<NP>(tm + kColBB0, tc::op_mnmajor(H->h1, 128), kPcH, op_ones, 0, tc::idesc_f16_mj(64, 8, 1, 1), 8, !first);
        if (c == 0) tc::mma_chain_np<NP>(tm + kColD0B, tc::op_kmajor(H->h1, 128), kPcH, tc::op_mnmajor(C->b1, 64), kPcB1, tc::idesc_f16_mj(128, 32, 0, 1), 4, false);
      });
      if (t == 0 && next_tile < n_tiles) {
        tc::mbar_expect_tx(&sm->ld_bar_b, NP * (8192 + 16384));
        load_h1(next_tile); load_e(next_tile);
      }
      // hash grid backward: d/d(position) here, the table scatter of this tile is handed to the scatter warps
      {
        float gp[3] = {0.f, 0.f, 0.f};
        tc::mbar_wait(&H->empty_bar, slot_phase);
        slot_phase ^= 1u;
        H->pbuf[t] = p[0]; H->pbuf[128 + t] = p[1]; H->pbuf[256 + t] = p[2];
#pragma unroll
        for (int l0 = 0; l0 < kMainLevels; l0 += 4) {
          float g[8];
          tc::tmem_ld8(taddr + kColD0B + 2 * l0, g);
          tc::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 8; ++k) H->denc[(2 * l0 + k) * 128 + t] = valid ? g[k] : 0.f;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float gx = g[2 * j], gy = g[2 * j + 1];
            // rows of level l0 + j: dx.xy dy.xy dz.xy = six consecutive floats of the row's 24 per four levels
            const uint4 qa = jn[(l0 >> 2) * 6 + (6 * j) / 4], qb = jn[(l0 >> 2) * 6 + (6 * j) / 4 + 1];
            const uint32_t w8[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
            const int b0 = (6 * j) % 4;
            const float dxx = __uint_as_float(w8[b0]), dxy = __uint_as_float(w8[b0 + 1]), dyx = __uint_as_float(w8[b0 + 2]),
                        dyy = __uint_as_float(w8[b0 + 3]), dzx = __uint_as_float(w8[b0 + 4]), dzy = __uint_as_float(w8[b0 + 5]);
            const float scs = M.main_grid.scal[l0 + j];
            gp[0] = fmaf(scs, gx * dxx + gy * dxy, gp[0]);
            gp[1] = fmaf(scs, gx * dyx + gy * dyy, gp[1]);
            gp[2] = fmaf(scs, gx * dzx + gy * dzy, gp[2]);
          }
        }
        float v6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (valid && sel) {
          float gx[3];
          contract_vjp(x, sel, gp, gx);
          if (rotated) {
            float gr[3];
            rot_transpose(A.rot, gx, gr);
            gx[0] = gr[0]; gx[1] = gr[1]; gx[2] = gr[2];
          }
          const float tmid = tsum * 0.5f;
#pragma unroll
          for (int k = 0; k < 3; ++k) { v6[k] = gx[k]; v6[3 + k] = gx[k] * tmid; }
        }
        {
          const long long r_any = s / kS2;
          const long long r_first = __shfl_sync(0xffffffffu, r_any, 0), r_last = __shfl_sync(0xffffffffu, r_any, 31);
          const bool lo = r_any == r_first;
          float a6[6], b6[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) { a6[k] = lo ? v6[k] : 0.f; b6[k] = lo ? 0.f : v6[k]; }
#pragma unroll
          for (int of = 16; of > 0; of >>= 1) {
#pragma unroll
            for (int k = 0; k < 6; ++k) {
              a6[k] += __shfl_xor_sync(0xffffffffu, a6[k], of);
              b6[k] += __shfl_xor_sync(0xffffffffu, b6[k], of);
            }
          }
          if (lane == 0 && r_first < A.n) {
#pragma unroll
            for (int k = 0; k < 6; ++k) atomicAdd(A.ray_grad + r_first * 6 + k, a6[k]);
          }
          if (lane == 31 && r_last != r_first && r_last < A.n) {
#pragma unroll
            for (int k = 0; k < 6; ++k) atomicAdd(A.ray_grad + r_last * 6 + k, b6[k]);
          }
        }
        mbar_arrive(&H->full_bar);
      }
      first = false;
    }
    // flush BASE's weight-gradient accumulators
    if (!first && A.G.main_table) {
      tc::tc_fence_before();
      half_sync(1);
      tc::tc_fence_after();
      const bool has = lane < 16;
      const int r = wq * 16 + (lane & 15);
      const float fs = A.loss_scale;
      float v[16];
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        tc::tmem_ld16(taddr + kColB0 + c * 16, v);
        tc::tmem_ld_wait();
        if (has)
          for (int i = 0; i < 16; ++i) atomicAdd(A.G.base_w0 + r * kMainEnc + c * 16 + i, fs * v[i]);
      }
      tc::tmem_ld16(taddr + kColB1T, v);
      tc::tmem_ld_wait();
      if (has)
        for (int n = 0; n < 16; ++n) atomicAdd(A.G.base_w1 + n * kHidden + r, fs * v[n]);
      {
        float u[8];
        tc::tmem_ld8(taddr + kColBB0, u);
        tc::tmem_ld_wait();
        if (has) atomicAdd(A.G.base_b0 + r, fs * u[0]);
      }
      tc::tmem_ld16(taddr + kColBB1, v);
      tc::tmem_ld_wait();
      if (t == 0)
        for (int n = 0; n < 16; ++n) atomicAdd(A.G.base_b1 + n, fs * v[n]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_dealloc<512>(C->tmem_slot);
}

template <int NP>
static cudaError_t launch_bwd_np(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream) {
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  const size_t smem = sizeof(BwdTcSmem<NP>) + 128;
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_backward_tc<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int64_t tiles = (A.n * kS2 + kTileM - 1) / kTileM;
  const int64_t per_cta = (tiles + BwdCfg<NP>::kHalves - 1) / BwdCfg<NP>::kHalves;
  const int64_t ctas = per_cta < 148 ? per_cta : 148;  // persistent: one CTA (its compute halves, all of TMEM) per SM
  k_field_backward_tc<NP><<<(unsigned)ctas, BwdCfg<NP>::kThreads, smem, stream>>>(M, A);
  return cudaGetLastError();
}
#ifndef NEMIT_SPLIT_BWD_PIPE
#define NEMIT_SPLIT_BWD_PIPE 1         // split operands: 1 = the pipelined kernel (HEAD / BASE warpgroups), 0 = one compute half
#endif
static cudaError_t launch_bwd_pipe(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream) {
  static DeviceOnce attr_once;
  const size_t smem = sizeof(BwdPipeSmem) + 128;
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_backward_tc_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int64_t tiles = (A.n * kS2 + kTileM - 1) / kTileM;
  k_field_backward_tc_pipe<<<(unsigned)(tiles < 148 ? tiles : 148), kPipeThreads, smem, stream>>>(M, A);
  return cudaGetLastError();
}
cudaError_t launch_field_backward_tc(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream, int pieces) {
  if (A.n <= 0) return cudaSuccess;
  if (pieces == 2) return NEMIT_SPLIT_BWD_PIPE ? launch_bwd_pipe(M, A, stream) : launch_bwd_np<2>(M, A, stream);
  return launch_bwd_np<1>(M, A, stream);
}

}  // namespace nemit
