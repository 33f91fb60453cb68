// field_tc.cu - k_field_forward_tc: the main field with the 64-wide MLP chain on the 5th-generation tensor
// cores (tcgen05.mma, fp16 operands, fp32 accumulation in TMEM).
//
// One CTA = 128 threads = one 128-sample tile at a time (thread m owns sample row m: it gathers the row's hash
// encoding, and after every layer reads ITS accumulator row back from TMEM lane m, applies bias / ReLU in fp32,
// and writes the fp16 row of the next layer's A operand).  The five GEMMs of a tile
//     enc[128x32] W_b0^T -> h1[128x64] W_b1^T -> [raw | geo15] ;  [SH16 | geo15 | 0] W_h0'^T -> h2 W_h1^T -> h3 W_h2^T
// are issued by one thread each (M = 128, N = 64 / 16, K = 32 / 64), the accumulator reuses the same 64 TMEM
// columns, the weights stay resident in shared memory in the UMMA K-major layout for the CTA's lifetime.
// Several CTAs per SM overlap one CTA's gather phase with another's MMA / epilogue phase.
// Density activation, SafeExp, weights scan and composite stay in fp32 (same code as the fp32 kernel).
//
// Reference: NerfactoField.get_density / get_outputs (NS/fields/nerfacto_field.py:228-334); the reference runs
// this MLP in fp16 under autocast / tiny-cuda-nn FullyFusedMLP (NS/models/base_model.py:305,
// NS/field_components/mlp.py:117-144) - this is that "fp16 MLP" mode of BASELINE.json configs[1].
#include "kernels.h"
#include "field_common.cuh"
#include "tc_common.cuh"

namespace nemit {

struct TcSmem {
  // fp16 B operands (UMMA K-major layout, rows = output features)
  alignas(128) uint8_t b1[64 * 32 * 2];   // base layer 0   (64 x 32)
  alignas(128) uint8_t b2[16 * 64 * 2];   // base layer 1   (16 x 64)
  alignas(128) uint8_t b3[64 * 32 * 2];   // head layer 0   (64 x [SH16 | geo15 | 0])
  alignas(128) uint8_t b4[64 * 64 * 2];   // head layer 1   (64 x 64)
  alignas(128) uint8_t b5[16 * 64 * 2];   // head layer 2   (3 rows used of 16)
  alignas(128) uint8_t a[128 * 64 * 2];   // A operand of the current layer (128 x K<=64)
  float bias_b0[64], bias_b1[16], bias_h0[64], bias_h1[64], bias_h2[4];
  float sigma[kGroupSamples];
  float rgb[kGroupSamples * 3];
  float t_edge[kRaysPerGroup * (kS2 + 1)];
  float ray_o[kRaysPerGroup * 3];
  float ray_d[kRaysPerGroup * 3];
  float app_tmp[kApp];
  float s_edge[kRaysPerGroup * (kS2 + 1)];   // adjoint-seed mode only
  alignas(8) uint64_t mbar;
  uint32_t tmem_slot;
  int seed_max_bits;                         // adjoint-seed mode: max seed magnitude of this CTA (float bits)
};

using tc::stage_weight;

// accumulator row (64 columns) -> +bias, ReLU -> fp16 row of the next A operand (8 K-chunks)
__device__ __forceinline__ void epilogue_hidden(uint32_t taddr, const float* __restrict__ bias, uint8_t* a, int row,
                                                uint4* keep = nullptr) {
  float v[4][16];
#pragma unroll
  for (int c = 0; c < 4; ++c) tc::tmem_ld16(taddr + c * 16, v[c]);
  tc::tmem_ld_wait();
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        int j = hh * 8 + q * 2;
        float x0 = fmaxf(v[c][j] + bias[c * 16 + j], 0.f), x1 = fmaxf(v[c][j + 1] + bias[c * 16 + j + 1], 0.f);
        w[q] = tc::pack_h2(x0, x1);
      }
      *reinterpret_cast<uint4*>(a + tc::chunk_off<128>(row, c * 2 + hh)) = make_uint4(w[0], w[1], w[2], w[3]);
      if (keep) __stcs(keep + (c * 2 + hh) * 128 + row, make_uint4(w[0], w[1], w[2], w[3]));   // same layout, for the backward
    }
  }
}

// make this thread's A-operand writes visible to the tensor core, then let one thread issue the layer
template <int N, int K>
__device__ __forceinline__ void run_layer(TcSmem* sm, const uint8_t* b, uint32_t tmem, uint32_t& phase, int tid,
                                          uint8_t* keep = nullptr, uint32_t keep_bytes = 0, uint64_t policy = 0) {
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc::tc_fence_after();
    tc::issue_layer<N, K>(tmem, sm->a, b, &sm->mbar);
    if (keep) {
      // adjoint-seed mode: the A operand (= this layer's input rows, already in the tile layout) goes to the
      // workspace through the TMA engine; the operand buffer is reused once the engine has read it
      tc::bulk_s2g(keep, sm->a, keep_bytes, policy);
      tc::bulk_wait_read();
    }
  }
  tc::mbar_wait(&sm->mbar, phase);
  phase ^= 1u;
  tc::tc_fence_after();
  if (keep) __syncthreads();
}

// MODE 0: forward.  MODE 1: first pass of the backward (same forward, then the per-ray composite adjoint).
template <int MODE>
__global__ void __launch_bounds__(128)
k_field_forward_tc(const __grid_constant__ ModelDev M, const __grid_constant__ FieldArgs A) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcSmem* sm = reinterpret_cast<TcSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

  // ---- one-time setup: weights -> fp16 operands, biases, effective head bias, TMEM, mbarrier ------------
  stage_weight<64, 32>(sm->b1, A.P.base_w0, 64, 32, 32, tid, 128);
  stage_weight<16, 64>(sm->b2, A.P.base_w1, 16, 64, 64, tid, 128);
  stage_weight<64, 32>(sm->b3, A.P.head_w0, 64, kHeadIn, kHeadK, tid, 128);
  stage_weight<64, 64>(sm->b4, A.P.head_w1, 64, 64, 64, tid, 128);
  stage_weight<16, 64>(sm->b5, A.P.head_w2, 3, 64, 64, tid, 128);
  for (int i = tid; i < 64; i += 128) { sm->bias_b0[i] = A.P.base_b0[i]; sm->bias_h1[i] = A.P.head_b1[i]; }
  for (int i = tid; i < 16; i += 128) sm->bias_b1[i] = A.P.base_b1[i];
  for (int i = tid; i < 4; i += 128) sm->bias_h2[i] = i < 3 ? A.P.head_b2[i] : 0.f;
  for (int i = tid; i < kApp; i += 128) {
    sm->app_tmp[i] = appearance_component(A.P.appearance, M.num_images, A.app_row, i);
  }
  if (tid == 0) {
    tc::mbar_init(&sm->mbar, 1);
    tc::fence_mbar_init();
    sm->seed_max_bits = 0;
  }
  float seed_max = 0.f;
  if (wid == 0) tc::tmem_alloc<64>(&sm->tmem_slot);
  __syncthreads();
  for (int n = tid; n < 64; n += 128) {
    float b = A.P.head_b0[n];
    for (int k = 0; k < kApp; ++k) b = fmaf(A.P.head_w0[n * kHeadIn + kHeadK + k], sm->app_tmp[k], b);
    sm->bias_h0[n] = b;
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = sm->tmem_slot;
  const uint32_t taddr = tmem + ((uint32_t)(wid * 32) << 16);   // this warp's 32 TMEM lanes
  uint32_t phase = 0;
  const uint64_t policy = MODE == 1 ? tc::l2_evict_first_policy() : 0;

  const int64_t n_groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t r0 = g * kRaysPerGroup;
    const int rays_here = (int)min((int64_t)kRaysPerGroup, A.n - r0);
    __syncthreads();
    if (tid < kRaysPerGroup) {
      float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
      if (tid < rays_here) load_ray(A.o, A.v, A.near_per_ray, A.near_scalar, A.n_total, A.ray_offset + r0 + tid,
                                    ray_scale(M, A), o, d);
#pragma unroll
      for (int k = 0; k < 3; ++k) { sm->ray_o[tid * 3 + k] = o[k]; sm->ray_d[tid * 3 + k] = d[k]; }
    }
    for (int i = tid; i < kRaysPerGroup * (kS2 + 1); i += 128) {
      int rr = i / (kS2 + 1);
      sm->t_edge[i] = rr < rays_here ? A.t_edges2[r0 * (kS2 + 1) + i] : 0.f;
      if (MODE == 1) sm->s_edge[i] = rr < rays_here ? A.s_edges2[r0 * (kS2 + 1) + i] : 0.f;
    }
    __syncthreads();

    for (int tile = 0; tile < kGroupSamples / kTileM; ++tile) {
      const int s = tile * kTileM + tid;
      const int rr = s / kS2, si = s % kS2;
      const float t0 = sm->t_edge[rr * (kS2 + 1) + si], t1 = sm->t_edge[rr * (kS2 + 1) + si + 1];
      const float tsum = add_rn(t0, t1);
      float x[3], p[3], dirq[3];
      rot_sample(A.rot, &sm->ray_o[rr * 3], &sm->ray_d[rr * 3], tsum, x, dirq);      // Rotater.apply_frustums when set for this stage
      const bool sel = contract_normalise(x, M, p) && rr < rays_here;
      const int64_t tile_id = g * (kGroupSamples / kTileM) + tile;
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
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 r = blend_cell(v[j], cell[j]);
          w[j] = tc::pack_h2(r.x, r.y);
        }
        *reinterpret_cast<uint4*>(sm->a + tc::chunk_off<128>(tid, l0 >> 2)) = make_uint4(w[0], w[1], w[2], w[3]);
        if (MODE == 1 && A.enc_tiles) {
          // keep the row and its Jacobian for the backward kernel (coalesced 16-byte stores, chunk-major tiles)
          uint32_t wj[12];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float2 dx, dy, dz;
            blend_cell_grad(v[j], cell[j], dx, dy, dz);
            wj[3 * j + 0] = tc::pack_h2_sat(dx.x, dx.y);
            wj[3 * j + 1] = tc::pack_h2_sat(dy.x, dy.y);
            wj[3 * j + 2] = tc::pack_h2_sat(dz.x, dz.y);
          }
          uint4* jp = reinterpret_cast<uint4*>(A.jac_tiles + tile_id * kJacTileBytes) + (l0 >> 2) * 3 * 128 + tid;
          __stcs(jp, make_uint4(wj[0], wj[1], wj[2], wj[3]));
          __stcs(jp + 128, make_uint4(wj[4], wj[5], wj[6], wj[7]));
          __stcs(jp + 256, make_uint4(wj[8], wj[9], wj[10], wj[11]));
        }
      }
      // adjoint-seed mode keeps the activations (fp16 operand rows) for k_field_backward_tc
      uint8_t* act = nullptr;
      if (MODE == 1 && A.act_tiles) act = A.act_tiles + tile_id * kActTileBytes;
      // ---- base MLP ------------------------------------------------------------------------------------
      run_layer<64, 32>(sm, sm->b1, tmem, phase, tid, act ? A.enc_tiles + tile_id * kEncTileBytes : nullptr, 8192, policy);
      epilogue_hidden(taddr, sm->bias_b0, sm->a, tid);
      run_layer<16, 64>(sm, sm->b2, tmem, phase, tid, act, 16384, policy);
      {
        float v[16];
        tc::tmem_ld16(taddr, v);
        tc::tmem_ld_wait();
        const float raw = v[0] + sm->bias_b1[0];
        sm->sigma[s] = sel ? expf(raw) : 0.f;                         // trunc_exp * selector, fp32
        float c[16];
        sh16((dirq[0] + 1.f) * 0.5f, (dirq[1] + 1.f) * 0.5f, (dirq[2] + 1.f) * 0.5f, c);
        uint32_t w[4];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
          for (int q = 0; q < 4; ++q) w[q] = tc::pack_h2(c[hh * 8 + q * 2], c[hh * 8 + q * 2 + 1]);
          *reinterpret_cast<uint4*>(sm->a + tc::chunk_off<128>(tid, hh)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        float gfe[16];
#pragma unroll
        for (int k = 0; k < 15; ++k) gfe[k] = v[k + 1] + sm->bias_b1[k + 1];
        gfe[15] = MODE == 1 ? 1.f : 0.f;   // pad column (zero weights); the backward's wgrad reads it as a ones column
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
          for (int q = 0; q < 4; ++q) w[q] = tc::pack_h2(gfe[hh * 8 + q * 2], gfe[hh * 8 + q * 2 + 1]);
          *reinterpret_cast<uint4*>(sm->a + tc::chunk_off<128>(tid, 2 + hh)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      // ---- head MLP ------------------------------------------------------------------------------------
      run_layer<64, 32>(sm, sm->b3, tmem, phase, tid, act ? act + 16384 : nullptr, 8192, policy);
      epilogue_hidden(taddr, sm->bias_h0, sm->a, tid);
      // H2 / H3 are not kept: the backward recomputes them from IN (kernels.h)
      run_layer<64, 64>(sm, sm->b4, tmem, phase, tid);
      epilogue_hidden(taddr, sm->bias_h1, sm->a, tid);
      run_layer<16, 64>(sm, sm->b5, tmem, phase, tid);
      {
        float v[16];
        tc::tmem_ld16(taddr, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 3; ++j)
          sm->rgb[s * 3 + j] = expf(fminf(v[j] + sm->bias_h2[j] + M.rgb_bias, 88.f));   // SafeExp
      }
      // the next tile's gather overwrites A: the last MMA has completed (waited above); TMEM loads are done
      tc::tc_fence_before();
      __syncthreads();
    }

    // composite: warp w handles rays w and w+4; lane handles samples lane and lane+32 (fp32, as field.cu)
    for (int rr = wid; rr < rays_here; rr += 4) {
      if (MODE == 1) {
        seed_max = fmaxf(seed_max, composite_adjoint_ray<true>(M, A, &sm->t_edge[rr * (kS2 + 1)], &sm->s_edge[rr * (kS2 + 1)],
                                                         &sm->sigma[rr * kS2], &sm->rgb[rr * kS2 * 3], r0 + rr, lane));
        continue;
      }
      float ds[2], w[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        ds[k] = 0.f;
        if (si < kS2) {
          float t0 = sm->t_edge[rr * (kS2 + 1) + si], t1 = sm->t_edge[rr * (kS2 + 1) + si + 1];
          ds[k] = mul_rn(sub_rn(t1, t0), sm->sigma[rr * kS2 + si]);
        }
      }
      weights_from_ds<2>(ds, w, lane);
      float cr = 0.f, cg = 0.f, cb = 0.f, acc = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        int si = k * 32 + lane;
        if (si < kS2) {
          const float* c = &sm->rgb[(rr * kS2 + si) * 3];
          float r = c[0], gg = c[1], b = c[2];
          if (!A.training) { r = nan_to_num(r); gg = nan_to_num(gg); b = nan_to_num(b); }
          cr = fmaf(w[k], r, cr); cg = fmaf(w[k], gg, cg); cb = fmaf(w[k], b, cb);
          acc += w[k];
          if (A.weights2) A.weights2[(r0 + rr) * kS2 + si] = w[k];
        }
      }
      cr = warp_sum(cr); cg = warp_sum(cg); cb = warp_sum(cb); acc = warp_sum(acc);
      if (lane == 0) {
        const float* last = &sm->rgb[(rr * kS2 + kS2 - 1) * 3];
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
  if (wid == 0) tc::tmem_dealloc<64>(tmem);
}

template <int MODE>
static cudaError_t launch_tc_mode(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  if (A.n <= 0) return cudaSuccess;
  static DeviceOnce attr_once;      // cudaFuncSetAttribute is per device
  size_t smem = sizeof(TcSmem) + 128;
  if (attr_once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_field_forward_tc<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int64_t groups = (A.n + kRaysPerGroup - 1) / kRaysPerGroup;
  int64_t ctas = groups < 148 * 4 ? groups : 148 * 4;     // persistent: 4 CTAs per SM share the tensor core
  k_field_forward_tc<MODE><<<(unsigned)ctas, 128, smem, stream>>>(M, A);
  return cudaGetLastError();
}
cudaError_t launch_field_forward_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  return launch_tc_mode<0>(M, A, stream);
}
cudaError_t launch_field_adjoint_seed_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream) {
  return launch_tc_mode<1>(M, A, stream);
}

}  // namespace nemit
