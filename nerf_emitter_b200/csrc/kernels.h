// kernels.h - argument blocks and launchers of the CUDA kernels (internal; the public ABI is include/nemit.h)
#pragma once
#include "common.cuh"

namespace nemit {

// rays are addressed as (3, n_total) SoA with a chunk [ray_offset, ray_offset + n); every OUTPUT pointer
// is already offset to the chunk by the host.
struct RayArgs {
  const float* o;
  const float* v;
  const float* near_per_ray;
  float near_scalar;
  int64_t n_total;
  int64_t ray_offset;
  int64_t n;
  int training;
  int camera_idx;
  const float* jitter;   // (n_total,3) or null
  float anneal;
};

struct SampleArgs : RayArgs {
  const float2* prop_table[2];
  const float* prop_w0[2];
  const float* prop_b0[2];
  const float* prop_w1[2];
  const float* prop_b1[2];
  const float* lin0;     // device copies of the host sampler tables
  const float* u1;
  const float* u2;
  float* weights0;
  float* weights1;
  float* s_edges1;
  float* s_edges2;
  float* t_edges2;
  int32_t* inds1;
  int32_t* inds2;
};

struct FieldParams {
  const float2* main_table;
  const float* base_w0; const float* base_b0;
  const float* base_w1; const float* base_b1;
  const float* head_w0; const float* head_b0;
  const float* head_w1; const float* head_b1;
  const float* head_w2; const float* head_b2;
  const float* appearance;
};

struct FieldArgs : RayArgs {
  FieldParams P;
  const float* t_edges2;   // (n, 49) chunk-local
  float* rgb;              // (n, 3)
  float* accumulation;     // (n) or null
  float* weights2;         // (n, 48) or null
  // forward-mode tangents (jvp kernel only)
  const float* d_o;        // (3, n_total)
  const float* d_v;
  float* d_rgb;            // (n, 3)
};

cudaError_t launch_sample_levels(const ModelDev& M, const SampleArgs& A, cudaStream_t stream);
cudaError_t launch_field_forward(const ModelDev& M, const FieldArgs& A, cudaStream_t stream);
cudaError_t launch_hash_encode(const GridDev& G, const float2* table, const float* x, int64_t n, float* out,
                               cudaStream_t stream);

}  // namespace nemit
