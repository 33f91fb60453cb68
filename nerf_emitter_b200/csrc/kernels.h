// kernels.h - argument blocks and launchers of the CUDA kernels (internal; the public ABI is include/nemit.h)
#pragma once
#include "common.cuh"

namespace nemit {

// "do once per device" latch for per-function attributes (one process may drive several devices through several handles)
struct DeviceOnce {
  bool done[64] = {};
  bool need() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
};

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
  int app_row;           // appearance row resolved by the host: >= 0 a row, kAppMean, kAppZero (field_common.cuh)
  const float* jitter;   // (n_total,3) or null
  float anneal;
  int nerf_layout;       // NemitRays.layout: 0 = (3,N) Mitsuba unit cube, 1 = (N,3) NeRF coordinates (also for tangents / ray grads)
  RotArgs rot;           // Rotater.apply_frustums for the samples THIS launch evaluates (mode 0 = off); set per stage by the
                         // host: the sampler stages always rotate, the final-sample stage only where the reference does
};
// the scene_scale argument of load_ray; 0 selects the (N,3) NeRF-coordinate layout
__host__ __device__ inline float ray_scale(const ModelDev& M, const RayArgs& A) { return A.nerf_layout ? 0.f : M.scene_scale; }

struct SampleArgs : RayArgs {
  const float2* prop_table[2];
  uint32_t* prop_table_h[2];   // fast sampler: packed-half copies of the two tables (handle-owned scratch, rewritten per launch)
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

constexpr int kEncTileBytes = 8192;     // 128 samples x 32 fp16
constexpr int kJacTileBytes = 24576;    // 128 samples x 96 fp16
constexpr int kActTileBytes = 24576;    // H1 (16384) | IN (8192)
constexpr int kKeptBytesPerSample = (kEncTileBytes + kJacTileBytes + kActTileBytes) / 128 + 16;   // 464 (round 1: 720)

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
  // adjoint-seed mode (first pass of the backward): composite adjoint per ray
  const float* s_edges2;   // (n, 49) chunk-local spacing edges (distortion loss)
  const float* grad_rgb;   // (n_total, 3)
  float* g_sigma;          // (n, 48)  dL/dsigma (post-selector density), gradient scaling applied
  float* g_rgbs;           // (n, 48, 3) dL/drgb_sample, gradient scaling applied
  int64_t reg_chunk;       // rays per backward chunk over which the regulariser means are taken
  float* g_absmax;         // or null: max |d/d(pre-activation) seed| of the launch (atomicMax on the float bits)
  // tensor-core seed pass only: the encoding rows and their Jacobians, kept for k_field_backward_tc so that the
  // backward never touches the parameter table (its L2 working set is the gradient table alone).
  // Tile = 128 consecutive samples of the chunk; 16-byte chunk c of row r of tile i lives at
  //   enc_tiles + i * 8192  + (c * 128 + r) * 16   (c < 4 : fp16 encoding, 8 values per chunk, = the UMMA A operand)
  //   jac_tiles + i * 24576 + (c * 128 + r) * 16   (c < 12: fp16 d enc / d offset, per level dx.xy dy.xy dz.xy)
  //   act_tiles + i * kActTileBytes + {0: H1 (8 chunks), 16384: IN = [SH16 | geo15 | 1] (4)} + (c * 128 + r) * 16
  //             fp16 operand rows of the base hidden layer and of the head input.  The two head hidden layers H2, H3 are
  //             NOT kept (round 2): k_field_backward_tc recomputes them from IN with two forward MMAs on the idle tensor
  //             pipe - same fp16 operands, same instruction shapes, hence the same bits and the same ReLU masks
  uint8_t* enc_tiles;
  uint8_t* jac_tiles;
  uint8_t* act_tiles;
  // d loss / d (rgb pre-activation xyz, density pre-activation w) per sample, i.e. the seeds with the SafeExp /
  // trunc_exp derivatives and the selector already applied (written instead of g_sigma / g_rgbs when set)
  float4* seeds4;
  // split-operand tensor-core mode only: the ReLU masks of the two head hidden layers per sample (bits 0..63 = H2 units,
  // 64..127 = H3 units), decided by the seed pass at 24 significant bits and applied by k_field_backward_tc, whose own
  // recompute of H2 / H3 has 16
  uint4* masks4;
  // light-point-cloud pass (SdfNerfactoModel.get_point_lights, NS/models/sdf_nerfacto.py:189-212): the adjoint seed is
  // the same luminance weight vector for every ray (calc_brightness_grad, NS/models/nerfacto.py:547-573), the pass also
  // writes the composited rgb and the "contrib" depth (DepthRenderer, NS/model_components/renderers.py:389-397)
  int point_lights;
  float lum_w[3];
  float* depth;            // (n)
};

// gradient accumulators in the torch parameter layouts (atomically accumulated)
struct FieldGrads {
  float2* main_table;
  float* base_w0; float* base_b0;
  float* base_w1; float* base_b1;
  float* head_w0; float* head_b0;
  float* head_w1; float* head_b1;
  float* head_w2; float* head_b2;
  float* appearance;
};

struct FieldBwdArgs : RayArgs {
  FieldParams P;
  FieldGrads G;
  const float* t_edges2;   // (n, 49)
  const float* g_sigma;    // (n, 48)
  const float* g_rgbs;     // (n, 48, 3)
  float* ray_grad;         // (n, 6) zero-initialised accumulators: d/d(o_nerf), d/d(d_nerf)
  float loss_scale;        // applied to parameter gradients only
  const float* g_absmax;   // tensor-core path: written by the seed pass (fp16 gradient operand scale)
  const uint8_t* enc_tiles;  // tensor-core path: see FieldArgs
  const uint8_t* jac_tiles;
  const uint8_t* act_tiles;
  const float4* seeds4;
  const uint4* masks4;           // split-operand mode: see FieldArgs
  unsigned long long* red_ops;   // or null: device counter of the reduction lane-ops the table scatter issues (profiling)
};

struct PropBwdArgs : SampleArgs {
  float2* g_table[2];
  float* g_w0[2]; float* g_b0[2]; float* g_w1[2]; float* g_b1[2];
  const float* weights2;   // (n, 48) final weights (detached in the interlevel loss)
  const float* s_edges2_in;// (n, 49)
  const float* s_edges1_in;// (n, 97) level-1 spacing edges of the SAME sampling pass (exact or fast sampler) that made s_edges2_in
  int64_t reg_chunk;
  float loss_scale;
};

struct FinalizeArgs : RayArgs {
  const float* ray_grad;   // (n, 6)
  float* grad_o;           // (3, n_total)
  float* grad_v;
};

// generic small MLP (component drop-in), torch (out,in) weights
struct MlpDev {
  int in_dim, num_layers;
  int out_dim[NEMIT_MLP_MAX_LAYERS];
  const float* weight[NEMIT_MLP_MAX_LAYERS];
  const float* bias[NEMIT_MLP_MAX_LAYERS];
  int activation, out_activation;
};

// optimizer step over the borrowed tensors (optim.cu)
constexpr int kOptMaxTensors = 24;
struct OptimGroup {
  float beta1, beta2, eps, max_norm, max_value;
  float omb1, omb2;     // (float)(1 - beta) with the subtraction in double, like torch's python-float arithmetic
  double lr, beta1_d, beta2_d;   // bias corrections are formed on the device from the group's own applied-step count
  int enabled;          // 0: the group has no gradients on this call (left untouched)
};
struct OptimArgs {
  int num_tensors;
  float* param[kOptMaxTensors];
  float* grad[kOptMaxTensors];
  float* exp_avg[kOptMaxTensors];
  float* exp_avg_sq[kOptMaxTensors];
  int64_t numel[kOptMaxTensors];
  int64_t chunk_end[kOptMaxTensors];   // filled by the launcher
  int group[kOptMaxTensors];           // 0 = fields, 1 = proposal_networks
  OptimGroup grp[2];
  float inv_scale;                     // 1 / GradScaler scale
  int zero_grads;
  double* stats;                       // device scratch (4): sum g^2 per group, non-finite flag per group
  float* stats_out;                    // optional user copy as float[4]
  long long* step_state;               // device int64[2]: applied steps per group (advanced by k_adam_finish)
};
cudaError_t launch_adam_step(OptimArgs& A, cudaStream_t stream);

// light point cloud -> spherical-Gaussian mixture (gmm.cu)
struct GmmArgs {
  const float* x;        // (n,3)
  const float* w;        // (n) sample weights (zero = point ignored)
  int64_t n;
  int k;
  const float* mean;     // (k,3) current parameters
  const float* var;      // (k)   isotropic variances
  const float* prior;    // (k)
  float min_var;
  double* stats;         // (k*6 + 2) scratch
  float* mean_out; float* var_out; float* prior_out;
  float* logp_out;       // (1) weighted log-likelihood under the input parameters
};
cudaError_t launch_compensate(const float* lum, int64_t n, float threshold, int use_threshold, double* scratch, float* w,
                              cudaStream_t stream);
cudaError_t launch_gmm_em(const GmmArgs& A, cudaStream_t stream);

// spherical-sensor bake (bake.cu)
struct SphericalArgs {
  float origin[3];
  float near_clip;
  int width, height, spp;
  uint32_t seed;
  int64_t first, n;        // samples [first, first + n) in pixel-major order
  const float* jitter;     // (n,2) or null -> counter-based RNG
  float* o;                // (3,n)
  float* v;                // (3,n)
  float* pos;              // (n,2) film positions in pixel units
};
struct FilmArgs {
  int width, height, filter;   // filter 0 = box, 1 = gaussian
  float stddev;
  const float* pos;        // (n,2)
  const float* rgb;        // (n,3)
  int64_t n;
  float4* accum;           // (H,W) of (sum w rgb, sum w)
};
cudaError_t launch_spherical_rays(const SphericalArgs& A, cudaStream_t stream);
cudaError_t launch_film_accumulate(const FilmArgs& A, cudaStream_t stream);
cudaError_t launch_film_develop(int64_t n, const float* accum, float* rgba, cudaStream_t stream);

cudaError_t launch_field_samples(const ModelDev& M, const FieldParams& P, const float* pos, const float* dir, int64_t n,
                                 int app_row, float* density, float* geo, float* rgb,
                                 cudaStream_t stream);
cudaError_t launch_prop_density(const ModelDev& M, const GridDev& G, const float2* table, const float* w0,
                                const float* b0, const float* w1, const float* b1, const float* pos, int64_t n,
                                float* density, cudaStream_t stream);
cudaError_t launch_sh_encode(const float* d, int64_t n, int levels, float* out, cudaStream_t stream);
cudaError_t launch_mlp(const MlpDev& L, const float* x, int64_t n, float* out, cudaStream_t stream);
cudaError_t launch_emitter_local_rays(const float* m34, const float* bmin, const float* bmax, float skip_distance,
                                      int outside_only, int mode, const float* o, const float* v, int64_t n,
                                      float* o_out, float* v_out, float* near_out, cudaStream_t stream);
cudaError_t launch_sample_levels(const ModelDev& M, const SampleArgs& A, cudaStream_t stream, bool fast = false);
cudaError_t launch_field_forward(const ModelDev& M, const FieldArgs& A, cudaStream_t stream);
// pieces = 1: fp16 operands; 2: fp32 arithmetic through split bf16 operands (forward: 3 pieces; reverse: 2 + the forward's ReLU masks)
cudaError_t launch_field_forward_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream, int pieces = 1);
cudaError_t launch_field_adjoint_seed_tc(const ModelDev& M, const FieldArgs& A, cudaStream_t stream, int pieces = 1);
cudaError_t launch_field_adjoint_seed(const ModelDev& M, const FieldArgs& A, cudaStream_t stream);
cudaError_t launch_field_jvp(const ModelDev& M, const FieldArgs& A, cudaStream_t stream);
cudaError_t launch_field_backward(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream);
cudaError_t launch_field_backward_tc(const ModelDev& M, const FieldBwdArgs& A, cudaStream_t stream, int pieces = 1);
cudaError_t launch_prop_backward(const ModelDev& M, const PropBwdArgs& A, cudaStream_t stream);
cudaError_t launch_finalize_ray_grads(const ModelDev& M, const FinalizeArgs& A, cudaStream_t stream);
cudaError_t launch_hash_encode(const GridDev& G, const float2* table, const float* x, int64_t n, float* out,
                               cudaStream_t stream);

}  // namespace nemit
