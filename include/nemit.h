/*
 * nemit.h - C ABI of libnerfemitter.so: the B200 (sm_100a) NeRF-emitter evaluation path.
 *
 * Drop-in boundary for gerwang/nerf-emitter's emitter query.  Every entry point
 * cites the reference interface it replaces (paths relative to the reference checkout;
 * NS = nerfstudio/, DS = differentiable-sdf-rendering/python/).
 *
 * Conventions
 *   - plain pointers and sizes; all data pointers are DEVICE pointers unless named *_host;
 *   - every call is asynchronous on the given cudaStream_t (passed as void*), no hidden host sync;
 *   - return value: 0 = ok, <0 = error (see NEMIT_E_*); nemit_last_error() gives the message;
 *   - the library BORROWS parameter / gradient memory (torch owns it); gradients are ACCUMULATED
 *     (+=) like torch's .backward() over chunks (NS/pipelines/mitsuba_sdf.py:1180-1190);
 *   - rays come in as struct-of-arrays (3,N) fp32 exactly as the dr.CustomOp receives them
 *     (DS/emitters/nerf.py:12-17, nerf_emitter_op.py:41-42) and radiance leaves as (N,3) row-major
 *     (nerf_emitter_op.py:35-36), so no transposes are needed on either side.
 */
#ifndef NEMIT_H_
#define NEMIT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NEMIT_OK 0
#define NEMIT_E_INVALID (-1)   /* bad argument / unsupported shape */
#define NEMIT_E_CUDA (-2)      /* CUDA runtime error */
#define NEMIT_E_UNBOUND (-3)   /* parameters (or gradients) not bound */
#define NEMIT_E_WORKSPACE (-4) /* workspace too small */

#define NEMIT_MAX_LEVELS 16

/* One multiresolution hash grid: NS/field_components/encodings.py:264-312.  `scalings` are the
 * per-level floor(min_res * g^l) values the host computed exactly as encodings.py:281-283 does
 * (main grid ends at 2047, not 2048).  Table layout: (num_levels << log2_table_size, 2) fp32. */
typedef struct NemitGrid {
  int32_t num_levels;
  int32_t log2_table_size;
  float scalings[NEMIT_MAX_LEVELS];
} NemitGrid;

/* Model constants: NS/models/nerfacto.py:58-157 as overridden by sdf-nerfacto
 * (NS/configs/method_configs.py:314-326). */
typedef struct NemitConfig {
  NemitGrid main_grid;          /* 16 levels, T=2^19 */
  NemitGrid prop_grid[2];       /* 5 levels, T=2^17 */
  int32_t num_prop_samples[2];  /* 256, 96   (NS/models/nerfacto.py:85-88) */
  int32_t num_nerf_samples;     /* 48 */
  int32_t hidden_dim;           /* 64 */
  int32_t geo_feat_dim;         /* 15 */
  int32_t appearance_dim;       /* 32 */
  int32_t prop_hidden_dim;      /* 16 */
  int32_t num_images;           /* rows of the appearance embedding */
  float near_plane;             /* 0.05, used in training mode only (scene_colliders.py:217-222) */
  float far_plane;              /* 1000 */
  float scene_scale;            /* DS/emitters/nerf_op.py:76-89 */
  float aabb[6];                /* min xyz, max xyz of the scene box */
  float rgb_bias;               /* SafeExp bias, -5 (NS/fields/nerfacto_field.py:40-57) */
  int32_t disable_inside_aabb;  /* 1 during emitter queries (NS/pipelines/mitsuba_sdf.py:922) */
  int32_t cpu_sum_width;        /* ATen CPU torch.sum vector width to reproduce (8); 0 = fp64 sum */
  float interlevel_mult;        /* 1.0   (NS/models/nerfacto.py:106) */
  float distortion_mult;        /* 7e-3  (NS/configs/method_configs.py:320) */
  /* sampler tables computed by the host with torch.linspace so they are bit-identical to
   * NS/model_components/ray_samplers.py:104 and :323-331 */
  const float* lin0_host;       /* num_prop_samples[0]+1 values of linspace(0,1,S0+1) */
  const float* u1_host;         /* num_prop_samples[1]+1 values of linspace(0,1-1/nb,nb)       */
  const float* u2_host;         /* num_nerf_samples+1 values                                   */
  /* Model variants this build does not implement: nemit_create fails with NEMIT_E_INVALID (and says which) when one is
   * set, instead of silently evaluating the default model.  use_fake_contraction: FakeContraction + AABBBoxCollider
   * (NS/field_components/spatial_distortions.py:94-103, NS/models/nerfacto.py:144-154,176,263-270); far_intersect:
   * AABBBoxFarIntersectCollider in get_point_lights (NS/model_components/scene_colliders.py:96-143,
   * NS/models/sdf_nerfacto.py:193-194).  Both default to False in sdf-nerfacto. */
  int32_t use_fake_contraction;
  int32_t far_intersect;
} NemitConfig;

/* Parameter pointers, named as in the reference state_dict (SURVEY.md 8b "Parameter ownership").
 * The same struct carries the gradient accumulators for nemit_bind_grads. */
typedef struct NemitParams {
  float* main_table;            /* field.mlp_base_grid.hash_table            (16*2^19, 2) */
  float* base_w0; float* base_b0; /* field.mlp_base_mlp.layers.0  (64,32),(64) */
  float* base_w1; float* base_b1; /* field.mlp_base_mlp.layers.1  (16,64),(16) */
  float* head_w0; float* head_b0; /* field.mlp_head.layers.0      (64,63),(64) */
  float* head_w1; float* head_b1; /* field.mlp_head.layers.1      (64,64),(64) */
  float* head_w2; float* head_b2; /* field.mlp_head.layers.2      (3,64),(3)   */
  float* appearance;            /* field.embedding_appearance.embedding.weight (num_images,32) */
  float* prop_table[2];         /* proposal_networks.i.encoding.hash_table   (5*2^17, 2) */
  float* prop_w0[2]; float* prop_b0[2]; /* proposal_networks.i.mlp_base.1.layers.0 (16,10),(16) */
  float* prop_w1[2]; float* prop_b1[2]; /* proposal_networks.i.mlp_base.1.layers.1 (1,16),(1)   */
} NemitParams;

/* Per-call ray batch: what NeRFEmitterOp.eval/forward/backward get from Mitsuba
 * (DS/emitters/nerf_emitter_op.py:18-21, 38-117). */
typedef struct NemitRays {
  const float* o;        /* (3,N) origins in Mitsuba unit-cube coordinates */
  const float* v;        /* (3,N) directions */
  const float* near_per_ray; /* (N) or NULL -> near_scalar; o' = o + near*v (nerf_emitter_op.py:28,43) */
  float near_scalar;
  int64_t n;
  int32_t camera_idx;    /* appearance row in training mode (nerfacto_field.py:272-285) */
  int32_t training;      /* 1: model.train() semantics (near plane 0.05, embedding row, no clamp) */
  const float* jitter;   /* (N,3) per-ray U[0,1) for the three single-jitter samplers, or NULL =
                            deterministic bins (ray_samplers.py:107-117, 323-333). The host draws
                            them from the samplers' own generators. */
  float anneal;          /* ProposalNetworkSampler._anneal (ray_samplers.py:609); 1.0 = no-op */
  int32_t layout;        /* NEMIT_RAYS_MITSUBA_3N (0, default): o, v as above, what the dr.CustomOp passes;
                            NEMIT_RAYS_NERF_N3: o, v are RayBundle.origins / .directions, (N,3) rows already in NeRF
                            coordinates (NS/cameras/rays.py:170-200), i.e. the model-level entry points
                            get_rgb_for_camera_ray_bundle / forward_grad_for_camera_ray_bundle /
                            backward_for_camera_ray_bundle (NS/models/base_model.py:290-344,
                            NS/pipelines/mitsuba_sdf.py:1144-1219); tangents and ray gradients then use (N,3) too */
  /* Optional: sample edges kept from an earlier nemit_forward / nemit_sample_levels call (NemitLevelsOut.s_edges1,
   * .s_edges2, .t_edges2) over the SAME rays, mode (training flag, near), anneal, jitter and parameters.  The reference
   * re-derives identical samples in the adjoint of one CustomOp instance because it restores the samplers' generator
   * state between the primal and the adjoint (NS/models/base_model.py:299,327-328, NS/pipelines/mitsuba_sdf.py:1152-1154);
   * with these set nemit_backward / nemit_jvp / nemit_point_lights skip the proposal sampling pass instead.
   * kept_t_edges2 and kept_s_edges2 (N,49) must come together; kept_s_edges1 (N,97) is needed in addition by
   * nemit_backward when proposal_requires_grad != 0.  All NULL (default) = sample again. */
  const float* kept_s_edges1;
  const float* kept_s_edges2;
  const float* kept_t_edges2;
} NemitRays;
#define NEMIT_RAYS_MITSUBA_3N 0
#define NEMIT_RAYS_NERF_N3 1

/* Optional per-level outputs of the proposal loop, for parity tests and for
 * ProposalNetworkSampler's return value (ray_samplers.py:584-626). Any pointer may be NULL. */
typedef struct NemitLevelsOut {
  float* weights0;   /* (N,S0) */
  float* weights1;   /* (N,S1) */
  float* s_edges1;   /* (N,S1+1) spacing-domain bin edges of level 1 */
  float* s_edges2;   /* (N,S2+1) */
  float* t_edges2;   /* (N,S2+1) euclidean bin edges of the final samples */
  int32_t* inds1;    /* (N,S1+1) searchsorted(cdf,u,right) of PDF stage 1 */
  int32_t* inds2;    /* (N,S2+1) */
  float* weights2;   /* (N,S2) final weights (filled by nemit_forward only) */
} NemitLevelsOut;

typedef struct NemitHandle NemitHandle;

const char* nemit_last_error(void);
int nemit_version(void);

/* lifetime: one handle per device (SURVEY.md 8b "Threading / streams") */
int nemit_create(const NemitConfig* cfg, int device, NemitHandle** out);
int nemit_destroy(NemitHandle* h);

/* Which appearance vector a call sees.  Default (no call): the reference without a Rotater, i.e. the camera's row of
 * field.embedding_appearance in training mode and the mean over rows in eval mode
 * (NS/fields/nerfacto_field.py:272-285, use_average_appearance_embedding NS/models/nerfacto.py:116).
 * always_use_camera_indices != 0 is the model built with a Rotater (turntable captures: num_images = number of
 * rotations and always_use_camera_indices = True, NS/models/nerfacto.py:186-206): the row is used in eval mode too.
 * rotation_ids_host (num_rotation_ids entries, copied; may be null) is Rotater.rotation_ids: every ray-level entry
 * maps NemitRays.camera_idx -> rotation_ids[min(camera_idx, num_rotation_ids - 1)] as Rotater.map_rotation_ids does
 * (NS/field_components/rotater.py:94-100, call sites NS/models/sdf_nerfacto.py:71-72,92-93,127-128,198-199);
 * nemit_field_samples takes an already mapped index.  use_average_appearance_embedding == 0 gives zeros in eval. */
int nemit_set_appearance_mode(NemitHandle* h, int32_t always_use_camera_indices,
                              int32_t use_average_appearance_embedding, const int32_t* rotation_ids_host,
                              int32_t num_rotation_ids);

/* Per-sample frustum rotation of turntable captures: Rotater.apply_frustums (NS/field_components/rotater.py:125-149,
 * hooked into RayBundle.get_ray_samples, NS/cameras/rays.py:285-286, by get_ray_bundle, DS/emitters/nerf_op.py:86-87).
 * transforms_host: num_rotations x (3 x 4) row-major [R | t] in NeRF coordinates = Rotater.get_torch_transforms(...)[:, :3, :4]
 * (call again when a rotation optimizer has moved them); the rotation of a call is the one of rotation id =
 * rotation_ids[min(camera_idx, len - 1)] (nemit_set_appearance_mode; identity mapping without a table).
 * bound_type NEMIT_ROT_SPHERE: bound_host = {centre xyz, radius} (|x - c| < r); NEMIT_ROT_BOX: {min xyz, max xyz} (inclusive).
 * A sample whose UNROTATED position lies inside the bound takes origin R o + t and direction R d.  Which stages rotate
 * follows the reference to the letter:
 *   proposal levels (sampling, and the proposal networks' backward)  always - every get_ray_samples call runs the hook;
 *   final 48 samples in nemit_forward                                rotated (positions and SH directions);
 *   final samples in nemit_jvp                                       primal rotated, tangents those of the unrotated ray
 *                                                                    (NS/models/sdf_nerfacto.py:96-104);
 *   final samples in nemit_backward                                  NOT rotated: get_backward_outputs puts the ray's own
 *                                                                    origins / directions back (sdf_nerfacto.py:129-131);
 *   final samples in nemit_point_lights                              rotated, the origin gradient goes back through R^T
 *                                                                    (calc_brightness_grad, NS/models/nerfacto.py:547-556).
 * num_rotations == 0 removes the rotater. */
#define NEMIT_ROT_SPHERE 1
#define NEMIT_ROT_BOX 2
int nemit_set_rotater(NemitHandle* h, int32_t num_rotations, const float* transforms_host, int32_t bound_type,
                      const float* bound_host);

/* borrow parameter / gradient pointers (never copied, never freed) */
int nemit_bind_params(NemitHandle* h, const NemitParams* p);
int nemit_bind_grads(NemitHandle* h, const NemitParams* g);

/* MLP arithmetic of the main field: NEMIT_PRECISION_FP32 = exact fp32 FMA path (parity mode, <=1e-3 vs the
 * reference's fp32 torch path); NEMIT_PRECISION_FP16_TC = fp16 operands on the tcgen05 tensor cores with fp32
 * accumulation - the reference's own mixed-precision mode (torch.autocast, NS/models/base_model.py:305;
 * tiny-cuda-nn FullyFusedMLP is fp16, NS/field_components/mlp.py:117-144).  Hash tables, density / colour
 * activations, weights and compositing are fp32 in both.  Default: FP32. */
#define NEMIT_PRECISION_FP32 0
#define NEMIT_PRECISION_FP16_TC 1
/* fp32 MLP arithmetic ON the tensor cores (round 2), same kernels as NEMIT_PRECISION_FP16_TC with split operands: every
 * operand element as bf16 pieces x0 + x1 (+ x2), a product = the sum of the piece products in one fp32 TMEM accumulator.
 * Forward (nemit_forward and the seed pass of nemit_backward): three pieces = 24 significant bits, six MMAs per layer.
 * Reverse (k_field_backward_tc): two pieces = 16 bits, three MMAs per product, with the ReLU masks decided by the
 * 24-bit forward - so the discontinuous part of the derivative is the fp32 one and the rest is smooth at ~1e-5.
 * nemit_jvp stays on the fp32 FMA path in every mode. */
#define NEMIT_PRECISION_SPLIT_TC 2
int nemit_set_precision(NemitHandle* h, int32_t precision);

/* Arithmetic of the proposal-network sampler (ProposalNetworkSampler, NS/model_components/ray_samplers.py:584-626):
 * NEMIT_SAMPLER_EXACT (default) reproduces the reference's fp32 CPU arithmetic operation by operation, so that the
 * searchsorted indices are bit-equal (parity mode); NEMIT_SAMPLER_FAST evaluates the two proposal density fields
 * (hash blend + 10->16->1 MLP) in packed half precision, which is what the reference does under torch.autocast /
 * tiny-cuda-nn - sample positions then agree to ~1e-3 of a bin, not bit for bit.  The PDF resampling itself
 * (cumsum / searchsorted) is the same code in both modes. */
#define NEMIT_SAMPLER_EXACT 0
#define NEMIT_SAMPLER_FAST 1
int nemit_set_sampler(NemitHandle* h, int32_t sampler);

/* bytes of scratch the calls below need for a batch of n rays (chunked internally) */
size_t nemit_workspace_bytes(const NemitHandle* h, int64_t n);

/* replaces NeRFEmitterOp.ns_render_ray -> Model.get_rgb_for_camera_ray_bundle ->
 * SdfNerfactoModel.get_rgb (nerf_emitter_op.py:38-58, NS/models/base_model.py:290-313,
 * NS/models/sdf_nerfacto.py:64-83).  rgb (N,3); accumulation (N) optional. */
int nemit_forward(NemitHandle* h, const NemitRays* rays, float* rgb, float* accumulation,
                  const NemitLevelsOut* levels_or_null, void* workspace, size_t workspace_bytes,
                  void* stream);

/* nemit_forward for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  The call allocates
 * nothing and takes no workspace: its scratch (the final sample edges, 49 floats per ray) is a static device array of
 * NEMIT_NOALLOC_MAX_RAYS rays = 3.06 MiB that every GPU loading the library holds until exit.  rays->n above the maximum
 * is rejected with NEMIT_E_INVALID.  Same results as nemit_forward (same kernels); rgb (N,3) only, no per-level outputs,
 * no profile records.  All calls share the one array: calls that overlap on a GPU (different streams) race - serialise
 * them. */
#define NEMIT_NOALLOC_MAX_RAYS 16384
int nemit_forward_noalloc(NemitHandle* h, const NemitRays* rays, float* rgb, void* stream);

/* proposal loop only: ProposalNetworkSampler.generate_ray_samples (ray_samplers.py:584-626) */
int nemit_sample_levels(NemitHandle* h, const NemitRays* rays, const NemitLevelsOut* out,
                        void* workspace, size_t workspace_bytes, void* stream);

/* replaces NeRFEmitterOp.ns_forward_ray -> forward_grad_for_camera_ray_bundle ->
 * SdfNerfactoModel.forward_grad (nerf_emitter_op.py:60-90, base_model.py:315-344,
 * sdf_nerfacto.py:85-113).  d_o, d_v (3,N) tangents; d_rgb (N,3), NaN tangents zeroed. */
int nemit_jvp(NemitHandle* h, const NemitRays* rays, const float* d_o, const float* d_v,
              float* d_rgb, void* workspace, size_t workspace_bytes, void* stream);

/* replaces NeRFEmitterOp.ns_backward_ray -> backward_for_camera_ray_bundle ->
 * get_backward_outputs + get_reg_loss_dict (nerf_emitter_op.py:92-117,
 * NS/pipelines/mitsuba_sdf.py:1144-1219, sdf_nerfacto.py:115-187).
 * grad_rgb (N,3); grad_o, grad_v (3,N) are OVERWRITTEN (already include the -near*grad_v
 * change of variables, nerf_emitter_op.py:112); parameter grads are accumulated into the bound
 * buffers.  chunk_rays = rays per backward chunk over which the regulariser means are taken
 * (backward_num_rays_per_chunk); loss_scale = GradScaler scale applied to parameter grads only.
 * proposal_requires_grad = the sampler's `updated` flag (ray_samplers.py:598,613-618). */
int nemit_backward(NemitHandle* h, const NemitRays* rays, const float* grad_rgb,
                   int64_t chunk_rays, float loss_scale, int32_t proposal_requires_grad,
                   float* grad_o, float* grad_v, void* workspace, size_t workspace_bytes,
                   void* stream);

/* component drop-in: HashEncoding.forward (NS/field_components/encodings.py:341-382).
 * x (n,3) in [0,1]; out (n, 2*num_levels). which: 0 = main grid, 1/2 = proposal grids. */
int nemit_hash_encode(NemitHandle* h, int32_t which, const float* x, int64_t n, float* out,
                      void* stream);

/* Light-point-cloud pass: SdfNerfactoModel.get_point_lights (NS/models/sdf_nerfacto.py:189-212) =
 * proposal sampling + calc_brightness_grad (NS/models/nerfacto.py:547-573) + DepthRenderer("contrib")
 * (NS/model_components/renderers.py:389-397).  rgb_weights: 3 HOST floats (0.299, 0.587, 0.114 in the reference);
 * rgb (N,3) composited radiance; depth (N) = midpoint of the sample with the largest weight * luminance;
 * brightness_grad = d sum(rgb * rgb_weights) / d origins in the rays' own layout ((3,N) or (N,3)), t^2 gradient
 * scaling applied as in the reference's reverse mode.  No parameter gradient is touched. */
int nemit_point_lights(NemitHandle* h, const NemitRays* rays, const float* rgb_weights, float* rgb, float* depth,
                       float* brightness_grad, void* workspace, size_t workspace_bytes, void* stream);

/* Optimizer step on the bound parameters / gradients: Optimizers.optimizer_scaler_step_all
 * (NS/engine/optimizers.py:122-137, called from NS/engine/trainer.py:517-537) for the two parameter groups of
 * NS/models/nerfacto.py:313-317 with the hyper-parameters of NS/configs/method_configs.py:331-342:
 * GradScaler.unscale_ (grad / grad_scale + non-finite check) -> clip_grad_norm_(group, max_norm) ->
 * clip_grad_value_(group, max_value) -> torch.optim.Adam step, in place on the borrowed tensors, PER GROUP as the
 * reference does it (one optimizer and one found_inf record per group):
 *   - a group with enabled == 0 is left untouched - parameters, moments and step count - which is what the reference does
 *     for an optimizer none of whose parameters has a .grad (optimizers.py:136: the proposal networks on every call
 *     whose backward chunks ran without proposal_requires_grad);
 *   - a group whose unscaled gradients contain a non-finite value skips ITS step only (GradScaler keeps found_inf per
 *     optimizer) and its step count does not advance (torch Adam's `step` only moves when optimizer.step() runs), so the
 *     bias corrections stay in sync after an overflow.
 * exp_avg / exp_avg_sq: Adam state in the parameter layouts (caller-owned, zero-initialised); step_state: DEVICE
 * int64[2] (caller-owned, zero-initialised) = applied steps of {fields, proposal_networks}, read and advanced on the
 * device so that a skipped step needs no host synchronisation; lr: the scheduler's current value
 * (NS/engine/schedulers.py:156-193); max_norm / max_value <= 0 disable the clip; zero_grads != 0 clears every gradient
 * in the same pass (optimizer.zero_grad).  stats_or_null: device float[4] = {sum g^2 fields, sum g^2 proposal_networks,
 * non-finite flag fields, non-finite flag proposal_networks} for GradScaler.update / logging. */
typedef struct NemitAdamGroup {
  double lr, beta1, beta2, eps;   /* python floats in torch.optim.Adam: 1 - beta is formed in double, as torch does */
  float max_norm, max_value;
  int32_t enabled;                /* 0: no parameter of the group has a gradient on this call */
  int32_t reserved;
} NemitAdamGroup;
int nemit_adam_step(NemitHandle* h, const NemitParams* exp_avg, const NemitParams* exp_avg_sq,
                    const NemitAdamGroup* fields, const NemitAdamGroup* proposal_networks, int64_t* step_state,
                    float grad_scale, int32_t zero_grads, float* stats_or_null, void* stream);

/* Headless spherical-sensor bake of the emitter radiance (BASELINE config 5): SphericalCamera.sample_ray
 * (DS/sensors/spherical_sensor.py:19-44 with assets/sensors/spherical_sensor.xml: to_world = translate, near_clip,
 * hdrfilm rgba with a gaussian rfilter of stddev 0.25) -> nemit_forward -> film accumulation -> develop, all on the
 * device.  Film samples are numbered pixel-major: sample s of pixel (x,y) has index (y*width + x)*spp + s. */
typedef struct NemitSphericalSensor {
  float origin[3];       /* to_world translation, Mitsuba unit-cube coordinates (0.5,0.5,0.5 in the reference XML) */
  float near_clip;
  int32_t width, height; /* film size */
  int32_t spp;           /* samples per pixel */
  uint32_t seed;         /* sub-pixel positions come from a counter-based RNG unless `jitter` is given */
} NemitSphericalSensor;
#define NEMIT_FILTER_BOX 0
#define NEMIT_FILTER_GAUSSIAN 1
/* rays of film samples [first, first+n): o, v (3,n) in the CustomOp layout, pos (n,2) film positions in pixels;
 * jitter_or_null (n,2) in [0,1) replaces the RNG (parity tests) */
int nemit_spherical_rays(const NemitSphericalSensor* s, int64_t first, int64_t n, const float* jitter_or_null,
                         float* o, float* v, float* pos, void* stream);
/* accum (height,width,4) += (w*rgb, w) for n samples; box: w = 1 in the pixel containing the sample; gaussian:
 * separable exp(-x^2/(2 s^2)) - exp(-r^2/(2 s^2)), r = 4 s, over the pixel centres within r */
int nemit_film_accumulate(int32_t width, int32_t height, int32_t filter, float stddev, const float* pos,
                          const float* rgb, int64_t n, float* accum, void* stream);
/* rgba (height,width,4) = (rgb / weight, 1) */
int nemit_film_develop(int32_t width, int32_t height, const float* accum, float* rgba, void* stream);
/* the whole bake for film rows [row_begin,row_end) (a rank's shard) into a zero-initialised full-size accum;
 * ranks sum their accum buffers, then develop.  workspace >= nemit_bake_workspace_bytes(h). */
size_t nemit_bake_workspace_bytes(const NemitHandle* h);
int nemit_bake_envmap(NemitHandle* h, const NemitSphericalSensor* s, int32_t filter, float stddev, int32_t row_begin,
                      int32_t row_end, int32_t camera_idx, float* accum, void* workspace, size_t workspace_bytes,
                      void* stream);

/* Light point cloud -> K spherical Gaussians, the producer of the vMF lobe table.
 * nemit_compensate_weights: compensate_pc (NS/model_components/output_light_pc.py:176-191): w_i = max(lum_i - mean(lum), 0),
 * zeroed unless lum_i > threshold (use_threshold != 0); points keep their place, a zero weight removes them from the fit.
 * nemit_gmm_em_step: one EM iteration of cluster_light's weighted mixture of isotropic normals
 * (NS/model_components/gmm_cluster_light.py:41-85, pomegranate GeneralMixtureModel + Normal('sphere'); pomegranate is not
 * in the reference tree: restated as the textbook weighted EM, parity unpinned).  x (n,3), w (n); (mean (k,3), var (k),
 * prior (k)) in -> out (may not alias); logp_out (1) = sum_i w_i log p(x_i) under the INPUT parameters;
 * scratch: device, >= (6k + 2) doubles; k <= 128; variances are floored at min_var. */
int nemit_compensate_weights(const float* lum, int64_t n, float threshold, int32_t use_threshold, float* w,
                             void* scratch, void* stream);
int nemit_gmm_em_step(const float* x, const float* w, int64_t n, int32_t k, const float* mean, const float* var,
                      const float* prior, float min_var, float* mean_out, float* var_out, float* prior_out,
                      float* logp_out, void* scratch, void* stream);

/* ---- component-level drop-ins (SURVEY.md 8b "Model-side API kept intact"): what the nerfstudio modules'
 * forward() dispatches to when a module is called on its own, outside the fused emitter query ---- */

/* HashEncoding.forward for any grid / table (NS/field_components/encodings.py:341-382): x (n,3) in [0,1],
 * table (num_levels << log2_table_size, 2), out (n, 2*num_levels). */
int nemit_hash_encode_grid(const NemitGrid* grid, const float* table, const float* x, int64_t n, float* out,
                           void* stream);

/* SHEncoding.forward, torch path (NS/field_components/encodings.py:708-716 -> components_from_spherical_harmonics,
 * NS/utils/math.py:27-90): d (n,3), out (n, levels^2), levels in 1..4. */
int nemit_sh_encode(const float* d, int64_t n, int32_t levels, float* out, void* stream);

/* NeRFEmitter.eval_nerf up to the op call (DS/emitters/nerf.py:69-82): world rays o, v (3,n) -> emitter-local rays
 * (o' = M o as a point, v' = M v as a vector; M = world_transform().inverse() as 3x4 row-major host floats) and, when
 * near_out != NULL, the distance to skip: the far slab intersection with [bbox_min, bbox_max] (the unit cube padded by
 * padding_size, nerf.py:33) for rays that hit it when outside_only, else skip_distance (nerf.py:75-79).
 * linear_only != 0 applies only the 3x3 part to both arrays (pass M^T to carry o / v gradients back to world space).
 * In-place (o_out == o) is allowed.  Stateless: no handle. */
int nemit_emitter_local_rays(const float* world_to_local_3x4, const float* bbox_min, const float* bbox_max,
                             float skip_distance, int32_t outside_only, int32_t linear_only, const float* o,
                             const float* v, int64_t n, float* o_out, float* v_out, float* near_out, void* stream);

/* MLP.forward, torch path (NS/field_components/mlp.py:163-183): nn.Linear layers in torch (out,in) layout,
 * `activation` between layers and `out_activation` after the last (0 = none, 1 = ReLU); widths <= 64. */
#define NEMIT_MLP_MAX_LAYERS 8
#define NEMIT_MLP_MAX_WIDTH 64
typedef struct NemitMlp {
  int32_t in_dim;
  int32_t num_layers;
  int32_t out_dim[NEMIT_MLP_MAX_LAYERS];
  const float* weight[NEMIT_MLP_MAX_LAYERS];   /* (out_dim[l], in) */
  const float* bias[NEMIT_MLP_MAX_LAYERS];     /* (out_dim[l]) or NULL */
  int32_t activation;
  int32_t out_activation;
} NemitMlp;
int nemit_mlp(const NemitMlp* mlp, const float* x, int64_t n, float* out, void* stream);

/* Field.density_fn (NS/fields/base_field.py:49-69) on explicit positions (n,3) in NeRF coordinates:
 * which = 0: NerfactoField.get_density (NS/fields/nerfacto_field.py:228-256), geo_or_null (n,15) receives the
 * geometry features; which = 1/2: HashMLPDensityField.get_density (NS/fields/density_fields.py:95-122). */
int nemit_density_fn(NemitHandle* h, int32_t which, const float* positions, int64_t n, float* density,
                     float* geo_or_null, void* stream);

/* NerfactoField.forward(ray_samples) = get_density + get_outputs (NS/fields/base_field.py:124-143,
 * NS/fields/nerfacto_field.py:228-334) on explicit sample positions and directions (n,3):
 * density (n), rgb (n,3) (SafeExp output, HDR).  training selects the appearance row camera_idx, otherwise
 * the mean embedding (nerfacto_field.py:272-285). */
int nemit_field_samples(NemitHandle* h, const float* positions, const float* directions, int64_t n,
                        int32_t camera_idx, int32_t training, float* density, float* rgb, void* stream);

/* vMF-mixture emitter: VonMisesFisherEmitter.sample_direction / pdf_direction
 * (DS/emitters/vMF.py:24-81).  lobes: position (K,3), weight (K) unnormalised, std (K).
 * p (N,3) shading points in emitter-local coordinates, u (N,2); d (N,3), pdf (N). */
typedef struct NemitVmfLobes {
  const float* position;
  const float* weight;
  const float* std;
  int32_t num_lobes;
} NemitVmfLobes;
int nemit_vmf_sample(const NemitVmfLobes* lobes, const float* p, const float* u, int64_t n,
                     float* d, float* pdf, void* stream);
int nemit_vmf_pdf(const NemitVmfLobes* lobes, const float* p, const float* d, int64_t n,
                  float* pdf, void* stream);

/* The same two calls as the emitter makes them, in WORLD space (vMF.py:24-56, 64-81): frame holds world_transform() and its
 * inverse as 3 x 4 row-major HOST floats (Transform4f.transform_affine: points take the translation, vectors do not) and the
 * emitter's bounding sphere m_bsphere.  it_p (N,3) world-space reference points.  sample: p = W^-1 it.p, lobe choice and warp
 * as above, d = normalize(W d_local), ds.pdf = pdf_direction(it, ds), dist = 2 max(r_bsphere, |it.p - c|), ds.p = it.p +
 * d dist (ds.n = -d, ds.uv = 0, ds.delta = false and the returned weight 0 carry no arithmetic); dist_or_null (N) and
 * ds_p_or_null (N,3) may be NULL.  pdf: d = normalize(W^-1 ds.d), p = W^-1 it.p. */
typedef struct NemitVmfFrame {
  float world_to_local[12];
  float local_to_world[12];
  float bsphere_center[3];
  float bsphere_radius;
} NemitVmfFrame;
int nemit_vmf_sample_direction(const NemitVmfLobes* lobes, const NemitVmfFrame* frame, const float* it_p, const float* u,
                               int64_t n, float* d, float* pdf, float* dist_or_null, float* ds_p_or_null, void* stream);
int nemit_vmf_pdf_direction(const NemitVmfLobes* lobes, const NemitVmfFrame* frame, const float* it_p, const float* ds_d,
                            int64_t n, float* pdf, void* stream);

/* end-to-end entry with HOST (pinned) ray and radiance buffers: H2D copy, forward, D2H copy,
 * all queued on `stream`; the device staging lives in the workspace. */
int nemit_forward_host(NemitHandle* h, const NemitRays* rays_host, float* rgb_host,
                       void* workspace, size_t workspace_bytes, void* stream);

/* Reverse-mode twin of nemit_forward_host (measurement honesty, SURVEY.md 8d: the metric's fwd+bwd mode through one
 * plugin call with HOST buffers): rays_host (o, v (3,N), optional near / jitter) and grad_rgb_host (N,3) are pinned
 * host arrays; H2D copies, nemit_forward (eval-mode primal when primal_training == 0, as a validation render, or the
 * training-mode primal of the same CustomOp instance when != 0 - its sample edges are then kept for the adjoint
 * unless proposal_requires_grad needs the proposal loop again), nemit_backward and the D2H copies of rgb (N,3),
 * grad_o, grad_v (3,N) are queued on `stream`; parameter gradients accumulate into the bound buffers. */
int nemit_backward_host(NemitHandle* h, const NemitRays* rays_host, const float* grad_rgb_host, int32_t primal_training,
                        int64_t chunk_rays, float loss_scale, int32_t proposal_requires_grad, float* rgb_host,
                        float* grad_o_host, float* grad_v_host, void* workspace, size_t workspace_bytes, void* stream);
size_t nemit_backward_host_workspace_bytes(const NemitHandle* h, int64_t n);

/* The one collective of the path behind the C ABI (reference: DDP's gradient all-reduce,
 * NS/pipelines/base_pipeline.py:355-357 with the no_sync accumulation of NS/pipelines/mitsuba_sdf.py:1179-1193).
 * nemit_comm_unique_id fills 128 bytes (ncclUniqueId) on one rank; the host broadcasts them (torch.distributed, MPI,
 * a file - the library does not care) and every rank calls nemit_comm_init with its rank.  nemit_allreduce_grads sums
 * `count` floats in place over the ranks with NCCL (NVLink / NVSwitch) and multiplies by `scale` IN THE SAME PASS
 * (ncclRedOpCreatePreMulSum: GradScaler unscale or 1/world folded into the collective, no second sweep over the 77.6 MB).
 * flat_grads: the contiguous gradient blob (NerfEmitter.bind_grads lays every .grad tensor out as a view of one buffer).
 * NCCL is loaded at run time (dlopen of the libnccl.so.2 already in the process or on the loader path); without it
 * nemit_comm_init fails with NEMIT_E_INVALID and says so. */
#define NEMIT_UNIQUE_ID_BYTES 128
int nemit_comm_unique_id(void* id_out_128);
int nemit_comm_init(NemitHandle* h, const void* id_128, int32_t rank, int32_t world_size);
int nemit_comm_destroy(NemitHandle* h);
int nemit_allreduce_grads(NemitHandle* h, float* flat_grads, int64_t count, float scale, void* stream);

/* Ceilings the gather / scatter kernels are measured against, re-measured on the device the handle lives on (bench.py
 * reports every kernel's rate next to the matching one).  kind: NEMIT_MB_GATHER_L1MISS = random 8-byte lane-loads from a
 * 64 MiB L2-resident table (one corner of a hashed level); NEMIT_MB_GATHER_L1HIT = the same from a 64 KiB table;
 * NEMIT_MB_GATHER_COHERENT = the 32 lanes of a warp read 32 consecutive entries (coarse level, ray-coherent samples);
 * NEMIT_MB_RED = random red.global.add.v2.f32 into a 64 MiB table (the hash-table scatter); NEMIT_MB_L2_READ = streaming
 * 128-bit reads of a 32 MiB L2-resident buffer.  result_out[0] = operations (lane-loads / lane-ops; bytes for L2_READ)
 * per second, result_out[1] = kernel milliseconds.  Synchronises the stream.  (The TMA tile::gather4 variant lives in
 * tools/gather_bench.cu: it needs a driver-API tensor map, which this library does not link.) */
#define NEMIT_MB_GATHER_L1MISS 0
#define NEMIT_MB_GATHER_L1HIT 1
#define NEMIT_MB_GATHER_COHERENT 2
#define NEMIT_MB_RED 3
#define NEMIT_MB_L2_READ 4
int nemit_microbench(NemitHandle* h, int32_t kind, double* result_out, void* stream);

/* number of kernels the library has launched on this handle since creation (bench evidence) */
int64_t nemit_launch_count(const NemitHandle* h);

/* Optional per-kernel timing: CUDA events recorded around every kernel launch ON THE LAUNCHING STREAM
 * (the reference's counterpart is the @profiler.time_function wall clock, NS/utils/profiler.py:68-131).
 * nemit_profile_read synchronises the pending events and returns, per kernel kind (arrays of NEMIT_PROFILE_KINDS),
 * accumulated milliseconds, launch counts and the RAYS those launches processed (so that rays per launch is what
 * was launched, not a guess); any of the three may be NULL.  No launch is ever dropped: the event ring holds 4096
 * launches and the oldest records are folded into the sums when it fills.  red_ops_or_null (1): reduction lane-ops
 * (red.global.add lanes) the hash-table scatter of k_field_backward_tc issued since the last reset, counted on the
 * device while profiling is on. */
#define NEMIT_K_SAMPLE 0      /* k_sample_levels */
#define NEMIT_K_FIELD_FWD 1   /* k_field_forward (fp32) */
#define NEMIT_K_FIELD_SEED 2  /* k_field_forward<adjoint seed> */
#define NEMIT_K_FIELD_BWD 3   /* k_field_backward */
#define NEMIT_K_PROP_BWD 4    /* k_prop_backward */
#define NEMIT_K_FINALIZE 5    /* k_finalize_ray_grads */
#define NEMIT_K_FIELD_JVP 6   /* k_field_jvp */
#define NEMIT_K_FIELD_TC 7    /* k_field_forward_tc (tensor-core MLP) */
#define NEMIT_K_FIELD_BWD_TC 8 /* k_field_backward_tc (tensor-core dgrad / wgrad) */
#define NEMIT_PROFILE_KINDS 9
int nemit_profile_enable(NemitHandle* h, int32_t on);
int nemit_profile_read(NemitHandle* h, double* ms_per_kind, int64_t* launches_per_kind, int64_t* rays_per_kind,
                       int64_t* red_ops_or_null, int32_t reset);

#ifdef __cplusplus
}
#endif
#endif /* NEMIT_H_ */
