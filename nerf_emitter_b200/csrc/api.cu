// api.cu - the C ABI of libnerfemitter.so (include/nemit.h): handle, argument checks, chunking,
// workspace carving, kernel launches.  No torch types, no allocation of parameter memory.
#include <mutex>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>

#include "handle.h"
#include "vmf.h"

using namespace nemit;

static thread_local std::string g_err;

int nemit_fail(int code, const char* fmt, const char* a, long long b) {
  char buf[512];
  snprintf(buf, sizeof(buf), fmt, a, b);
  g_err = buf;
  return code;
}
int nemit_cuda_fail(cudaError_t e, const char* where) {
  return nemit_fail(NEMIT_E_CUDA, "CUDA error in %s: ", where), g_err += cudaGetErrorString(e), NEMIT_E_CUDA;
}
void nemit_append_error(const char* more) { g_err += more; }
static int fail(int code, const char* fmt, const char* a = "", long long b = 0) { return nemit_fail(code, fmt, a, b); }
static int cuda_fail(cudaError_t e, const char* where) { return nemit_cuda_fail(e, where); }

// fold the oldest pending launch record into the per-kind sums (blocks until that launch has finished)
static void profile_fold_oldest(NemitHandle* h) {
  const int i = h->ev_head;
  cudaEventSynchronize(h->ev[2 * i + 1]);
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, h->ev[2 * i], h->ev[2 * i + 1]) == cudaSuccess) h->kind_ms[h->ev_kind[i]] += ms;
  h->ev_head = (i + 1) % kEvRing;
  h->ev_count--;
}

// launch with launch counting and optional event bracketing; `rays` = rays this launch processes
template <class F>
static cudaError_t timed_launch(NemitHandle* h, int kind, cudaStream_t st, int64_t rays, F&& f) {
  const bool rec = h->profile && h->ev;
  int slot = 0;
  if (rec) {
    if (h->ev_count == kEvRing) profile_fold_oldest(h);
    slot = (h->ev_head + h->ev_count) % kEvRing;
    if (!h->ev[2 * slot]) { cudaEventCreate(&h->ev[2 * slot]); cudaEventCreate(&h->ev[2 * slot + 1]); }
    cudaEventRecord(h->ev[2 * slot], st);
  }
  cudaError_t e = f();
  if (rec) {
    cudaEventRecord(h->ev[2 * slot + 1], st);
    h->ev_kind[slot] = kind;
    h->ev_count++;
    h->kind_launches[kind]++;
    h->kind_rays[kind] += rays;
  }
  h->launches++;
  return e;
}

extern "C" int nemit_profile_enable(NemitHandle* h, int32_t on) {
  if (!h) return NEMIT_E_INVALID;
  h->profile = on != 0;
  return NEMIT_OK;
}
extern "C" int nemit_profile_read(NemitHandle* h, double* ms_per_kind, int64_t* launches_per_kind, int64_t* rays_per_kind,
                                  int64_t* red_ops_or_null, int32_t reset) {
  if (!h) return NEMIT_E_INVALID;
  DeviceGuard guard(h->device);
  while (h->ev_count > 0) profile_fold_oldest(h);
  for (int k = 0; k < NEMIT_PROFILE_KINDS; ++k) {
    if (ms_per_kind) ms_per_kind[k] = h->kind_ms[k];
    if (launches_per_kind) launches_per_kind[k] = h->kind_launches[k];
    if (rays_per_kind) rays_per_kind[k] = h->kind_rays[k];
    if (reset) { h->kind_ms[k] = 0.0; h->kind_launches[k] = 0; h->kind_rays[k] = 0; }
  }
  if (red_ops_or_null || reset) {
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long c = 0;
    if (e == cudaSuccess) e = cudaMemcpy(&c, h->counters, sizeof(c), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && reset) e = cudaMemset(h->counters, 0, sizeof(unsigned long long) * 8);
    if (e != cudaSuccess) return cuda_fail(e, "nemit_profile_read");
    if (red_ops_or_null) *red_ops_or_null = (int64_t)c;
  }
  return NEMIT_OK;
}

static const int64_t kChunkRays = 1 << 18;
static const int64_t kBwdChunkRays = 1 << 16;   // reverse mode keeps 256 B / sample of encoding tiles in the workspace

extern "C" const char* nemit_last_error(void) { return g_err.c_str(); }
extern "C" int nemit_version(void) { return 100; }

static bool grid_ok(const NemitGrid& g, int levels) {
  return g.num_levels == levels && g.log2_table_size >= 1 && g.log2_table_size <= 26;
}
static void grid_to_dev(const NemitGrid& g, GridDev& d) {
  d.num_levels = g.num_levels;
  d.table_size = 1u << g.log2_table_size;
  d.table_mask = d.table_size - 1u;
  for (int i = 0; i < NEMIT_MAX_LEVELS; ++i) d.scal[i] = i < g.num_levels ? g.scalings[i] : 0.f;
}

extern "C" int nemit_create(const NemitConfig* cfg, int device, NemitHandle** out) {
  if (!cfg || !out) return fail(NEMIT_E_INVALID, "nemit_create: null argument%s");
  if (!grid_ok(cfg->main_grid, kMainLevels) || !grid_ok(cfg->prop_grid[0], kPropLevels) ||
      !grid_ok(cfg->prop_grid[1], kPropLevels))
    return fail(NEMIT_E_INVALID, "nemit_create: this build supports a 16-level main grid and 5-level proposal grids%s");
  if (cfg->num_prop_samples[0] != kS0 || cfg->num_prop_samples[1] != kS1 || cfg->num_nerf_samples != kS2)
    return fail(NEMIT_E_INVALID, "nemit_create: this build supports 256/96/48 samples per ray%s");
  if (cfg->hidden_dim != kHidden || cfg->geo_feat_dim != kGeo || cfg->appearance_dim != kApp ||
      cfg->prop_hidden_dim != kPropHidden)
    return fail(NEMIT_E_INVALID, "nemit_create: this build supports hidden 64 / geo 15 / appearance 32 / proposal hidden 16%s");
  if (cfg->num_images < 1) return fail(NEMIT_E_INVALID, "nemit_create: num_images must be >= 1%s");
  if (cfg->use_fake_contraction)
    return fail(NEMIT_E_INVALID, "nemit_create: use_fake_contraction (FakeContraction + AABBBoxCollider, NS/models/nerfacto.py:144-154,263-270) "
                                 "is not implemented by this build; sdf-nerfacto runs with SceneContraction(order=inf)%s");
  if (cfg->far_intersect)
    return fail(NEMIT_E_INVALID, "nemit_create: far_intersect (AABBBoxFarIntersectCollider in get_point_lights, NS/models/sdf_nerfacto.py:193-194) "
                                 "is not implemented by this build%s");
  if (!cfg->lin0_host || !cfg->u1_host || !cfg->u2_host)
    return fail(NEMIT_E_INVALID, "nemit_create: sampler tables missing%s");
  if (cfg->cpu_sum_width != 0 && cfg->cpu_sum_width != 8 && cfg->cpu_sum_width != 16)
    return fail(NEMIT_E_INVALID, "nemit_create: cpu_sum_width must be 0, 8 or 16%s");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
  if (device < 0 || device >= ndev) return fail(NEMIT_E_INVALID, "nemit_create: no such CUDA device%s");
  DeviceGuard guard(device);
  NemitHandle* h = new (std::nothrow) NemitHandle();
  if (!h) return fail(NEMIT_E_INVALID, "nemit_create: out of host memory%s");
  memset(h, 0, sizeof(*h));
  h->device = device;
  h->use_average_embedding = 1;       // NerfactoModelConfig.use_average_appearance_embedding (NS/models/nerfacto.py:116)
  h->cfg = *cfg;
  h->cfg.lin0_host = h->cfg.u1_host = h->cfg.u2_host = nullptr;
  grid_to_dev(cfg->main_grid, h->M.main_grid);
  grid_to_dev(cfg->prop_grid[0], h->M.prop_grid[0]);
  grid_to_dev(cfg->prop_grid[1], h->M.prop_grid[1]);
  for (int k = 0; k < 3; ++k) { h->M.aabb_min[k] = cfg->aabb[k]; h->M.aabb_max[k] = cfg->aabb[3 + k]; }
  h->M.scene_scale = cfg->scene_scale;
  h->M.near_plane = cfg->near_plane;
  h->M.far_plane = cfg->far_plane;
  h->M.rgb_bias = cfg->rgb_bias;
  h->M.disable_inside_aabb = cfg->disable_inside_aabb;
  h->M.cpu_sum_width = cfg->cpu_sum_width;
  h->M.num_images = cfg->num_images;
  h->M.interlevel_mult = cfg->interlevel_mult;
  h->M.distortion_mult = cfg->distortion_mult;
  const int n0 = kS0 + 1, n1 = kS1 + 1, n2 = kS2 + 1;
  if ((e = cudaMalloc(&h->tables_dev, sizeof(float) * (n0 + n1 + n2 + 17))) != cudaSuccess) {
    delete h;
    return cuda_fail(e, "cudaMalloc(sampler tables)");
  }
  cudaMemcpy(h->tables_dev, cfg->lin0_host, sizeof(float) * n0, cudaMemcpyHostToDevice);
  cudaMemcpy(h->tables_dev + n0, cfg->u1_host, sizeof(float) * n1, cudaMemcpyHostToDevice);
  e = cudaMemcpy(h->tables_dev + n0 + n1, cfg->u2_host, sizeof(float) * n2, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cudaFree(h->tables_dev); delete h; return cuda_fail(e, "cudaMemcpy(sampler tables)"); }
  h->lin0 = h->tables_dev;
  h->u1 = h->tables_dev + n0;
  h->u2 = h->tables_dev + n0 + n1;
  h->opt_stats = reinterpret_cast<double*>(h->tables_dev + ((n0 + n1 + n2 + 1) & ~1));   // 8-byte aligned tail
  h->ev = static_cast<cudaEvent_t*>(calloc(2 * kEvRing, sizeof(cudaEvent_t)));
  h->ev_kind = static_cast<int*>(calloc(kEvRing, sizeof(int)));
  for (int i = 0; i < 2 && e == cudaSuccess; ++i)
    e = cudaMalloc(&h->prop_half[i], sizeof(uint32_t) * ((size_t)cfg->prop_grid[i].num_levels << cfg->prop_grid[i].log2_table_size));
  if (e == cudaSuccess) e = cudaMalloc(&h->counters, sizeof(unsigned long long) * 8);
  if (e == cudaSuccess) e = cudaMemset(h->counters, 0, sizeof(unsigned long long) * 8);
  if (e != cudaSuccess || !h->ev || !h->ev_kind) {
    cudaFree(h->tables_dev); cudaFree(h->counters); cudaFree(h->prop_half[0]); cudaFree(h->prop_half[1]);
    free(h->ev); free(h->ev_kind); delete h;
    return e != cudaSuccess ? cuda_fail(e, "cudaMalloc(counters)") : fail(NEMIT_E_INVALID, "nemit_create: out of host memory%s");
  }
  *out = h;
  return NEMIT_OK;
}

extern "C" int nemit_destroy(NemitHandle* h) {
  if (!h) return NEMIT_OK;
  DeviceGuard guard(h->device);
  nemit_comm_destroy(h);
  cudaFree(h->tables_dev);
  cudaFree(h->counters);
  cudaFree(h->prop_half[0]);
  cudaFree(h->prop_half[1]);
  free(h->rotation_ids);
  free(h->rot_transforms);
  if (h->ev)
    for (int i = 0; i < 2 * kEvRing; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  free(h->ev);
  free(h->ev_kind);
  delete h;
  return NEMIT_OK;
}

extern "C" int nemit_set_appearance_mode(NemitHandle* h, int32_t always_use_camera_indices,
                                         int32_t use_average_appearance_embedding, const int32_t* rotation_ids_host,
                                         int32_t num_rotation_ids) {
  if (!h) return fail(NEMIT_E_INVALID, "nemit_set_appearance_mode: null handle%s");
  if (num_rotation_ids < 0 || (num_rotation_ids > 0 && !rotation_ids_host))
    return fail(NEMIT_E_INVALID, "nemit_set_appearance_mode: bad rotation id table%s");
  for (int32_t i = 0; i < num_rotation_ids; ++i)
    if (rotation_ids_host[i] < 0 || rotation_ids_host[i] >= h->cfg.num_images)
      return fail(NEMIT_E_INVALID, "nemit_set_appearance_mode: rotation id outside the appearance embedding%s");
  int32_t* ids = nullptr;
  if (num_rotation_ids > 0) {
    ids = static_cast<int32_t*>(malloc(sizeof(int32_t) * num_rotation_ids));
    if (!ids) return fail(NEMIT_E_INVALID, "nemit_set_appearance_mode: out of host memory%s");
    memcpy(ids, rotation_ids_host, sizeof(int32_t) * num_rotation_ids);
  }
  free(h->rotation_ids);
  h->rotation_ids = ids;
  h->num_rotation_ids = num_rotation_ids;
  h->always_use_camera_indices = always_use_camera_indices != 0;
  h->use_average_embedding = use_average_appearance_embedding != 0;
  return NEMIT_OK;
}

extern "C" int nemit_set_rotater(NemitHandle* h, int32_t num_rotations, const float* transforms_host, int32_t bound_type,
                                 const float* bound_host) {
  if (!h) return fail(NEMIT_E_INVALID, "nemit_set_rotater: null handle%s");
  if (num_rotations < 0) return fail(NEMIT_E_INVALID, "nemit_set_rotater: negative rotation count%s");
  if (num_rotations == 0) {
    free(h->rot_transforms);
    h->rot_transforms = nullptr; h->num_rotations = 0; h->rot_mode = 0;
    return NEMIT_OK;
  }
  if (!transforms_host || !bound_host) return fail(NEMIT_E_INVALID, "nemit_set_rotater: null argument%s");
  if (bound_type != NEMIT_ROT_SPHERE && bound_type != NEMIT_ROT_BOX) return fail(NEMIT_E_INVALID, "nemit_set_rotater: unknown bound type%s");
  float* t = static_cast<float*>(malloc(sizeof(float) * 12 * (size_t)num_rotations));
  if (!t) return fail(NEMIT_E_INVALID, "nemit_set_rotater: out of host memory%s");
  memcpy(t, transforms_host, sizeof(float) * 12 * (size_t)num_rotations);
  free(h->rot_transforms);
  h->rot_transforms = t; h->num_rotations = num_rotations; h->rot_mode = bound_type;
  for (int k = 0; k < 6; ++k) h->rot_bound[k] = k < (bound_type == NEMIT_ROT_SPHERE ? 4 : 6) ? bound_host[k] : 0.f;
  return NEMIT_OK;
}

static bool params_complete(const NemitParams* p) {
  const float* const* q = reinterpret_cast<const float* const*>(p);
  for (size_t i = 0; i < sizeof(NemitParams) / sizeof(float*); ++i)
    if (!q[i]) return false;
  return true;
}

extern "C" int nemit_bind_params(NemitHandle* h, const NemitParams* p) {
  if (!h || !p) return fail(NEMIT_E_INVALID, "nemit_bind_params: null argument%s");
  if (!params_complete(p)) return fail(NEMIT_E_INVALID, "nemit_bind_params: a parameter pointer is null%s");
  // table rows are fetched as aligned 16-byte pairs, MLP tensors as float4 streams by the optimizer step
  if (((uintptr_t)p->main_table | (uintptr_t)p->prop_table[0] | (uintptr_t)p->prop_table[1]) & 15)
    return fail(NEMIT_E_INVALID, "nemit_bind_params: hash tables must be 16-byte aligned%s");
  h->P = *p;
  h->params_bound = true;
  return NEMIT_OK;
}
extern "C" int nemit_bind_grads(NemitHandle* h, const NemitParams* g) {
  if (!h || !g) return fail(NEMIT_E_INVALID, "nemit_bind_grads: null argument%s");
  if (!params_complete(g)) return fail(NEMIT_E_INVALID, "nemit_bind_grads: a gradient pointer is null%s");
  // the table scatter uses red.global.add.v4.f32 on aligned pairs of rows
  if (((uintptr_t)g->main_table | (uintptr_t)g->prop_table[0] | (uintptr_t)g->prop_table[1]) & 15)
    return fail(NEMIT_E_INVALID, "nemit_bind_grads: hash-table gradients must be 16-byte aligned%s");
  h->G = *g;
  h->grads_bound = true;
  return NEMIT_OK;
}
extern "C" int64_t nemit_launch_count(const NemitHandle* h) { return h ? h->launches : 0; }
extern "C" int nemit_set_sampler(NemitHandle* h, int32_t sampler) {
  if (!h) return fail(NEMIT_E_INVALID, "nemit_set_sampler: null handle%s");
  if (sampler != NEMIT_SAMPLER_EXACT && sampler != NEMIT_SAMPLER_FAST)
    return fail(NEMIT_E_INVALID, "nemit_set_sampler: unknown sampler%s");
  h->sampler = sampler;
  return NEMIT_OK;
}
extern "C" int nemit_set_precision(NemitHandle* h, int32_t precision) {
  if (!h) return fail(NEMIT_E_INVALID, "nemit_set_precision: null handle%s");
  if (precision != NEMIT_PRECISION_FP32 && precision != NEMIT_PRECISION_FP16_TC && precision != NEMIT_PRECISION_SPLIT_TC)
    return fail(NEMIT_E_INVALID, "nemit_set_precision: unknown precision%s");
  h->precision = precision;
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------------------
static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

struct WsLayout {
  size_t t_edges2, s_edges2, s_edges1, host_stage, bwd, total;
  size_t g_sigma, g_rgbs, weights2, ray_grad, g_absmax, enc_tiles, jac_tiles, act_tiles, seeds4, masks4;   // inside the backward block
};
static size_t backward_chunk_bytes(int64_t c, WsLayout* L, size_t base) {
  size_t off = base;
  if (L) L->g_sigma = off;
  off += align_up(sizeof(float) * (size_t)c * kS2);
  if (L) L->g_rgbs = off;
  off += align_up(sizeof(float) * (size_t)c * kS2 * 3);
  if (L) L->weights2 = off;
  off += align_up(sizeof(float) * (size_t)c * kS2);
  if (L) L->ray_grad = off;
  off += align_up(sizeof(float) * (size_t)c * 6);
  if (L) L->g_absmax = off;
  off += align_up(sizeof(float));
  // tensor-core backward: per sample of the backward sub-chunk the encoding row (64 B), its Jacobian (192 B), the
  // operand rows H1 / IN (192 B) and the pre-activation seeds (16 B), kept by the seed pass (H2 / H3 are recomputed)
  const size_t bc = (size_t)(c < kBwdChunkRays ? c : kBwdChunkRays);
  // (+8 tiles of slack: a launch may write up to 3 tiles past its rays, and the split-operand mode runs two launches of
  // half the rays with tiles of twice the size in the same space, see bwd_sub_chunk)
  const size_t tiles = (bc * kS2 + 127) / 128 + 8;
  if (L) L->enc_tiles = off;
  off += align_up(tiles * (size_t)kEncTileBytes);
  if (L) L->jac_tiles = off;
  off += align_up(tiles * (size_t)kJacTileBytes);
  if (L) L->act_tiles = off;
  off += align_up(tiles * (size_t)kActTileBytes);
  if (L) L->seeds4 = off;
  off += align_up(tiles * 128 * 16);
  if (L) L->masks4 = off;
  off += align_up(tiles * 128 * 16);
  return off - base;
}
// rays per reverse-mode launch group: the kept tiles of NEMIT_PRECISION_SPLIT_TC are twice as large (bf16 hi | lo pieces,
// fp32 Jacobian), so that mode walks the chunk in sub-chunks of half the rays inside the same workspace
static int64_t bwd_sub_chunk(const NemitHandle* h, int64_t n_total) {
  int64_t bc = n_total < kBwdChunkRays ? n_total : kBwdChunkRays;
  if (h->precision == NEMIT_PRECISION_SPLIT_TC) bc = bc / 2 > 0 ? bc / 2 : 1;
  return bc;
}
static int tc_pieces(const NemitHandle* h) { return h->precision == NEMIT_PRECISION_SPLIT_TC ? 2 : 1; }
static bool tc_mode(const NemitHandle* h) { return h->precision != NEMIT_PRECISION_FP32; }
static WsLayout ws_layout(int64_t n) {
  int64_t c = n < kChunkRays ? n : kChunkRays;
  if (c < 1) c = 1;
  WsLayout L;
  size_t off = 0;
  L.t_edges2 = off; off += align_up(sizeof(float) * (size_t)c * (kS2 + 1));
  L.s_edges2 = off; off += align_up(sizeof(float) * (size_t)c * (kS2 + 1));
  L.s_edges1 = off; off += align_up(sizeof(float) * (size_t)(c < kBwdChunkRays ? c : kBwdChunkRays) * (kS1 + 1));   // backward sub-chunk
  L.bwd = off;      off += backward_chunk_bytes(c, &L, off);
  L.host_stage = off; off += align_up(sizeof(float) * (size_t)(n < 1 ? 1 : n) * 13);   // o,v (6) near (1) jitter (3) rgb (3)
  L.total = off;
  return L;
}

extern "C" size_t nemit_workspace_bytes(const NemitHandle* h, int64_t n) {
  (void)h;
  return ws_layout(n).total;
}

// Which appearance vector a call sees (NS/fields/nerfacto_field.py:272-285, NS/models/sdf_nerfacto.py:71-72,
// NS/field_components/rotater.py:94-100): a row of the embedding in training mode or when the model was built with a
// Rotater (always_use_camera_indices, NS/models/nerfacto.py:186-206), after camera -> rotation id mapping; else
// the mean of the rows (use_average_appearance_embedding) or zeros.  Returns false when the row is out of range.
static bool appearance_row(const NemitHandle* h, int training, int camera_idx, bool map_ids, int* row) {
  if (!(training || h->always_use_camera_indices)) {
    *row = h->use_average_embedding ? kAppMean : kAppZero;
    return true;
  }
  int idx = camera_idx;
  if (map_ids && h->rotation_ids) {
    if (idx < 0) return false;
    if (idx > h->num_rotation_ids - 1) idx = h->num_rotation_ids - 1;      // camera_indices.clamp(max = len - 1)
    idx = h->rotation_ids[idx];
  }
  *row = idx;
  return idx >= 0 && idx < h->cfg.num_images;
}

// rotation id of a call: Rotater.map_rotation_ids(camera_idx) (rotater.py:94-100), -1 when it has no transform
static int rotation_id(const NemitHandle* h, int camera_idx) {
  int idx = camera_idx;
  if (h->rotation_ids) {
    if (idx < 0) return -1;
    if (idx > h->num_rotation_ids - 1) idx = h->num_rotation_ids - 1;
    idx = h->rotation_ids[idx];
  }
  return (idx >= 0 && idx < h->num_rotations) ? idx : -1;
}

static int check_rays(const NemitHandle* h, const NemitRays* r, const char* who) {
  if (!h || !r) return fail(NEMIT_E_INVALID, "%s: null argument", who);
  if (!h->params_bound) return fail(NEMIT_E_UNBOUND, "%s: parameters not bound (nemit_bind_params)", who);
  if (r->n < 0) return fail(NEMIT_E_INVALID, "%s: negative ray count", who);
  if (r->n > 0 && (!r->o || !r->v)) return fail(NEMIT_E_INVALID, "%s: null ray pointers", who);
  int row;
  if (!appearance_row(h, r->training, r->camera_idx, true, &row))
    return fail(NEMIT_E_INVALID, "%s: camera_idx outside the appearance embedding", who);
  if (h && h->rot_mode && rotation_id(h, r->camera_idx) < 0)
    return fail(NEMIT_E_INVALID, "%s: camera_idx has no rotation transform (nemit_set_rotater)", who);
  if (!(r->anneal > 0.f)) return fail(NEMIT_E_INVALID, "%s: anneal must be > 0", who);
  if (r->layout != NEMIT_RAYS_MITSUBA_3N && r->layout != NEMIT_RAYS_NERF_N3)
    return fail(NEMIT_E_INVALID, "%s: unknown ray layout", who);
  if ((r->kept_t_edges2 != nullptr) != (r->kept_s_edges2 != nullptr) || (r->kept_s_edges1 && !r->kept_t_edges2))
    return fail(NEMIT_E_INVALID, "%s: kept_t_edges2 and kept_s_edges2 come together (kept_s_edges1 only with them)", who);
  return NEMIT_OK;
}

// The final samples of rays [off, off + n) of a reverse- / forward-mode call: the edges kept from the primal of the same
// CustomOp instance (NemitRays.kept_*), or a proposal sampling pass into the workspace.
struct ChunkEdges {
  const float* t_edges2;
  const float* s_edges2;
  const float* s_edges1;   // only when asked for
};
static int chunk_edges(NemitHandle* h, const ModelDev& M, const NemitRays* rays, int64_t off, int64_t n, char* ws,
                       const struct WsLayout& L, bool need_s1, cudaStream_t st, ChunkEdges* out);

// rotate: whether THIS stage's samples go through Rotater.apply_frustums (see nemit_set_rotater in nemit.h)
static void fill_ray_args(const NemitHandle* h, RayArgs& A, const NemitRays* r, int64_t off, int64_t n, bool rotate = false) {
  A.rot.mode = 0;
  if (rotate && h->rot_mode) {
    const int id = rotation_id(h, r->camera_idx);
    if (id >= 0) {
      A.rot.mode = h->rot_mode;
      memcpy(A.rot.m, h->rot_transforms + 12 * (size_t)id, sizeof(float) * 12);
      memcpy(A.rot.b, h->rot_bound, sizeof(float) * 6);
    }
  }
  A.o = r->o; A.v = r->v; A.near_per_ray = r->near_per_ray; A.near_scalar = r->near_scalar;
  A.n_total = r->n; A.ray_offset = off; A.n = n; A.training = r->training;
  appearance_row(h, r->training, r->camera_idx, true, &A.app_row);
  A.jitter = r->jitter; A.anneal = r->anneal; A.nerf_layout = r->layout == NEMIT_RAYS_NERF_N3;
}

static void fill_sample_args(const NemitHandle* h, SampleArgs& S, const NemitRays* r, int64_t off, int64_t n) {
  memset(&S, 0, sizeof(S));
  fill_ray_args(h, S, r, off, n, true);          // every proposal level's samples are rotated
  for (int i = 0; i < 2; ++i) {
    S.prop_table[i] = reinterpret_cast<const float2*>(h->P.prop_table[i]);
    S.prop_w0[i] = h->P.prop_w0[i]; S.prop_b0[i] = h->P.prop_b0[i];
    S.prop_w1[i] = h->P.prop_w1[i]; S.prop_b1[i] = h->P.prop_b1[i];
    S.prop_table_h[i] = h->prop_half[i];
  }
  S.lin0 = h->lin0; S.u1 = h->u1; S.u2 = h->u2;
}

static void fill_field_params(const NemitParams& P, FieldParams& F) {
  F.main_table = reinterpret_cast<const float2*>(P.main_table);
  F.base_w0 = P.base_w0; F.base_b0 = P.base_b0; F.base_w1 = P.base_w1; F.base_b1 = P.base_b1;
  F.head_w0 = P.head_w0; F.head_b0 = P.head_b0; F.head_w1 = P.head_w1; F.head_b1 = P.head_b1;
  F.head_w2 = P.head_w2; F.head_b2 = P.head_b2; F.appearance = P.appearance;
}

static void offset_levels(SampleArgs& S, const NemitLevelsOut* L, int64_t off) {
  if (!L) return;
  if (L->weights0) S.weights0 = L->weights0 + off * kS0;
  if (L->weights1) S.weights1 = L->weights1 + off * kS1;
  if (L->s_edges1) S.s_edges1 = L->s_edges1 + off * (kS1 + 1);
  if (L->s_edges2) S.s_edges2 = L->s_edges2 + off * (kS2 + 1);
  if (L->t_edges2) S.t_edges2 = L->t_edges2 + off * (kS2 + 1);
  if (L->inds1) S.inds1 = L->inds1 + off * (kS1 + 1);
  if (L->inds2) S.inds2 = L->inds2 + off * (kS2 + 1);
}

// the proposal sampling pass of one chunk (the fast sampler first refreshes its packed-half table copies: one more launch)
static cudaError_t launch_sampler(NemitHandle* h, const ModelDev& M, const SampleArgs& S, cudaStream_t st, int64_t n) {
  const bool fast = h->sampler == NEMIT_SAMPLER_FAST;
  if (fast) h->launches++;
  return timed_launch(h, NEMIT_K_SAMPLE, st, n, [&] { return launch_sample_levels(M, S, st, fast); });
}

static int chunk_edges(NemitHandle* h, const ModelDev& M, const NemitRays* rays, int64_t off, int64_t n, char* ws,
                       const WsLayout& L, bool need_s1, cudaStream_t st, ChunkEdges* out) {
  if (rays->kept_t_edges2 && (!need_s1 || rays->kept_s_edges1)) {
    out->t_edges2 = rays->kept_t_edges2 + off * (kS2 + 1);
    out->s_edges2 = rays->kept_s_edges2 + off * (kS2 + 1);
    out->s_edges1 = need_s1 ? rays->kept_s_edges1 + off * (kS1 + 1) : nullptr;
    return NEMIT_OK;
  }
  SampleArgs S;
  fill_sample_args(h, S, rays, off, n);
  S.t_edges2 = reinterpret_cast<float*>(ws + L.t_edges2);
  S.s_edges2 = reinterpret_cast<float*>(ws + L.s_edges2);
  S.s_edges1 = need_s1 ? reinterpret_cast<float*>(ws + L.s_edges1) : nullptr;
  cudaError_t e = launch_sampler(h, M, S, st, n);
  if (e != cudaSuccess) return cuda_fail(e, "k_sample_levels");
  out->t_edges2 = S.t_edges2; out->s_edges2 = S.s_edges2; out->s_edges1 = S.s_edges1;
  return NEMIT_OK;
}

extern "C" int nemit_sample_levels(NemitHandle* h, const NemitRays* rays, const NemitLevelsOut* out,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  (void)workspace; (void)workspace_bytes;
  int rc = check_rays(h, rays, "nemit_sample_levels");
  if (rc) return rc;
  if (!out) return fail(NEMIT_E_INVALID, "nemit_sample_levels: null outputs%s");
  DeviceGuard guard(h->device);
  for (int64_t off = 0; off < rays->n; off += kChunkRays) {
    int64_t n = rays->n - off < kChunkRays ? rays->n - off : kChunkRays;
    SampleArgs S;
    fill_sample_args(h, S, rays, off, n);
    offset_levels(S, out, off);
    cudaError_t e = launch_sampler(h, h->M, S, (cudaStream_t)stream, n);
    if (e != cudaSuccess) return cuda_fail(e, "k_sample_levels");
  }
  return NEMIT_OK;
}

extern "C" int nemit_forward(NemitHandle* h, const NemitRays* rays, float* rgb, float* accumulation,
                             const NemitLevelsOut* levels, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_rays(h, rays, "nemit_forward");
  if (rc) return rc;
  if (rays->n == 0) return NEMIT_OK;
  if (!rgb) return fail(NEMIT_E_INVALID, "nemit_forward: null rgb output%s");
  WsLayout L = ws_layout(rays->n);
  if (!workspace || workspace_bytes < L.total)
    return fail(NEMIT_E_WORKSPACE, "nemit_forward: workspace too small%s (need %lld bytes)", "", (long long)L.total);
  DeviceGuard guard(h->device);
  char* ws = static_cast<char*>(workspace);
  for (int64_t off = 0; off < rays->n; off += kChunkRays) {
    int64_t n = rays->n - off < kChunkRays ? rays->n - off : kChunkRays;
    SampleArgs S;
    fill_sample_args(h, S, rays, off, n);
    offset_levels(S, levels, off);
    float* t_edges = reinterpret_cast<float*>(ws + L.t_edges2);
    float* user_t = S.t_edges2;
    S.t_edges2 = t_edges;
    cudaError_t e = launch_sampler(h, h->M, S, (cudaStream_t)stream, n);
    if (e != cudaSuccess) return cuda_fail(e, "k_sample_levels");
    if (user_t) {
      e = cudaMemcpyAsync(user_t, t_edges, sizeof(float) * n * (kS2 + 1), cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
      if (e != cudaSuccess) return cuda_fail(e, "copy t_edges2");
    }
    FieldArgs F;
    memset(&F, 0, sizeof(F));
    fill_ray_args(h, F, rays, off, n, true);
    fill_field_params(h->P, F.P);
    F.t_edges2 = t_edges;
    F.rgb = rgb + off * 3;
    F.accumulation = accumulation ? accumulation + off : nullptr;
    F.weights2 = (levels && levels->weights2) ? levels->weights2 + off * kS2 : nullptr;
    if (tc_mode(h))
      e = timed_launch(h, NEMIT_K_FIELD_TC, (cudaStream_t)stream, n, [&] { return launch_field_forward_tc(h->M, F, (cudaStream_t)stream, tc_pieces(h)); });
    else
      e = timed_launch(h, NEMIT_K_FIELD_FWD, (cudaStream_t)stream, n, [&] { return launch_field_forward(h->M, F, (cudaStream_t)stream); });
    if (e != cudaSuccess) return cuda_fail(e, "k_field_forward");
  }
  return NEMIT_OK;
}

extern "C" int nemit_hash_encode(NemitHandle* h, int32_t which, const float* x, int64_t n, float* out, void* stream) {
  if (!h || (n > 0 && (!x || !out))) return fail(NEMIT_E_INVALID, "nemit_hash_encode: null argument%s");
  if (!h->params_bound) return fail(NEMIT_E_UNBOUND, "nemit_hash_encode: parameters not bound%s");
  if (which < 0 || which > 2) return fail(NEMIT_E_INVALID, "nemit_hash_encode: which must be 0, 1 or 2%s");
  DeviceGuard guard(h->device);
  const GridDev& G = which == 0 ? h->M.main_grid : h->M.prop_grid[which - 1];
  const float* t = which == 0 ? h->P.main_table : h->P.prop_table[which - 1];
  cudaError_t e = launch_hash_encode(G, reinterpret_cast<const float2*>(t), x, n, out, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_hash_encode");
  h->launches++;
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// component-level drop-ins (components.cu)
// ---------------------------------------------------------------------------------------------
extern "C" int nemit_hash_encode_grid(const NemitGrid* grid, const float* table, const float* x, int64_t n, float* out,
                                      void* stream) {
  if (!grid || n < 0 || (n > 0 && (!table || !x || !out))) return fail(NEMIT_E_INVALID, "nemit_hash_encode_grid: null argument%s");
  if (grid->num_levels < 1 || grid->num_levels > NEMIT_MAX_LEVELS || grid->log2_table_size < 1 || grid->log2_table_size > 31)
    return fail(NEMIT_E_INVALID, "nemit_hash_encode_grid: 1..16 levels and 2^1..2^31 table entries per level%s");
  GridDev G;
  grid_to_dev(*grid, G);
  cudaError_t e = launch_hash_encode(G, reinterpret_cast<const float2*>(table), x, n, out, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_hash_encode");
  return NEMIT_OK;
}

extern "C" int nemit_sh_encode(const float* d, int64_t n, int32_t levels, float* out, void* stream) {
  if (n < 0 || (n > 0 && (!d || !out))) return fail(NEMIT_E_INVALID, "nemit_sh_encode: null argument%s");
  if (levels < 1 || levels > 4) return fail(NEMIT_E_INVALID, "nemit_sh_encode: levels must be in 1..4%s");
  cudaError_t e = launch_sh_encode(d, n, levels, out, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_sh_encode");
  return NEMIT_OK;
}

extern "C" int nemit_emitter_local_rays(const float* world_to_local_3x4, const float* bbox_min, const float* bbox_max,
                                        float skip_distance, int32_t outside_only, int32_t linear_only, const float* o,
                                        const float* v, int64_t n, float* o_out, float* v_out, float* near_out,
                                        void* stream) {
  if (!world_to_local_3x4 || n < 0 || (n > 0 && (!o || !v || !o_out || !v_out)))
    return fail(NEMIT_E_INVALID, "nemit_emitter_local_rays: null argument%s");
  if (outside_only && (!bbox_min || !bbox_max))
    return fail(NEMIT_E_INVALID, "nemit_emitter_local_rays: outside_only needs the bounding box%s");
  const float zero[3] = {0.f, 0.f, 0.f};
  cudaError_t e = launch_emitter_local_rays(world_to_local_3x4, bbox_min ? bbox_min : zero, bbox_max ? bbox_max : zero,
                                            skip_distance, outside_only != 0, linear_only != 0, o, v, n, o_out, v_out,
                                            near_out, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_emitter_local_rays");
  return NEMIT_OK;
}

extern "C" int nemit_mlp(const NemitMlp* mlp, const float* x, int64_t n, float* out, void* stream) {
  if (!mlp || n < 0 || (n > 0 && (!x || !out))) return fail(NEMIT_E_INVALID, "nemit_mlp: null argument%s");
  if (mlp->num_layers < 1 || mlp->num_layers > NEMIT_MLP_MAX_LAYERS || mlp->in_dim < 1 || mlp->in_dim > NEMIT_MLP_MAX_WIDTH)
    return fail(NEMIT_E_INVALID, "nemit_mlp: 1..8 layers, widths 1..64%s");
  MlpDev L;
  memset(&L, 0, sizeof(L));
  L.in_dim = mlp->in_dim; L.num_layers = mlp->num_layers;
  L.activation = mlp->activation; L.out_activation = mlp->out_activation;
  if ((L.activation | L.out_activation) & ~1) return fail(NEMIT_E_INVALID, "nemit_mlp: activation must be 0 (none) or 1 (ReLU)%s");
  for (int l = 0; l < mlp->num_layers; ++l) {
    if (mlp->out_dim[l] < 1 || mlp->out_dim[l] > NEMIT_MLP_MAX_WIDTH) return fail(NEMIT_E_INVALID, "nemit_mlp: 1..8 layers, widths 1..64%s");
    if (!mlp->weight[l]) return fail(NEMIT_E_INVALID, "nemit_mlp: null weight%s");
    L.out_dim[l] = mlp->out_dim[l]; L.weight[l] = mlp->weight[l]; L.bias[l] = mlp->bias[l];
  }
  cudaError_t e = launch_mlp(L, x, n, out, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_mlp");
  return NEMIT_OK;
}

extern "C" int nemit_density_fn(NemitHandle* h, int32_t which, const float* positions, int64_t n, float* density,
                                float* geo_or_null, void* stream) {
  if (!h || n < 0 || (n > 0 && (!positions || !density))) return fail(NEMIT_E_INVALID, "nemit_density_fn: null argument%s");
  if (!h->params_bound) return fail(NEMIT_E_UNBOUND, "nemit_density_fn: parameters not bound%s");
  if (which < 0 || which > 2) return fail(NEMIT_E_INVALID, "nemit_density_fn: which must be 0, 1 or 2%s");
  DeviceGuard guard(h->device);
  cudaError_t e;
  if (which == 0) {
    FieldParams P;
    fill_field_params(h->P, P);
    e = launch_field_samples(h->M, P, positions, nullptr, n, kAppMean, density, geo_or_null, nullptr, (cudaStream_t)stream);
  } else {
    const int i = which - 1;
    e = launch_prop_density(h->M, h->M.prop_grid[i], reinterpret_cast<const float2*>(h->P.prop_table[i]), h->P.prop_w0[i],
                            h->P.prop_b0[i], h->P.prop_w1[i], h->P.prop_b1[i], positions, n, density, (cudaStream_t)stream);
  }
  if (e != cudaSuccess) return cuda_fail(e, "nemit_density_fn");
  h->launches++;
  return NEMIT_OK;
}

extern "C" int nemit_field_samples(NemitHandle* h, const float* positions, const float* directions, int64_t n,
                                   int32_t camera_idx, int32_t training, float* density, float* rgb, void* stream) {
  if (!h || n < 0 || (n > 0 && (!positions || !directions || !density || !rgb)))
    return fail(NEMIT_E_INVALID, "nemit_field_samples: null argument%s");
  if (!h->params_bound) return fail(NEMIT_E_UNBOUND, "nemit_field_samples: parameters not bound%s");
  int app_row;      // the model has already mapped camera -> rotation ids on ray_samples (sdf_nerfacto.py:71-72)
  if (!appearance_row(h, training, camera_idx, false, &app_row))
    return fail(NEMIT_E_INVALID, "nemit_field_samples: camera_idx outside the appearance embedding%s");
  DeviceGuard guard(h->device);
  FieldParams P;
  fill_field_params(h->P, P);
  cudaError_t e = launch_field_samples(h->M, P, positions, directions, n, app_row, density, nullptr, rgb,
                                       (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_field_samples");
  h->launches++;
  return NEMIT_OK;
}

extern "C" int nemit_forward_host(NemitHandle* h, const NemitRays* rh, float* rgb_host, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  int rc = check_rays(h, rh, "nemit_forward_host");
  if (rc) return rc;
  if (rh->n == 0) return NEMIT_OK;
  if (!rgb_host) return fail(NEMIT_E_INVALID, "nemit_forward_host: null rgb output%s");
  WsLayout L = ws_layout(rh->n);
  if (!workspace || workspace_bytes < L.total)
    return fail(NEMIT_E_WORKSPACE, "nemit_forward_host: workspace too small%s (need %lld bytes)", "", (long long)L.total);
  DeviceGuard guard(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = rh->n;
  float* stage = reinterpret_cast<float*>(static_cast<char*>(workspace) + L.host_stage);
  float *d_o = stage, *d_v = stage + 3 * n, *d_near = stage + 6 * n, *d_jit = stage + 7 * n, *d_rgb = stage + 10 * n;
  cudaError_t e;
  if ((e = cudaMemcpyAsync(d_o, rh->o, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D o");
  if ((e = cudaMemcpyAsync(d_v, rh->v, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D v");
  NemitRays rd = *rh;
  rd.o = d_o; rd.v = d_v;
  if (rh->near_per_ray) {
    if ((e = cudaMemcpyAsync(d_near, rh->near_per_ray, sizeof(float) * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D near");
    rd.near_per_ray = d_near;
  }
  if (rh->jitter) {
    if ((e = cudaMemcpyAsync(d_jit, rh->jitter, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D jitter");
    rd.jitter = d_jit;
  }
  rc = nemit_forward(h, &rd, d_rgb, nullptr, nullptr, workspace, workspace_bytes, stream);
  if (rc) return rc;
  if ((e = cudaMemcpyAsync(rgb_host, d_rgb, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return cuda_fail(e, "D2H rgb");
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// forward-mode and reverse-mode entries
// ---------------------------------------------------------------------------------------------
extern "C" int nemit_jvp(NemitHandle* h, const NemitRays* rays, const float* d_o, const float* d_v, float* d_rgb,
                         void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_rays(h, rays, "nemit_jvp");
  if (rc) return rc;
  if (rays->n == 0) return NEMIT_OK;
  if (!d_o || !d_v || !d_rgb) return fail(NEMIT_E_INVALID, "nemit_jvp: null tangent pointers%s");
  WsLayout L = ws_layout(rays->n);
  if (!workspace || workspace_bytes < L.total)
    return fail(NEMIT_E_WORKSPACE, "nemit_jvp: workspace too small%s (need %lld bytes)", "", (long long)L.total);
  DeviceGuard guard(h->device);
  char* ws = static_cast<char*>(workspace);
  cudaStream_t st = (cudaStream_t)stream;
  for (int64_t off = 0; off < rays->n; off += kChunkRays) {
    int64_t n = rays->n - off < kChunkRays ? rays->n - off : kChunkRays;
    ChunkEdges E;
    if ((rc = chunk_edges(h, h->M, rays, off, n, ws, L, false, st, &E))) return rc;
    cudaError_t e;
    FieldArgs F;
    memset(&F, 0, sizeof(F));
    fill_ray_args(h, F, rays, off, n, true);
    fill_field_params(h->P, F.P);
    F.t_edges2 = E.t_edges2;
    F.d_o = d_o; F.d_v = d_v;
    F.d_rgb = d_rgb + off * 3;
    e = timed_launch(h, NEMIT_K_FIELD_JVP, st, n, [&] { return launch_field_jvp(h->M, F, st); });
    if (e != cudaSuccess) return cuda_fail(e, "k_field_jvp");
  }
  return NEMIT_OK;
}

extern "C" int nemit_backward(NemitHandle* h, const NemitRays* rays, const float* grad_rgb, int64_t chunk_rays,
                              float loss_scale, int32_t proposal_requires_grad, float* grad_o, float* grad_v,
                              void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_rays(h, rays, "nemit_backward");
  if (rc) return rc;
  if (!h->grads_bound) return fail(NEMIT_E_UNBOUND, "nemit_backward: gradient buffers not bound (nemit_bind_grads)%s");
  if (rays->n == 0) return NEMIT_OK;
  if (!grad_rgb || !grad_o || !grad_v) return fail(NEMIT_E_INVALID, "nemit_backward: null gradient pointers%s");
  if (chunk_rays <= 0) chunk_rays = rays->n;
  WsLayout L = ws_layout(rays->n);
  if (!workspace || workspace_bytes < L.total)
    return fail(NEMIT_E_WORKSPACE, "nemit_backward: workspace too small%s (need %lld bytes)", "", (long long)L.total);
  DeviceGuard guard(h->device);
  char* ws = static_cast<char*>(workspace);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  const int64_t sub = bwd_sub_chunk(h, rays->n);
  ChunkEdges E0;
  int64_t edges_of = -1;
  for (int64_t off = 0; off < rays->n; off += sub) {
    int64_t n = rays->n - off < sub ? rays->n - off : sub;
    float* g_sigma = reinterpret_cast<float*>(ws + L.g_sigma);
    float* g_rgbs = reinterpret_cast<float*>(ws + L.g_rgbs);
    float* weights2 = reinterpret_cast<float*>(ws + L.weights2);
    float* ray_grad = reinterpret_cast<float*>(ws + L.ray_grad);
    const bool prop_pass = proposal_requires_grad && h->M.interlevel_mult != 0.f;
    // the sample edges are produced once per chunk of kBwdChunkRays rays, also when the mode walks it in smaller sub-chunks
    const int64_t outer = off / kBwdChunkRays * kBwdChunkRays;
    if (outer != edges_of) {
      const int64_t n_outer = rays->n - outer < kBwdChunkRays ? rays->n - outer : kBwdChunkRays;
      if ((rc = chunk_edges(h, h->M, rays, outer, n_outer, ws, L, prop_pass, st, &E0))) return rc;
      edges_of = outer;
    }
    ChunkEdges E = E0;
    E.t_edges2 += (off - outer) * (kS2 + 1);
    E.s_edges2 += (off - outer) * (kS2 + 1);
    if (E.s_edges1) E.s_edges1 += (off - outer) * (kS1 + 1);
    const float* t_edges = E.t_edges2;
    const float* s_edges = E.s_edges2;
    // pass 1: forward of the 48-sample stage + composite adjoint
    FieldArgs F;
    memset(&F, 0, sizeof(F));
    fill_ray_args(h, F, rays, off, n);
    fill_field_params(h->P, F.P);
    F.t_edges2 = t_edges; F.s_edges2 = s_edges; F.grad_rgb = grad_rgb;
    F.g_sigma = g_sigma; F.g_rgbs = g_rgbs; F.weights2 = weights2; F.reg_chunk = chunk_rays;
    float* g_absmax = reinterpret_cast<float*>(ws + L.g_absmax);
    if (tc_mode(h)) {
      if ((e = cudaMemsetAsync(g_absmax, 0, sizeof(float), st)) != cudaSuccess) return cuda_fail(e, "memset g_absmax");
      F.g_absmax = g_absmax;
      F.enc_tiles = reinterpret_cast<uint8_t*>(ws + L.enc_tiles);
      F.jac_tiles = reinterpret_cast<uint8_t*>(ws + L.jac_tiles);
      F.act_tiles = reinterpret_cast<uint8_t*>(ws + L.act_tiles);
      F.seeds4 = reinterpret_cast<float4*>(ws + L.seeds4);
      F.masks4 = reinterpret_cast<uint4*>(ws + L.masks4);
    }
    if (tc_mode(h))
      e = timed_launch(h, NEMIT_K_FIELD_SEED, st, n, [&] { return launch_field_adjoint_seed_tc(h->M, F, st, tc_pieces(h)); });
    else
      e = timed_launch(h, NEMIT_K_FIELD_SEED, st, n, [&] { return launch_field_adjoint_seed(h->M, F, st); });
    if (e != cudaSuccess) return cuda_fail(e, "k_field_forward<adjoint>");
    // pass 2: MLP / hash-grid backward
    if ((e = cudaMemsetAsync(ray_grad, 0, sizeof(float) * 6 * (size_t)n, st)) != cudaSuccess) return cuda_fail(e, "memset ray_grad");
    FieldBwdArgs B;
    memset(&B, 0, sizeof(B));
    fill_ray_args(h, B, rays, off, n);
    fill_field_params(h->P, B.P);
    B.G.main_table = reinterpret_cast<float2*>(h->G.main_table);
    B.G.base_w0 = h->G.base_w0; B.G.base_b0 = h->G.base_b0; B.G.base_w1 = h->G.base_w1; B.G.base_b1 = h->G.base_b1;
    B.G.head_w0 = h->G.head_w0; B.G.head_b0 = h->G.head_b0; B.G.head_w1 = h->G.head_w1; B.G.head_b1 = h->G.head_b1;
    B.G.head_w2 = h->G.head_w2; B.G.head_b2 = h->G.head_b2; B.G.appearance = h->G.appearance;
    B.t_edges2 = t_edges; B.g_sigma = g_sigma; B.g_rgbs = g_rgbs; B.ray_grad = ray_grad; B.loss_scale = loss_scale;
    if (tc_mode(h)) {
      B.g_absmax = g_absmax;
      B.enc_tiles = reinterpret_cast<const uint8_t*>(ws + L.enc_tiles);
      B.jac_tiles = reinterpret_cast<const uint8_t*>(ws + L.jac_tiles);
      B.act_tiles = reinterpret_cast<const uint8_t*>(ws + L.act_tiles);
      B.seeds4 = reinterpret_cast<const float4*>(ws + L.seeds4);
      B.masks4 = reinterpret_cast<const uint4*>(ws + L.masks4);
      B.red_ops = h->profile ? h->counters : nullptr;
      e = timed_launch(h, NEMIT_K_FIELD_BWD_TC, st, n, [&] { return launch_field_backward_tc(h->M, B, st, tc_pieces(h)); });
    } else {
      e = timed_launch(h, NEMIT_K_FIELD_BWD, st, n, [&] { return launch_field_backward(h->M, B, st); });
    }
    if (e != cudaSuccess) return cuda_fail(e, "k_field_backward");
    // pass 3: interlevel loss -> proposal networks (only when the sampler says they carry autograd)
    if (prop_pass) {
      PropBwdArgs Q;
      memset(&Q, 0, sizeof(Q));
      fill_sample_args(h, Q, rays, off, n);
      for (int i = 0; i < 2; ++i) {
        Q.g_table[i] = reinterpret_cast<float2*>(h->G.prop_table[i]);
        Q.g_w0[i] = h->G.prop_w0[i]; Q.g_b0[i] = h->G.prop_b0[i];
        Q.g_w1[i] = h->G.prop_w1[i]; Q.g_b1[i] = h->G.prop_b1[i];
      }
      Q.weights2 = weights2; Q.s_edges2_in = s_edges; Q.s_edges1_in = E.s_edges1; Q.reg_chunk = chunk_rays; Q.loss_scale = loss_scale;
      if ((e = timed_launch(h, NEMIT_K_PROP_BWD, st, n, [&] { return launch_prop_backward(h->M, Q, st); })) != cudaSuccess) return cuda_fail(e, "k_prop_backward");
    }
    FinalizeArgs Z;
    memset(&Z, 0, sizeof(Z));
    fill_ray_args(h, Z, rays, off, n);
    Z.ray_grad = ray_grad; Z.grad_o = grad_o; Z.grad_v = grad_v;
    if ((e = timed_launch(h, NEMIT_K_FINALIZE, st, n, [&] { return launch_finalize_ray_grads(h->M, Z, st); })) != cudaSuccess) return cuda_fail(e, "k_finalize_ray_grads");
  }
  return NEMIT_OK;
}

// stage of nemit_backward_host behind the ordinary workspace: o v (6) near (1) jitter (3) grad_rgb (3) rgb (3) grad_o (3)
// grad_v (3) = 22 floats per ray, then the primal's kept edges (49 + 49 + 97 floats per ray)
static size_t host_bwd_stage_bytes(int64_t n) {
  const size_t m = (size_t)(n < 1 ? 1 : n);
  return align_up(sizeof(float) * m * 22) + align_up(sizeof(float) * m * (kS2 + 1)) * 2 + align_up(sizeof(float) * m * (kS1 + 1));
}
extern "C" size_t nemit_backward_host_workspace_bytes(const NemitHandle* h, int64_t n) {
  (void)h;
  return ws_layout(n).total + host_bwd_stage_bytes(n);
}

extern "C" int nemit_backward_host(NemitHandle* h, const NemitRays* rh, const float* grad_rgb_host, int32_t primal_training,
                                   int64_t chunk_rays, float loss_scale, int32_t proposal_requires_grad, float* rgb_host,
                                   float* grad_o_host, float* grad_v_host, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  int rc = check_rays(h, rh, "nemit_backward_host");
  if (rc) return rc;
  if (!h->grads_bound) return fail(NEMIT_E_UNBOUND, "nemit_backward_host: gradient buffers not bound (nemit_bind_grads)%s");
  if (rh->n == 0) return NEMIT_OK;
  if (!grad_rgb_host || !rgb_host || !grad_o_host || !grad_v_host) return fail(NEMIT_E_INVALID, "nemit_backward_host: null host buffer%s");
  if (rh->kept_t_edges2) return fail(NEMIT_E_INVALID, "nemit_backward_host: kept edges are managed by the call itself%s");
  const int64_t n = rh->n;
  const size_t base = ws_layout(n).total;
  if (!workspace || workspace_bytes < base + host_bwd_stage_bytes(n))
    return fail(NEMIT_E_WORKSPACE, "nemit_backward_host: workspace too small%s (need %lld bytes)", "", (long long)(base + host_bwd_stage_bytes(n)));
  DeviceGuard guard(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  char* p = static_cast<char*>(workspace) + base;
  float* stage = reinterpret_cast<float*>(p);
  float *d_o = stage, *d_v = stage + 3 * n, *d_near = stage + 6 * n, *d_jit = stage + 7 * n, *d_g = stage + 10 * n,
        *d_rgb = stage + 13 * n, *d_go = stage + 16 * n, *d_gv = stage + 19 * n;
  p += align_up(sizeof(float) * (size_t)n * 22);
  float* k_t2 = reinterpret_cast<float*>(p); p += align_up(sizeof(float) * (size_t)n * (kS2 + 1));
  float* k_s2 = reinterpret_cast<float*>(p); p += align_up(sizeof(float) * (size_t)n * (kS2 + 1));
  float* k_s1 = reinterpret_cast<float*>(p);
  cudaError_t e;
  if ((e = cudaMemcpyAsync(d_o, rh->o, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D o");
  if ((e = cudaMemcpyAsync(d_v, rh->v, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D v");
  if ((e = cudaMemcpyAsync(d_g, grad_rgb_host, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D grad_rgb");
  NemitRays rd = *rh;
  rd.o = d_o; rd.v = d_v;
  if (rh->near_per_ray) {
    if ((e = cudaMemcpyAsync(d_near, rh->near_per_ray, sizeof(float) * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D near");
    rd.near_per_ray = d_near;
  }
  if (rh->jitter) {
    if ((e = cudaMemcpyAsync(d_jit, rh->jitter, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D jitter");
    rd.jitter = d_jit;
  }
  // primal (NeRFEmitterOp.eval): the same CustomOp instance's adjoint sees the same samples when the mode is the same
  const bool keep = (primal_training != 0) == (rh->training != 0);
  NemitRays rp = rd;
  rp.training = primal_training != 0;
  NemitLevelsOut LO;
  memset(&LO, 0, sizeof(LO));
  if (keep) { LO.t_edges2 = k_t2; LO.s_edges2 = k_s2; if (proposal_requires_grad) LO.s_edges1 = k_s1; }
  if ((rc = nemit_forward(h, &rp, d_rgb, nullptr, keep ? &LO : nullptr, workspace, base, stream))) return rc;
  if (keep) { rd.kept_t_edges2 = k_t2; rd.kept_s_edges2 = k_s2; rd.kept_s_edges1 = proposal_requires_grad ? k_s1 : nullptr; }
  if ((rc = nemit_backward(h, &rd, d_g, chunk_rays, loss_scale, proposal_requires_grad, d_go, d_gv, workspace, base, stream))) return rc;
  if ((e = cudaMemcpyAsync(rgb_host, d_rgb, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return cuda_fail(e, "D2H rgb");
  if ((e = cudaMemcpyAsync(grad_o_host, d_go, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return cuda_fail(e, "D2H grad_o");
  if ((e = cudaMemcpyAsync(grad_v_host, d_gv, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return cuda_fail(e, "D2H grad_v");
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// light-point-cloud pass: SdfNerfactoModel.get_point_lights (NS/models/sdf_nerfacto.py:189-212)
// ---------------------------------------------------------------------------------------------
extern "C" int nemit_point_lights(NemitHandle* h, const NemitRays* rays, const float* rgb_weights, float* rgb,
                                  float* depth, float* brightness_grad, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  int rc = check_rays(h, rays, "nemit_point_lights");
  if (rc) return rc;
  if (rays->n == 0) return NEMIT_OK;
  if (!rgb_weights || !rgb || !depth || !brightness_grad) return fail(NEMIT_E_INVALID, "nemit_point_lights: null argument%s");
  WsLayout L = ws_layout(rays->n);
  if (!workspace || workspace_bytes < L.total)
    return fail(NEMIT_E_WORKSPACE, "nemit_point_lights: workspace too small%s (need %lld bytes)", "", (long long)L.total);
  DeviceGuard guard(h->device);
  char* ws = static_cast<char*>(workspace);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  ModelDev M = h->M;
  M.distortion_mult = 0.f;           // the objective is the brightness alone (nerfacto.py:562-568): no regularisers
  const bool tc = tc_mode(h);
  const int64_t sub = bwd_sub_chunk(h, rays->n);
  for (int64_t off = 0; off < rays->n; off += sub) {
    int64_t n = rays->n - off < sub ? rays->n - off : sub;
    float* g_sigma = reinterpret_cast<float*>(ws + L.g_sigma);
    float* g_rgbs = reinterpret_cast<float*>(ws + L.g_rgbs);
    float* ray_grad = reinterpret_cast<float*>(ws + L.ray_grad);
    float* g_absmax = reinterpret_cast<float*>(ws + L.g_absmax);
    ChunkEdges E;
    if ((rc = chunk_edges(h, M, rays, off, n, ws, L, false, st, &E))) return rc;
    const float* t_edges = E.t_edges2;
    const float* s_edges = E.s_edges2;
    FieldArgs F;
    memset(&F, 0, sizeof(F));
    fill_ray_args(h, F, rays, off, n, true);
    fill_field_params(h->P, F.P);
    F.t_edges2 = t_edges; F.s_edges2 = s_edges; F.g_sigma = g_sigma; F.g_rgbs = g_rgbs; F.reg_chunk = rays->n;
    F.point_lights = 1;
    for (int k = 0; k < 3; ++k) F.lum_w[k] = rgb_weights[k];
    F.rgb = rgb + off * 3; F.depth = depth + off;
    if (tc) {
      if ((e = cudaMemsetAsync(g_absmax, 0, sizeof(float), st)) != cudaSuccess) return cuda_fail(e, "memset g_absmax");
      F.g_absmax = g_absmax;
      F.enc_tiles = reinterpret_cast<uint8_t*>(ws + L.enc_tiles);
      F.jac_tiles = reinterpret_cast<uint8_t*>(ws + L.jac_tiles);
      F.act_tiles = reinterpret_cast<uint8_t*>(ws + L.act_tiles);
      F.seeds4 = reinterpret_cast<float4*>(ws + L.seeds4);
      F.masks4 = reinterpret_cast<uint4*>(ws + L.masks4);
      e = timed_launch(h, NEMIT_K_FIELD_SEED, st, n, [&] { return launch_field_adjoint_seed_tc(M, F, st, tc_pieces(h)); });
    } else {
      e = timed_launch(h, NEMIT_K_FIELD_SEED, st, n, [&] { return launch_field_adjoint_seed(M, F, st); });
    }
    if (e != cudaSuccess) return cuda_fail(e, "k_field_forward<adjoint>");
    if ((e = cudaMemsetAsync(ray_grad, 0, sizeof(float) * 6 * (size_t)n, st)) != cudaSuccess) return cuda_fail(e, "memset ray_grad");
    FieldBwdArgs B;
    memset(&B, 0, sizeof(B));       // B.G stays null: ray gradients only
    fill_ray_args(h, B, rays, off, n, true);
    fill_field_params(h->P, B.P);
    B.t_edges2 = t_edges; B.g_sigma = g_sigma; B.g_rgbs = g_rgbs; B.ray_grad = ray_grad; B.loss_scale = 1.f;
    if (tc) {
      B.g_absmax = g_absmax;
      B.enc_tiles = reinterpret_cast<const uint8_t*>(ws + L.enc_tiles);
      B.jac_tiles = reinterpret_cast<const uint8_t*>(ws + L.jac_tiles);
      B.act_tiles = reinterpret_cast<const uint8_t*>(ws + L.act_tiles);
      B.seeds4 = reinterpret_cast<const float4*>(ws + L.seeds4);
      B.masks4 = reinterpret_cast<const uint4*>(ws + L.masks4);
      e = timed_launch(h, NEMIT_K_FIELD_BWD_TC, st, n, [&] { return launch_field_backward_tc(M, B, st, tc_pieces(h)); });
    } else {
      e = timed_launch(h, NEMIT_K_FIELD_BWD, st, n, [&] { return launch_field_backward(M, B, st); });
    }
    if (e != cudaSuccess) return cuda_fail(e, "k_field_backward");
    FinalizeArgs Z;
    memset(&Z, 0, sizeof(Z));
    fill_ray_args(h, Z, rays, off, n);
    Z.ray_grad = ray_grad; Z.grad_o = brightness_grad; Z.grad_v = nullptr;
    // brightness_grad = d / d ray_bundle.origins (nerfacto.py:565-568): the op's near change of variables
    // (o.grad -= near * v.grad, nerf_emitter_op.py:112) does not belong to this pass
    Z.near_per_ray = nullptr; Z.near_scalar = 0.f;
    if ((e = timed_launch(h, NEMIT_K_FINALIZE, st, n, [&] { return launch_finalize_ray_grads(M, Z, st); })) != cudaSuccess) return cuda_fail(e, "k_finalize_ray_grads");
  }
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// optimizer step (optim.cu)
// ---------------------------------------------------------------------------------------------
extern "C" int nemit_adam_step(NemitHandle* h, const NemitParams* exp_avg, const NemitParams* exp_avg_sq,
                               const NemitAdamGroup* fields, const NemitAdamGroup* proposal_networks, int64_t* step_state,
                               float grad_scale, int32_t zero_grads, float* stats_or_null, void* stream) {
  if (!h || !exp_avg || !exp_avg_sq || !fields || !proposal_networks || !step_state) return fail(NEMIT_E_INVALID, "nemit_adam_step: null argument%s");
  if (!h->params_bound || !h->grads_bound) return fail(NEMIT_E_UNBOUND, "nemit_adam_step: parameters and gradients must be bound%s");
  if (!params_complete(exp_avg) || !params_complete(exp_avg_sq)) return fail(NEMIT_E_INVALID, "nemit_adam_step: an optimizer-state pointer is null%s");
  if ((uintptr_t)step_state & 7) return fail(NEMIT_E_INVALID, "nemit_adam_step: step_state must be an 8-byte aligned device int64[2]%s");
  if (!(grad_scale > 0.f)) return fail(NEMIT_E_INVALID, "nemit_adam_step: grad_scale must be > 0%s");
  DeviceGuard guard(h->device);
  const NemitConfig& c = h->cfg;
  const int64_t main_n = ((int64_t)c.main_grid.num_levels << c.main_grid.log2_table_size) * 2;
  const int64_t enc = c.main_grid.num_levels * 2, hd = c.hidden_dim, head_in = 16 + c.geo_feat_dim + c.appearance_dim;
  const int64_t numel_main[12] = {main_n, hd * enc, hd, (1 + c.geo_feat_dim) * hd, 1 + c.geo_feat_dim, hd * head_in, hd,
                                  hd * hd, hd, 3 * hd, 3, (int64_t)c.num_images * c.appearance_dim};
  OptimArgs A;
  memset(&A, 0, sizeof(A));
  float* const* P = reinterpret_cast<float* const*>(&h->P);
  float* const* G = reinterpret_cast<float* const*>(&h->G);
  float* const* M1 = reinterpret_cast<float* const*>(exp_avg);
  float* const* M2 = reinterpret_cast<float* const*>(exp_avg_sq);
  int t = 0;
  for (int i = 0; i < 12; ++i, ++t) {
    A.param[t] = P[i]; A.grad[t] = G[i]; A.exp_avg[t] = M1[i]; A.exp_avg_sq[t] = M2[i]; A.numel[t] = numel_main[i]; A.group[t] = 0;
  }
  // NemitParams tail: prop_table[2], prop_w0[2], prop_b0[2], prop_w1[2], prop_b1[2]
  for (int f = 0; f < 5; ++f)
    for (int i = 0; i < 2; ++i, ++t) {
      const NemitGrid& g = c.prop_grid[i];
      const int64_t pe = g.num_levels * 2, ph = c.prop_hidden_dim;
      const int64_t nn[5] = {((int64_t)g.num_levels << g.log2_table_size) * 2, ph * pe, ph, ph, 1};
      const int k = 12 + f * 2 + i;
      A.param[t] = P[k]; A.grad[t] = G[k]; A.exp_avg[t] = M1[k]; A.exp_avg_sq[t] = M2[k]; A.numel[t] = nn[f]; A.group[t] = 1;
    }
  A.num_tensors = t;
  for (int i = 0; i < t; ++i)
    if ((A.numel[i] & 3) == 0 &&
        (((uintptr_t)A.param[i] | (uintptr_t)A.grad[i] | (uintptr_t)A.exp_avg[i] | (uintptr_t)A.exp_avg_sq[i]) & 15))
      return fail(NEMIT_E_INVALID, "nemit_adam_step: tensors are streamed as float4 and must be 16-byte aligned%s");
  const NemitAdamGroup* src[2] = {fields, proposal_networks};
  for (int g = 0; g < 2; ++g) {
    const NemitAdamGroup& s = *src[g];
    if (!(s.beta1 >= 0.0 && s.beta1 < 1.0 && s.beta2 >= 0.0 && s.beta2 < 1.0 && s.eps >= 0.0))
      return fail(NEMIT_E_INVALID, "nemit_adam_step: betas must be in [0,1) and eps >= 0%s");
    OptimGroup& d = A.grp[g];
    d.beta1 = (float)s.beta1; d.beta2 = (float)s.beta2; d.eps = (float)s.eps; d.max_norm = s.max_norm; d.max_value = s.max_value;
    d.omb1 = (float)(1.0 - s.beta1); d.omb2 = (float)(1.0 - s.beta2);
    d.lr = s.lr; d.beta1_d = s.beta1; d.beta2_d = s.beta2;
    d.enabled = s.enabled != 0;
  }
  A.step_state = reinterpret_cast<long long*>(step_state);
  A.inv_scale = 1.f / grad_scale;
  A.zero_grads = zero_grads;
  A.stats = h->opt_stats;
  A.stats_out = stats_or_null;
  cudaError_t e = launch_adam_step(A, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "k_adam_step");
  h->launches += 3;
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// spherical-sensor bake (bake.cu)
// ---------------------------------------------------------------------------------------------
static int check_sensor(const NemitSphericalSensor* s, const char* who) {
  if (!s) return fail(NEMIT_E_INVALID, "%s: null sensor", who);
  if (s->width < 1 || s->height < 1 || s->spp < 1) return fail(NEMIT_E_INVALID, "%s: film size and spp must be >= 1", who);
  if ((int64_t)s->width * s->height * s->spp >= ((int64_t)1 << 31)) return fail(NEMIT_E_INVALID, "%s: more than 2^31 film samples", who);
  return NEMIT_OK;
}
static void fill_spherical(SphericalArgs& A, const NemitSphericalSensor* s, int64_t first, int64_t n, const float* jitter,
                           float* o, float* v, float* pos) {
  memset(&A, 0, sizeof(A));
  for (int k = 0; k < 3; ++k) A.origin[k] = s->origin[k];
  A.near_clip = s->near_clip; A.width = s->width; A.height = s->height; A.spp = s->spp; A.seed = s->seed;
  A.first = first; A.n = n; A.jitter = jitter; A.o = o; A.v = v; A.pos = pos;
}
extern "C" int nemit_spherical_rays(const NemitSphericalSensor* s, int64_t first, int64_t n, const float* jitter_or_null,
                                    float* o, float* v, float* pos, void* stream) {
  int rc = check_sensor(s, "nemit_spherical_rays");
  if (rc) return rc;
  if (first < 0 || n < 0 || first + n > (int64_t)s->width * s->height * s->spp) return fail(NEMIT_E_INVALID, "nemit_spherical_rays: sample range outside the film%s");
  if (n > 0 && (!o || !v || !pos)) return fail(NEMIT_E_INVALID, "nemit_spherical_rays: null output%s");
  SphericalArgs A;
  fill_spherical(A, s, first, n, jitter_or_null, o, v, pos);
  cudaError_t e = launch_spherical_rays(A, (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_spherical_rays");
}
extern "C" int nemit_film_accumulate(int32_t width, int32_t height, int32_t filter, float stddev, const float* pos,
                                     const float* rgb, int64_t n, float* accum, void* stream) {
  if (width < 1 || height < 1 || n < 0 || (n > 0 && (!pos || !rgb || !accum))) return fail(NEMIT_E_INVALID, "nemit_film_accumulate: bad argument%s");
  if (filter != NEMIT_FILTER_BOX && filter != NEMIT_FILTER_GAUSSIAN) return fail(NEMIT_E_INVALID, "nemit_film_accumulate: unknown filter%s");
  if (filter == NEMIT_FILTER_GAUSSIAN && !(stddev > 0.f)) return fail(NEMIT_E_INVALID, "nemit_film_accumulate: stddev must be > 0%s");
  FilmArgs A;
  memset(&A, 0, sizeof(A));
  A.width = width; A.height = height; A.filter = filter; A.stddev = stddev; A.pos = pos; A.rgb = rgb; A.n = n;
  A.accum = reinterpret_cast<float4*>(accum);
  cudaError_t e = launch_film_accumulate(A, (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_film_accumulate");
}
extern "C" int nemit_film_develop(int32_t width, int32_t height, const float* accum, float* rgba, void* stream) {
  if (width < 1 || height < 1 || !accum || !rgba) return fail(NEMIT_E_INVALID, "nemit_film_develop: bad argument%s");
  cudaError_t e = launch_film_develop((int64_t)width * height, accum, rgba, (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_film_develop");
}
extern "C" size_t nemit_bake_workspace_bytes(const NemitHandle* h) {
  (void)h;
  return ws_layout(kChunkRays).total + align_up(sizeof(float) * 11 * (size_t)kChunkRays);
}
extern "C" int nemit_bake_envmap(NemitHandle* h, const NemitSphericalSensor* s, int32_t filter, float stddev,
                                 int32_t row_begin, int32_t row_end, int32_t camera_idx, float* accum, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  int rc = check_sensor(s, "nemit_bake_envmap");
  if (rc) return rc;
  if (!h || !accum) return fail(NEMIT_E_INVALID, "nemit_bake_envmap: null argument%s");
  if (row_begin < 0 || row_end > s->height || row_begin > row_end) return fail(NEMIT_E_INVALID, "nemit_bake_envmap: bad row range%s");
  if (!workspace || workspace_bytes < nemit_bake_workspace_bytes(h)) return fail(NEMIT_E_WORKSPACE, "nemit_bake_envmap: workspace too small%s");
  DeviceGuard guard(h->device);      // the ray-generation and film kernels below carry no handle: run them on its device
  char* ws = static_cast<char*>(workspace);
  const size_t fwd_bytes = ws_layout(kChunkRays).total;
  float* stage = reinterpret_cast<float*>(ws + fwd_bytes);
  float *o = stage, *v = stage + 3 * (size_t)kChunkRays, *pos = stage + 6 * (size_t)kChunkRays, *rgb = stage + 8 * (size_t)kChunkRays;
  const int64_t first = (int64_t)row_begin * s->width * s->spp, last = (int64_t)row_end * s->width * s->spp;
  for (int64_t off = first; off < last; off += kChunkRays) {
    const int64_t n = last - off < kChunkRays ? last - off : kChunkRays;
    if ((rc = nemit_spherical_rays(s, off, n, nullptr, o, v, pos, stream))) return rc;
    NemitRays R;
    memset(&R, 0, sizeof(R));
    R.o = o; R.v = v; R.n = n; R.near_scalar = 0.f; R.camera_idx = camera_idx; R.training = 0; R.anneal = 1.f;
    if ((rc = nemit_forward(h, &R, rgb, nullptr, nullptr, workspace, fwd_bytes, stream))) return rc;
    if ((rc = nemit_film_accumulate(s->width, s->height, filter, stddev, pos, rgb, n, accum, stream))) return rc;
    h->launches += 2;
  }
  return NEMIT_OK;
}

// ---------------------------------------------------------------------------------------------
// light point cloud -> spherical-Gaussian mixture (gmm.cu)
// ---------------------------------------------------------------------------------------------
extern "C" int nemit_compensate_weights(const float* lum, int64_t n, float threshold, int32_t use_threshold, float* w,
                                        void* scratch, void* stream) {
  if (n < 0 || (n > 0 && (!lum || !w)) || !scratch) return fail(NEMIT_E_INVALID, "nemit_compensate_weights: bad argument%s");
  cudaError_t e = launch_compensate(lum, n, threshold, use_threshold, static_cast<double*>(scratch), w, (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_compensate");
}
extern "C" int nemit_gmm_em_step(const float* x, const float* w, int64_t n, int32_t k, const float* mean, const float* var,
                                 const float* prior, float min_var, float* mean_out, float* var_out, float* prior_out,
                                 float* logp_out, void* scratch, void* stream) {
  if (n < 0 || (n > 0 && (!x || !w)) || !mean || !var || !prior || !mean_out || !var_out || !prior_out || !logp_out || !scratch)
    return fail(NEMIT_E_INVALID, "nemit_gmm_em_step: null argument%s");
  if (k < 1 || k > 128) return fail(NEMIT_E_INVALID, "nemit_gmm_em_step: 1..128 components%s");
  if (mean == mean_out || var == var_out || prior == prior_out) return fail(NEMIT_E_INVALID, "nemit_gmm_em_step: outputs may not alias inputs%s");
  GmmArgs A;
  memset(&A, 0, sizeof(A));
  A.x = x; A.w = w; A.n = n; A.k = k; A.mean = mean; A.var = var; A.prior = prior; A.min_var = min_var;
  A.stats = static_cast<double*>(scratch);
  A.mean_out = mean_out; A.var_out = var_out; A.prior_out = prior_out; A.logp_out = logp_out;
  cudaError_t e = launch_gmm_em(A, (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_gmm_em");
}

// ---------------------------------------------------------------------------------------------
// vMF mixture (DS/emitters/vMF.py)
// ---------------------------------------------------------------------------------------------
static int check_lobes(const NemitVmfLobes* l, const char* who) {
  if (!l || !l->position || !l->weight || !l->std) return fail(NEMIT_E_INVALID, "%s: null lobe table", who);
  if (l->num_lobes < 1 || l->num_lobes > 1024) return fail(NEMIT_E_INVALID, "%s: num_lobes must be in [1,1024]", who);
  return NEMIT_OK;
}
extern "C" int nemit_vmf_sample(const NemitVmfLobes* lobes, const float* p, const float* u, int64_t n, float* d,
                                float* pdf, void* stream) {
  int rc = check_lobes(lobes, "nemit_vmf_sample");
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!p || !u || !d || !pdf))) return fail(NEMIT_E_INVALID, "nemit_vmf_sample: bad sample arguments%s");
  cudaError_t e = launch_vmf_sample(lobes->position, lobes->weight, lobes->std, lobes->num_lobes, p, u, n, d, pdf,
                                    (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_vmf_sample");
}
extern "C" int nemit_vmf_pdf(const NemitVmfLobes* lobes, const float* p, const float* d, int64_t n, float* pdf,
                             void* stream) {
  int rc = check_lobes(lobes, "nemit_vmf_pdf");
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!p || !d || !pdf))) return fail(NEMIT_E_INVALID, "nemit_vmf_pdf: bad sample arguments%s");
  cudaError_t e = launch_vmf_pdf(lobes->position, lobes->weight, lobes->std, lobes->num_lobes, p, d, n, pdf,
                                 (cudaStream_t)stream);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_vmf_pdf");
}

static void frame_to_dev(const NemitVmfFrame* f, VmfFrame& F) {
  memset(&F, 0, sizeof(F));
  F.on = 1;
  memcpy(F.w2l, f->world_to_local, sizeof(F.w2l));
  memcpy(F.l2w, f->local_to_world, sizeof(F.l2w));
  memcpy(F.centre, f->bsphere_center, sizeof(F.centre));
  F.radius = f->bsphere_radius;
}
extern "C" int nemit_vmf_sample_direction(const NemitVmfLobes* lobes, const NemitVmfFrame* frame, const float* it_p,
                                          const float* u, int64_t n, float* d, float* pdf, float* dist_or_null,
                                          float* ds_p_or_null, void* stream) {
  int rc = check_lobes(lobes, "nemit_vmf_sample_direction");
  if (rc) return rc;
  if (!frame) return fail(NEMIT_E_INVALID, "nemit_vmf_sample_direction: null frame%s");
  if (n < 0 || (n > 0 && (!it_p || !u || !d || !pdf))) return fail(NEMIT_E_INVALID, "nemit_vmf_sample_direction: bad sample arguments%s");
  VmfFrame F;
  frame_to_dev(frame, F);
  cudaError_t e = launch_vmf_sample(lobes->position, lobes->weight, lobes->std, lobes->num_lobes, it_p, u, n, d, pdf,
                                    (cudaStream_t)stream, &F, dist_or_null, ds_p_or_null);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_vmf_sample");
}
extern "C" int nemit_vmf_pdf_direction(const NemitVmfLobes* lobes, const NemitVmfFrame* frame, const float* it_p,
                                       const float* ds_d, int64_t n, float* pdf, void* stream) {
  int rc = check_lobes(lobes, "nemit_vmf_pdf_direction");
  if (rc) return rc;
  if (!frame) return fail(NEMIT_E_INVALID, "nemit_vmf_pdf_direction: null frame%s");
  if (n < 0 || (n > 0 && (!it_p || !ds_d || !pdf))) return fail(NEMIT_E_INVALID, "nemit_vmf_pdf_direction: bad sample arguments%s");
  VmfFrame F;
  frame_to_dev(frame, F);
  cudaError_t e = launch_vmf_pdf(lobes->position, lobes->weight, lobes->std, lobes->num_lobes, it_p, ds_d, n, pdf,
                                 (cudaStream_t)stream, &F);
  return e == cudaSuccess ? NEMIT_OK : cuda_fail(e, "k_vmf_pdf");
}

// ---------------------------------------------------------------------------------------------
// allocation-free forward (no counterpart in the reference)
// ---------------------------------------------------------------------------------------------
// nemit_forward without a caller-provided workspace, for callers that cannot allocate device memory (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  The one piece of scratch the forward needs - the final sample
// edges t_edges2 that k_sample_levels writes and k_field_forward(_tc) reads, (n, 49) floats - lives in a file-scope device
// array of NEMIT_NOALLOC_MAX_RAYS x 49 floats = 3.06 MiB, held by every GPU that loads the library until exit.
// In host code `g_noalloc_edges` is only a stand-in for the symbol: its device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and it is that pointer which goes to the kernels.
// Nothing in the array has to start at zero (the sampler writes rows [0, n) before the field kernel reads exactly those
// rows), so no clearing pass; every call shares the array and calls that overlap on one GPU (two streams) race on it.
static __device__ __align__(256) float g_noalloc_edges[(size_t)NEMIT_NOALLOC_MAX_RAYS * (kS2 + 1)];

static cudaError_t noalloc_edges_ptr(int device, float** out) {
  static float* ptr[64] = {};
  static std::mutex mu;
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_edges);      // the calling thread's current device = `device`
    if (e != cudaSuccess) return e;
    ptr[device] = static_cast<float*>(p);
  }
  *out = ptr[device];
  return cudaSuccess;
}

extern "C" int nemit_forward_noalloc(NemitHandle* h, const NemitRays* rays, float* rgb, void* stream) {
  int rc = check_rays(h, rays, "nemit_forward_noalloc");
  if (rc) return rc;
  if (rays->n > NEMIT_NOALLOC_MAX_RAYS)
    return fail(NEMIT_E_INVALID, "nemit_forward_noalloc: %s", "more rays than NEMIT_NOALLOC_MAX_RAYS (16384): split the batch or use nemit_forward");
  if (rays->n == 0) return NEMIT_OK;
  if (!rgb) return fail(NEMIT_E_INVALID, "nemit_forward_noalloc: null rgb output%s");
  DeviceGuard guard(h->device);
  float* t_edges = nullptr;
  cudaError_t e = noalloc_edges_ptr(h->device, &t_edges);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_edges)");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = rays->n;
  SampleArgs S;
  fill_sample_args(h, S, rays, 0, n);
  S.t_edges2 = t_edges;
  const bool fast = h->sampler == NEMIT_SAMPLER_FAST;
  if ((e = launch_sample_levels(h->M, S, st, fast)) != cudaSuccess) return cuda_fail(e, "k_sample_levels");
  h->launches += fast ? 2 : 1;       // (no event bracketing here: the profile ring creates its events on first use)
  FieldArgs F;
  memset(&F, 0, sizeof(F));
  fill_ray_args(h, F, rays, 0, n, true);
  fill_field_params(h->P, F.P);
  F.t_edges2 = t_edges;
  F.rgb = rgb;
  e = tc_mode(h) ? launch_field_forward_tc(h->M, F, st, tc_pieces(h)) : launch_field_forward(h->M, F, st);
  if (e != cudaSuccess) return cuda_fail(e, "k_field_forward");
  h->launches++;
  return NEMIT_OK;
}
