// comm.cu - the one collective of the emitter path behind the C ABI: sum all-reduce of the flat parameter-gradient blob
// over the ray-sharded ranks, with the GradScaler unscale (or a 1/world average) folded into the collective itself
// through ncclRedOpCreatePreMulSum - one pass over the 77.6 MB instead of all-reduce + a separate multiply.
//
// Reference: DDP's gradient all-reduce (NS/pipelines/base_pipeline.py:355-357) after the no_sync accumulation of the
// backward chunks (NS/pipelines/mitsuba_sdf.py:1179-1190), then GradScaler.unscale_ (:1193).
//
// NCCL is bound at run time with dlopen: the process that hosts the library (torch) has already mapped its own
// libnccl.so.2, which is then shared; a bare C host gets the system one.  No link-time dependency.
#include <dlfcn.h>

#include <cstring>

#include "handle.h"

namespace {

struct NcclId { char internal[NEMIT_UNIQUE_ID_BYTES]; };
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(void**, int, NcclId, int);
typedef int (*fn_comm_destroy)(void*);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_redop_premul)(int*, void*, int, int, void*);
typedef int (*fn_redop_destroy)(int, void*);
typedef const char* (*fn_error_string)(int);

constexpr int kNcclFloat = 7;              // ncclFloat32 (nccl.h)
constexpr int kNcclSum = 0;                // ncclSum
constexpr int kNcclScalarHostImmediate = 1;

struct NcclApi {
  void* lib = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_redop_premul redop_premul = nullptr;
  fn_redop_destroy redop_destroy = nullptr;
  fn_error_string error_string = nullptr;
  bool ok = false;
};

NcclApi& nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // the copy already in the process (torch's)
  if (!api.lib) api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!api.lib) api.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!api.lib) return api;
  api.get_unique_id = (fn_get_unique_id)dlsym(api.lib, "ncclGetUniqueId");
  api.comm_init_rank = (fn_comm_init_rank)dlsym(api.lib, "ncclCommInitRank");
  api.comm_destroy = (fn_comm_destroy)dlsym(api.lib, "ncclCommDestroy");
  api.all_reduce = (fn_all_reduce)dlsym(api.lib, "ncclAllReduce");
  api.redop_premul = (fn_redop_premul)dlsym(api.lib, "ncclRedOpCreatePreMulSum");
  api.redop_destroy = (fn_redop_destroy)dlsym(api.lib, "ncclRedOpDestroy");
  api.error_string = (fn_error_string)dlsym(api.lib, "ncclGetErrorString");
  api.ok = api.get_unique_id && api.comm_init_rank && api.comm_destroy && api.all_reduce && api.redop_premul &&
           api.redop_destroy && api.error_string;
  return api;
}

int nccl_fail(int rc, const char* where) {
  nemit_fail(NEMIT_E_CUDA, "NCCL error in %s: ", where);
  nemit_append_error(nccl().error_string ? nccl().error_string(rc) : "?");
  return NEMIT_E_CUDA;
}

__global__ void k_scale_inplace(float* __restrict__ x, int64_t n, float s) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n4 = n >> 2;
  float4* x4 = reinterpret_cast<float4*>(x);
  for (int64_t j = i; j < n4; j += stride) {
    float4 v = x4[j];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    x4[j] = v;
  }
  for (int64_t j = (n4 << 2) + i; j < n; j += stride) x[j] *= s;
}

}  // namespace

extern "C" int nemit_comm_unique_id(void* id_out_128) {
  if (!id_out_128) return nemit_fail(NEMIT_E_INVALID, "nemit_comm_unique_id: null argument%s");
  NcclApi& N = nccl();
  if (!N.ok) return nemit_fail(NEMIT_E_INVALID, "nemit_comm_unique_id: libnccl.so.2 could not be loaded%s");
  NcclId id;
  int rc = N.get_unique_id(&id);
  if (rc) return nccl_fail(rc, "ncclGetUniqueId");
  memcpy(id_out_128, &id, sizeof(id));
  return NEMIT_OK;
}

extern "C" int nemit_comm_init(NemitHandle* h, const void* id_128, int32_t rank, int32_t world_size) {
  if (!h || !id_128) return nemit_fail(NEMIT_E_INVALID, "nemit_comm_init: null argument%s");
  if (world_size < 1 || rank < 0 || rank >= world_size) return nemit_fail(NEMIT_E_INVALID, "nemit_comm_init: bad rank / world size%s");
  NcclApi& N = nccl();
  if (!N.ok) return nemit_fail(NEMIT_E_INVALID, "nemit_comm_init: libnccl.so.2 could not be loaded%s");
  DeviceGuard guard(h->device);
  nemit_comm_destroy(h);
  NcclId id;
  memcpy(&id, id_128, sizeof(id));
  void* comm = nullptr;
  int rc = N.comm_init_rank(&comm, world_size, id, rank);
  if (rc) return nccl_fail(rc, "ncclCommInitRank");
  h->nccl_comm = comm;
  h->comm_world = world_size;
  return NEMIT_OK;
}

extern "C" int nemit_comm_destroy(NemitHandle* h) {
  if (!h) return NEMIT_OK;
  if (h->nccl_comm && nccl().ok) {
    DeviceGuard guard(h->device);
    nccl().comm_destroy(h->nccl_comm);
  }
  h->nccl_comm = nullptr;
  h->comm_world = 0;
  return NEMIT_OK;
}

extern "C" int nemit_allreduce_grads(NemitHandle* h, float* flat_grads, int64_t count, float scale, void* stream) {
  if (!h || count < 0 || (count > 0 && !flat_grads)) return nemit_fail(NEMIT_E_INVALID, "nemit_allreduce_grads: bad argument%s");
  if (count == 0) return NEMIT_OK;
  DeviceGuard guard(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (!h->nccl_comm || h->comm_world == 1) {
    // single rank: the sum is the identity, only the scale remains.  (A one-rank NCCL communicator is not used for it:
    // NCCL 2.28.9's in-place PreMulSum all-reduce on ONE rank leaves a tail of count % 4 elements unscaled - seen on
    // B200, tools/diag_round2.py; with two or more ranks the collective kernels run and the tail is handled.)
    if (scale != 1.f) {
      if ((uintptr_t)flat_grads & 15) return nemit_fail(NEMIT_E_INVALID, "nemit_allreduce_grads: the gradient blob must be 16-byte aligned%s");
      k_scale_inplace<<<148 * 8, 256, 0, st>>>(flat_grads, count, scale);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return nemit_cuda_fail(e, "k_scale_inplace");
      h->launches++;
    }
    return NEMIT_OK;
  }
  NcclApi& N = nccl();
  int op = kNcclSum;
  const bool premul = scale != 1.f;
  if (premul) {
    int rc = N.redop_premul(&op, &scale, kNcclFloat, kNcclScalarHostImmediate, h->nccl_comm);
    if (rc) return nccl_fail(rc, "ncclRedOpCreatePreMulSum");
  }
  int rc = N.all_reduce(flat_grads, flat_grads, (size_t)count, kNcclFloat, op, h->nccl_comm, st);
  if (premul) N.redop_destroy(op, h->nccl_comm);      // NCCL keeps the op alive until the enqueued collective has used it
  if (rc) return nccl_fail(rc, "ncclAllReduce");
  h->launches++;
  return NEMIT_OK;
}
