// handle.h - the library handle and the helpers every C-ABI translation unit shares (internal)
#pragma once
#include <cstdint>
#include <string>

#include "kernels.h"

constexpr int kEvRing = 4096;      // pending launch records of the optional per-kernel timing

struct NemitHandle {
  int device;
  nemit::ModelDev M;
  NemitConfig cfg;
  float* tables_dev;        // lin0 | u1 | u2 copies of the host sampler tables
  const float* lin0;
  const float* u1;
  const float* u2;
  double* opt_stats;        // optimizer-step scratch (tail of tables_dev)
  uint32_t* prop_half[2];   // packed-half copies of the proposal tables for the fast sampler (scratch, rewritten per launch)
  unsigned long long* counters;   // device counters: [0] red lane-ops of the tensor-core backward (profiling only)
  bool params_bound, grads_bound;
  // nemit_set_appearance_mode
  int always_use_camera_indices, use_average_embedding;
  int32_t* rotation_ids;    // host copy of Rotater.rotation_ids, or null
  int32_t num_rotation_ids;
  // nemit_set_rotater: Rotater.apply_frustums (bound + one 3x4 transform per rotation id), host copies
  int rot_mode;             // 0 = none, NEMIT_ROT_SPHERE, NEMIT_ROT_BOX
  int num_rotations;
  float* rot_transforms;    // num_rotations x 12
  float rot_bound[6];
  NemitParams P, G;
  int64_t launches;
  int precision;
  int sampler;
  // optional per-kernel timing (nemit_profile_*): CUDA events around every launch on the launching stream, kept in a
  // ring; when it fills the oldest record is folded into the sums (one event synchronise), so no launch is dropped
  bool profile;
  int ev_head, ev_count;
  cudaEvent_t* ev;          // 2 * kEvRing, created lazily
  int* ev_kind;
  double kind_ms[NEMIT_PROFILE_KINDS];
  int64_t kind_launches[NEMIT_PROFILE_KINDS];
  int64_t kind_rays[NEMIT_PROFILE_KINDS];
  // nemit_comm_*: NCCL communicator (comm.cu), null = single rank
  void* nccl_comm;
  int comm_world;
};

// error reporting shared by the translation units (api.cu owns the thread-local message)
int nemit_fail(int code, const char* fmt, const char* a = "", long long b = 0);
int nemit_cuda_fail(cudaError_t e, const char* where);
void nemit_append_error(const char* more);

// makes the handle's device current for the duration of an entry point and restores the caller's device afterwards
struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) changed = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};
