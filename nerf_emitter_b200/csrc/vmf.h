// vmf.h - launchers of the vMF-mixture kernels (vmf.cu) and backward/jvp entry points (internal)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace nemit {
size_t backward_chunk_bytes(int64_t chunk_rays);
// world <-> emitter-local frame of a VonMisesFisherEmitter and its bounding sphere (vMF.py:25,37-42,65-67); on == 0: the
// points and directions are emitter-local already
struct VmfFrame {
  int on;
  float w2l[12];      // world_transform().inverse(), 3 x 4 row-major
  float l2w[12];      // world_transform()
  float centre[3];    // m_bsphere
  float radius;
};
cudaError_t launch_vmf_sample(const float* pos, const float* weight, const float* std, int K, const float* p,
                              const float* u, int64_t n, float* d, float* pdf, cudaStream_t stream,
                              const VmfFrame* frame = nullptr, float* dist = nullptr, float* ds_p = nullptr);
cudaError_t launch_vmf_pdf(const float* pos, const float* weight, const float* std, int K, const float* p,
                           const float* d, int64_t n, float* pdf, cudaStream_t stream, const VmfFrame* frame = nullptr);
}  // namespace nemit
