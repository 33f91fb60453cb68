// vmf.h - launchers of the vMF-mixture kernels (vmf.cu) and backward/jvp entry points (internal)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace nemit {
size_t backward_chunk_bytes(int64_t chunk_rays);
cudaError_t launch_vmf_sample(const float* pos, const float* weight, const float* std, int K, const float* p,
                              const float* u, int64_t n, float* d, float* pdf, cudaStream_t stream);
cudaError_t launch_vmf_pdf(const float* pos, const float* weight, const float* std, int K, const float* p,
                           const float* d, int64_t n, float* pdf, cudaStream_t stream);
}  // namespace nemit
