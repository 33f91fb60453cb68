// temporary: jvp / backward entry points until their kernels land
#include "kernels.h"
#include "vmf.h"
namespace nemit { size_t backward_chunk_bytes(int64_t) { return 0; } }
extern "C" int nemit_jvp(NemitHandle*, const NemitRays*, const float*, const float*, float*, void*, size_t, void*) { return NEMIT_E_INVALID; }
extern "C" int nemit_backward(NemitHandle*, const NemitRays*, const float*, int64_t, float, int32_t, float*, float*, void*, size_t, void*) { return NEMIT_E_INVALID; }
