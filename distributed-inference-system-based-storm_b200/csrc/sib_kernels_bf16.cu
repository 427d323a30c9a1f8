// sib_kernels_bf16.cu -- placeholder until the tensor-core kernel lands: reports "unsupported" loudly.
#include <cuda_runtime.h>

#include "sib_internal.hpp"

namespace sib {
bool bf16_supported(const ModelDesc&) { return false; }
int bf16_prepare(const ModelDesc&, Bf16Weights*) { set_error("BF16 path not built"); return SIB_EINVAL; }
void bf16_release(Bf16Weights* w) { if (w && w->d_packed) cudaFree(w->d_packed); if (w) *w = Bf16Weights{}; }
int launch_forward_bf16(const Bf16Weights&, const DeviceModel&, const float*, int, float*, float*, int32_t*,
                        GuardScratch*, float, void*) { return (int)cudaErrorNotSupported; }
}  // namespace sib
