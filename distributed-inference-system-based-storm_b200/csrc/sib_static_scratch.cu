// sib_static_scratch.cu -- a forward entry point that performs no device allocation of its own.
//
// Who needs it. sib_forward_device() keeps the margin guard's scratch (per-image margins, the compaction list of
// low-margin images, its counter) in memory the engine cudaMalloc'ed for itself at open time, one set per engine.
// Some users of this library drive the kernels from places where asking the driver for more device memory is not
// an option:
//   * a bolt task that attaches to the GPU after the worker's start-up phase, in a process whose device memory the
//     operator has already handed out in full to the other executors' engines -- cudaMalloc would simply fail there;
//   * a latency-critical executor loop sharing its GPU with up to eight other executors: cudaMalloc and cudaFree
//     synchronise the whole device, so one call would stall every other executor's streams;
//   * code that runs between "everything is set up" and shutdown under a no-allocation rule (real-time or
//     certification style), where only pre-existing memory may be touched.
// For them the scratch lives in the library image itself: a file-scope __device__ array that every GPU which loads
// libsib.so carries from load to exit (kStaticBytes, about 128 KB). The caller passes inputs and outputs as usual.
//
// Things to know (also in include/sib.h):
//   * capacity is fixed: at most SIB_STATIC_MAX_IMAGES (16384) images per call, larger n is refused with SIB_ECAP;
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that first
//     call before starting any stream capture on that device;
//   * every call on a device uses the same array and finds the previous call's bytes in it; the counter is zeroed at
//     the start of each call (stream-ordered), the rest is fully overwritten before it is read;
//   * because the array is shared, two calls whose GPU work overlaps (different streams, or two engines on one
//     device) race on it -- order them on one stream or with events.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "sib_internal.hpp"

namespace {

constexpr int kStaticMaxImages = SIB_STATIC_MAX_IMAGES;
constexpr size_t kMarginOff = 0;
constexpr size_t kListOff = kMarginOff + sizeof(float) * kStaticMaxImages;
constexpr size_t kCountOff = kListOff + sizeof(int32_t) * kStaticMaxImages;
constexpr size_t kStaticBytes = kCountOff + 256;

// The scratch itself. In host code this name is only a handle for cudaGetSymbolAddress, never an address to pass on.
static __device__ __align__(256) unsigned char g_guard_scratch[kStaticBytes];

std::mutex g_mu;
unsigned char* g_dev_ptr[64] = {};  // device address of g_guard_scratch, per CUDA ordinal

// Returns the device address of the array on `device` (which must be current), resolving it on first use.
int scratch_for(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (device < 0 || device >= 64) { sib::set_error("device ordinal out of range for the static scratch"); return SIB_EINVAL; }
    if (!g_dev_ptr[device]) {
        unsigned char* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_guard_scratch);
        if (e != cudaSuccess || !p) {
            sib::set_error(std::string("cudaGetSymbolAddress(static guard scratch): ") + cudaGetErrorString(e));
            return SIB_ECUDA;
        }
        g_dev_ptr[device] = p;
    }
    *out = g_dev_ptr[device];
    return SIB_OK;
}

}  // namespace

extern "C" int sib_forward_device_static(sib_engine* e, const float* d_nhwc, int n, float* d_probs, float* d_logits,
                                         int32_t* d_argmax, void* stream) {
    if (!e || !d_nhwc || !d_probs || !d_argmax || n < 0) return SIB_EINVAL;
    if (n > kStaticMaxImages) {
        sib::set_error("sib_forward_device_static: n = " + std::to_string(n) + " exceeds the static scratch capacity of " +
                       std::to_string(kStaticMaxImages) + " images per call; split the batch or use sib_forward_device");
        return SIB_ECAP;
    }
    sib::EngineView v;
    if (!sib::engine_view(e, &v)) return SIB_EINVAL;
    if (v.precision == SIB_PREC_FP32 || !v.guard) {
        sib::set_error("sib_forward_device_static serves the guarded tensor-core path (SIB_PREC_FP16/BF16 with guard on); "
                       "the fp32 path uses no scratch, call sib_forward_device");
        return SIB_EINVAL;
    }
    if (v.preset != 0 && v.preset != 1) {
        sib::set_error("sib_forward_device_static serves the LeNet-5 presets' fused kernel; this model runs on the any-shape kernel, call sib_forward_device");
        return SIB_EINVAL;
    }
    if (n == 0) return SIB_OK;
    if (cudaSetDevice(v.device) != cudaSuccess) { sib::set_error("cudaSetDevice failed"); return SIB_ECUDA; }
    unsigned char* base = nullptr;
    int rc = scratch_for(v.device, &base);
    if (rc != SIB_OK) return rc;
    sib::GuardScratch g;
    g.d_margin = reinterpret_cast<float*>(base + kMarginOff);
    g.d_list = reinterpret_cast<int32_t*>(base + kListOff);
    g.d_count = reinterpret_cast<int32_t*>(base + kCountOff);  // zeroed by launch_forward_tc at the head of the call
    int ce = sib::launch_forward_tc(*v.tc, d_nhwc, n, d_probs, d_logits, d_argmax, &g, v.tau, nullptr, stream);
    if (!ce) ce = sib::launch_guard_fp32(v.preset, *v.dm, d_nhwc, n, g, d_probs, d_logits, d_argmax, stream);
    if (ce) {
        sib::set_error(std::string("sib_forward_device_static: ") + cudaGetErrorString((cudaError_t)ce));
        return SIB_ECUDA;
    }
    return SIB_OK;
}
