// sib_internal.hpp -- shared declarations inside libsib.so (not part of the C ABI).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <atomic>
#include <mutex>
#include <vector>

#include "../../include/sib.h"

namespace sib {

constexpr uint32_t kKindConv = 1;   // conv(kh,kw) VALID s1 + bias + ReLU + maxpool 2x2/2
constexpr uint32_t kKindDense = 2;  // x.W + b (+ReLU when flags&1)
constexpr int kMaxLayers = 16;
constexpr int kMaxClasses = 32;

struct LayerDesc {
    uint32_t kind, a, b, c, d, flags, w_off, b_off;  // conv: a=kh b=kw c=cin d=cout; dense: a=in b=out
};

// Parsed model blob: stands where the reference keeps model/saved_model.pb (InferenceBolt.java:49).
struct ModelDesc {
    uint32_t in_h = 0, in_w = 0, in_c = 0, n_layers = 0, n_classes = 0, n_weights = 0;
    LayerDesc layer[kMaxLayers]{};
    std::vector<float> weights;
    uint32_t img_elems = 0;
    uint32_t max_store_act = 0;  // largest activation kept between layers (input, pooled maps, dense outputs)
    int64_t flops_per_image = 0;
    int preset = -1;  // 0 mnist-lenet5, 1 cifar-lenet5, 2 mnist-wide, 3 cifar-wide, -1 = other shape
};

int parse_model(const void* blob, size_t n, ModelDesc* out, std::string* err);  // blob, or saved_model.pb bytes
// TensorFlow SavedModel / frozen GraphDef bytes -> model blob (sib_pbimport.cpp)
int import_pb(const void* bytes, size_t n, std::vector<uint8_t>* blob, std::string* err);

void set_error(const std::string& s);

// ---- codec (sib_json.cpp)
long json_decode(const char* utf8, size_t len, float* out, long cap, int64_t dims[4]);
long json_encode(const float* probs, long n, int n_classes, char* buf, size_t cap);
int float_to_string(float x, char* buf);

// Per-device bookkeeping for cudaFuncAttributeMaxDynamicSharedMemorySize of ONE kernel (one static instance per
// kernel instantiation). Engines on several host threads launch concurrently: the limit only ever grows, and the
// attribute call and the record of it happen under one lock, so a launch never finds the limit below what it needs.
struct SmemLimit {
    std::atomic<int> have[64];
    std::mutex mu;
    SmemLimit() { for (auto& h : have) h.store(-1, std::memory_order_relaxed); }
    // returns cudaSuccess (0) or the cudaError_t of the attribute call; `set` performs it for `bytes`
    template <typename Set>
    int ensure(int device, int bytes, Set set) {
        std::atomic<int>& h = have[device & 63];
        if (h.load(std::memory_order_acquire) >= bytes) return 0;
        std::lock_guard<std::mutex> lk(mu);
        if (h.load(std::memory_order_relaxed) >= bytes) return 0;
        const int rc = set(bytes);
        if (rc == 0) h.store(bytes, std::memory_order_release);
        return rc;
    }
};

// ---- device side (sib_kernels_*.cu)
struct DeviceLayer {
    uint32_t kind, a, b, c, d, flags, w_off, b_off;
};
struct DeviceModel {
    uint32_t in_h, in_w, in_c, n_layers, n_classes, img_elems, max_store_act, pad;
    DeviceLayer layer[kMaxLayers];
    const float* weights;  // device pointer, fp32, blob order
};

// fp32 CUDA-core forward. d_idx (may be null): run only images d_idx[0..n) (guard recompute), writing in place.
// Any of d_logits/d_argmax/d_margin may be null. Returns cudaError_t as int.
int launch_forward_fp32(const DeviceModel& m, const float* d_in, int n, const int32_t* d_idx, const int32_t* d_count,
                        float* d_probs, float* d_logits, int32_t* d_argmax, float* d_margin, void* stream);
int fp32_smem_bytes(const DeviceModel& m);

// Tensor-core whole-network kernel (sib_kernels_tc.cu). Scratch for the margin guard: per-image top1-top2 logit
// margin, the compacted list of images below tau, and its length.
struct GuardScratch {
    float* d_margin = nullptr;
    int32_t* d_list = nullptr;
    int32_t* d_count = nullptr;      // [0] list length, [1] ticket of the guard kernel's CTAs (self_reset)
    int32_t* d_out_count = nullptr;  // self_reset: where the guard kernel leaves the launch's list length (may be null)
    // true: the guard kernel's last CTA clears the counter for the next launch (no cudaMemsetAsync between batches);
    // false: launch_forward_tc clears it stream-ordered at the head of every call (callers that share the scratch)
    bool self_reset = false;
};
// Any-shape tensor-core kernel (sib_kernels_tcgen.cu): what the launch needs to know about the model.
struct TcgLayer {
    uint32_t kind, a, b, c, d, flags, w_off, b_off;  // as LayerDesc; w_off / b_off index the packed image (32-bit words)
    int ks, nt;                                      // K steps of 16 (per kernel row for conv) and N tiles of 8
};
struct TcgPlan {
    int in_h, in_w, in_c, n_layers, n_classes, img_elems;
    int act_stride;  // halves per image and activation buffer (largest kept activation + slack)
    int wsm_words;   // shared memory for the largest conv layer's fragments
    int smem_bytes;
    TcgLayer layer[kMaxLayers];
};
struct TcWeights {
    void* d_packed = nullptr;  // weights pre-swizzled into mma fragment order (half or bf16) + fp32 biases
    size_t bytes = 0;
    int preset = -1;
    int precision = SIB_PREC_FP16;
    bool generic = false;      // true: the any-shape kernel (sib_kernels_tcgen.cu) with `plan`; false: the fused LeNet-5 kernel
    TcgPlan plan{};
};
bool tcg_plan(const ModelDesc& m, TcgPlan* pl, std::string* why);
std::vector<uint32_t> tcg_pack(const ModelDesc& m, int precision, TcgPlan* pl);
int launch_forward_tcgen(const TcWeights& w, const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax,
                         GuardScratch* guard, float tau, void* stream);
bool tc_supported(const ModelDesc& m);
int tc_prepare(const ModelDesc& m, int precision, TcWeights* out);  // uploads to the current device
void tc_release(TcWeights* w);
size_t tc_debug_floats(int preset);
int launch_forward_tc(const TcWeights& w, const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax,
                      GuardScratch* guard, float tau, float* d_dbg, void* stream);

// GPU-side JSON decode (sib_kernels_json.cu): one CTA per tuple; desc = {text offset, length, image index, 0} x n.
size_t json_gpu_max_text();
int launch_json_decode(const char* d_text, const int32_t* d_desc, int n_tuples, float* d_out, int32_t* d_flags, int H,
                       int W, int C, void* stream);

// Read-only view of an open engine for translation units that add entry points next to sib_engine.cpp.
struct EngineView {
    int device = 0, precision = 0, guard = 0, preset = -1;
    float tau = 0.f;
    const TcWeights* tc = nullptr;
    const DeviceModel* dm = nullptr;
};
bool engine_view(sib_engine* e, EngineView* out);

// fp32 recompute of the listed images, one image per CTA (sib_kernels_guard.cu). LeNet-5 presets only.
int launch_guard_fp32(int preset, const DeviceModel& m, const float* d_in, int n, const GuardScratch& g, float* d_probs,
                      float* d_logits, int32_t* d_argmax, void* stream);

}  // namespace sib
