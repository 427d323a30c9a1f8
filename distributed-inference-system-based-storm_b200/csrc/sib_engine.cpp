// sib_engine.cpp -- the per-executor engine behind the C ABI: device selection, weight upload, and the pinned-host
// micro-batcher keyed on batch size and deadline.
//
// What it replaces in the reference (file:line relative to src/main/java/dke/model/):
//   prepare(): SavedModelBundle.load + session()                     InferenceBolt.java:43-62  -> Engine::open
//   execute(): one Session.run per tuple, strictly serial, batch = 1  InferenceBolt.java:69-100 -> submit_* / poll
// Tuples from one executor are packed into fp32 NHWC batches in pinned host slots; a slot is launched when it holds
// max_batch images (flush-by-size) or when its oldest tuple is deadline_us old (flush-by-deadline, helper thread).
// Each launch is H2D copy -> forward kernel(s) -> D2H copy on the slot's own stream, so consecutive slots overlap
// copy and compute. Results come back in submission order, one per input tuple (emit+ack order, :98-99).
// There is no CPU fallback: without a CUDA device open() fails with SIB_ECUDA.
#include <cuda_runtime_api.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>

#include "sib_internal.hpp"

namespace sib {

using Clock = std::chrono::steady_clock;

#define CU_TRY(expr)                                                                                  \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                            \
            return SIB_ECUDA;                                                                         \
        }                                                                                             \
    } while (0)

struct Entry {
    uint64_t tag;
    int n, offset;
    int text = -1;        // index into the slot's text-tuple table when the tuple went in as raw JSON
    int side = -1;        // index into the slot's side results (tuples the host decoder had to redo)
    int status = SIB_OK;  // per-tuple status (json_on_gpu: SIB_EJSON / SIB_ESHAPE found after the fact)
};

struct SideResult {
    std::vector<float> probs;
    std::vector<int32_t> argmax;
};

struct DevBufs {  // one forward's worth of device memory
    float* d_in = nullptr;
    float* d_probs = nullptr;
    float* d_logits = nullptr;
    int32_t* d_argmax = nullptr;
    GuardScratch guard;
};

// An atomic that a std::vector element may hold: copies copy the value (slots are only copied while the engine is built).
template <typename T>
struct Atom {
    std::atomic<T> v;
    Atom(T x = T()) : v(x) {}
    Atom(const Atom& o) : v(o.v.load()) {}
    Atom& operator=(const Atom& o) { v.store(o.v.load()); return *this; }
    Atom& operator=(T x) { v.store(x); return *this; }
    operator T() const { return v.load(); }
};

struct Slot {
    // FILLING: accepting tuples. SEALED: closed to new tuples and queued for launch, but some submitter is still
    // copying its images in (writers > 0). INFLIGHT: copies and kernels issued. READY: results in h_probs.
    // HANDED: results given out by sib_poll, slot recycled at the next poll.
    // REDO: complete on the GPU, a poller is re-reading the tuples the decode kernel declined (engine lock released).
    enum State { FREE, FILLING, SEALED, INFLIGHT, REDO, READY, HANDED };
    Atom<int> state{FREE};       // written under the engine lock; the last writer of a tuple reads it without (see written())
    float* h_in = nullptr;       // pinned, max_batch x img
    float* h_probs = nullptr;    // pinned, max_batch x ncls
    int32_t* h_argmax = nullptr; // pinned
    int32_t* h_guard_n = nullptr; // pinned, 1
    int32_t* r_argmax = nullptr;  // where the last launch's argmax / guard count landed (inside the h_probs block)
    int32_t* r_guard_n = nullptr;
    DevBufs dev;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    std::vector<Entry> entries;
    size_t next_entry = 0;
    int images = 0;
    int status = SIB_OK;
    // raw-JSON tuples (json_on_gpu): text bytes + {offset, length, image index, 0} per tuple, decoded by a kernel
    char* h_text = nullptr;
    char* d_text = nullptr;
    int32_t* h_tdesc = nullptr;
    int32_t* d_tdesc = nullptr;
    int32_t* h_tflags = nullptr;
    int32_t* d_tflags = nullptr;
    size_t text_cap = 0, text_used = 0;
    int n_text = 0;
    std::vector<SideResult> side;
    std::vector<std::pair<int, int>> staged;  // (offset, n) image ranges that sit in h_in and need the H2D at launch
    size_t direct_bytes = 0;                  // bytes already sent device-ward straight from caller-pinned memory
    Atom<int> writers{0};  // submitters that reserved room and are copying outside the lock; they leave without it
    int flush_reason = 0;  // why it was sealed: 0 size, 1 deadline, 2 explicit
    Clock::time_point first_ts;
};

}  // namespace sib

using namespace sib;

struct sib_engine {
    ModelDesc model;
    DeviceModel dm{};
    sib_config cfg{};
    int device = 0;
    float* d_weights = nullptr;
    TcWeights tc{};
    std::vector<Slot> slots;
    int filling = -1;
    std::deque<int> fifo;  // sealed/launched slots, submission order
    std::mutex mu;
    std::condition_variable cv;
    std::thread deadline_thread;
    Atom<bool> closing{false};
    sib_stats stats{};
    // synchronous path
    std::mutex sync_mu;
    Slot sync_slot;
    DevBufs dev_scratch;  // guard scratch for sib_forward_device

    int alloc_dev(DevBufs& d, bool with_io);
    void free_dev(DevBufs& d);
    int alloc_slot(Slot& s);
    void free_slot(Slot& s);
    int run_forward(const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax, GuardScratch* g,
                    cudaStream_t st);
    void redo_flagged(Slot& s);           // json_on_gpu: tuples the decode kernel declined go through the host decoder (no lock held)
    std::atomic<uint64_t> launches{0};    // kernels launched; bumped from the batch path and the synchronous paths alike
    int seal_locked(int si, int reason);  // stop filling; launches at once unless a writer is still copying
    int launch_locked(int si);
    void deadline_main();
};

int sib_engine::alloc_dev(DevBufs& d, bool with_io) {
    const size_t mb = (size_t)cfg.max_batch, nc = model.n_classes;
    if (with_io) {
        CU_TRY(cudaMalloc((void**)&d.d_in, mb * model.img_elems * sizeof(float)));
        // room for the packed result block of a launch: probs[n][nc] | argmax[n] | guard count (one D2H copy)
        CU_TRY(cudaMalloc((void**)&d.d_probs, (mb * (nc + 1) + 4) * sizeof(float)));
        CU_TRY(cudaMalloc((void**)&d.d_logits, mb * nc * sizeof(float)));
        CU_TRY(cudaMalloc((void**)&d.d_argmax, mb * sizeof(int32_t)));
    }
    if (cfg.precision != SIB_PREC_FP32) {
        CU_TRY(cudaMalloc((void**)&d.guard.d_margin, mb * sizeof(float)));
        CU_TRY(cudaMalloc((void**)&d.guard.d_list, mb * sizeof(int32_t)));
        CU_TRY(cudaMalloc((void**)&d.guard.d_count, 2 * sizeof(int32_t)));
        CU_TRY(cudaMemset(d.guard.d_count, 0, 2 * sizeof(int32_t)));
        d.guard.self_reset = true;  // one stream per scratch: the guard kernel clears the counter behind itself
    }
    return SIB_OK;
}

void sib_engine::free_dev(DevBufs& d) {
    cudaFree(d.d_in); cudaFree(d.d_probs); cudaFree(d.d_logits); cudaFree(d.d_argmax);
    cudaFree(d.guard.d_margin); cudaFree(d.guard.d_list); cudaFree(d.guard.d_count);
    d = DevBufs{};
}

int sib_engine::alloc_slot(Slot& s) {
    const size_t mb = (size_t)cfg.max_batch, nc = model.n_classes;
    CU_TRY(cudaHostAlloc((void**)&s.h_in, mb * model.img_elems * sizeof(float), cudaHostAllocDefault));
    CU_TRY(cudaHostAlloc((void**)&s.h_probs, (mb * (nc + 1) + 4) * sizeof(float), cudaHostAllocDefault));
    CU_TRY(cudaHostAlloc((void**)&s.h_argmax, mb * sizeof(int32_t), cudaHostAllocDefault));
    CU_TRY(cudaHostAlloc((void**)&s.h_guard_n, 2 * sizeof(int32_t), cudaHostAllocDefault));
    s.h_guard_n[0] = 0;
    s.r_argmax = s.h_argmax;
    s.r_guard_n = s.h_guard_n;
    int r = alloc_dev(s.dev, true);
    if (r) return r;
    if (cfg.json_on_gpu) {
        // room for ~14 characters per pixel value; a slot also seals when its text area is full
        const size_t per_img = std::min<size_t>(json_gpu_max_text(), 16 + 14 * (size_t)model.img_elems);
        s.text_cap = std::min<size_t>(mb, 8192) * per_img + 64;
        CU_TRY(cudaHostAlloc((void**)&s.h_text, s.text_cap, cudaHostAllocDefault));
        CU_TRY(cudaMalloc((void**)&s.d_text, s.text_cap));
        CU_TRY(cudaHostAlloc((void**)&s.h_tdesc, mb * 4 * sizeof(int32_t), cudaHostAllocDefault));
        CU_TRY(cudaMalloc((void**)&s.d_tdesc, mb * 4 * sizeof(int32_t)));
        CU_TRY(cudaHostAlloc((void**)&s.h_tflags, mb * sizeof(int32_t), cudaHostAllocDefault));
        CU_TRY(cudaMalloc((void**)&s.d_tflags, mb * sizeof(int32_t)));
    }
    CU_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CU_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    s.entries.reserve(256);
    return SIB_OK;
}

void sib_engine::free_slot(Slot& s) {
    if (s.stream) cudaStreamSynchronize(s.stream);
    cudaFreeHost(s.h_in); cudaFreeHost(s.h_probs); cudaFreeHost(s.h_argmax); cudaFreeHost(s.h_guard_n);
    cudaFreeHost(s.h_text); cudaFreeHost(s.h_tdesc); cudaFreeHost(s.h_tflags);
    cudaFree(s.d_text); cudaFree(s.d_tdesc); cudaFree(s.d_tflags);
    free_dev(s.dev);
    if (s.done) cudaEventDestroy(s.done);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Slot{};
}

// One forward pass on device-resident data. FP16/BF16: tensor-core kernel, then the fp32 guard over the
// low-margin images it listed (same stream, so the list and its count are ready).
int sib_engine::run_forward(const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax,
                            GuardScratch* g, cudaStream_t st) {
    int e;
    if (cfg.precision == SIB_PREC_FP32) {
        e = launch_forward_fp32(dm, d_in, n, nullptr, nullptr, d_probs, d_logits, d_argmax, nullptr, st);
        launches.fetch_add(1, std::memory_order_relaxed);
    } else {
        const bool guard = cfg.guard != 0;
        e = launch_forward_tc(tc, d_in, n, d_probs, d_logits, d_argmax, guard ? g : nullptr, cfg.guard_tau, nullptr, st);
        launches.fetch_add(1, std::memory_order_relaxed);
        if (!e && guard && !tc.generic) {
            e = launch_guard_fp32(model.preset, dm, d_in, n, *g, d_probs, d_logits, d_argmax, st);
            launches.fetch_add(1, std::memory_order_relaxed);
        } else if (!e && guard) {
            // any-shape models: the listed images go through the generic fp32 kernel; it does not know the counter
            // protocol of guard_fp32_kernel, so the length is handed on and cleared by two small stream-ordered operations
            e = launch_forward_fp32(dm, d_in, n, g->d_list, g->d_count, d_probs, d_logits, d_argmax, nullptr, st);
            launches.fetch_add(1, std::memory_order_relaxed);
            if (!e && g->d_out_count) e = (int)cudaMemcpyAsync(g->d_out_count, g->d_count, sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
            if (!e) e = (int)cudaMemsetAsync(g->d_count, 0, sizeof(int32_t), st);
        }
    }
    if (e) {
        set_error(std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)e));
        return SIB_ECUDA;
    }
    return SIB_OK;
}

int sib_engine::seal_locked(int si, int reason) {
    Slot& s = slots[si];
    s.state = Slot::SEALED;
    s.flush_reason = reason;
    s.next_entry = 0;
    fifo.push_back(si);  // result order is fixed here, at seal time
    if (filling == si) filling = -1;
    if (s.writers == 0) return launch_locked(si);
    return SIB_OK;  // the last writer launches it
}

int sib_engine::launch_locked(int si) {
    Slot& s = slots[si];
    const int reason = s.flush_reason;
    cudaSetDevice(device);
    const size_t nc = model.n_classes;
    int rc = SIB_OK;
    auto chk = [&](cudaError_t e, const char* what) {
        if (e != cudaSuccess && rc == SIB_OK) {
            set_error(std::string(what) + ": " + cudaGetErrorString(e));
            rc = SIB_ECUDA;
        }
    };
    const size_t img_bytes = (size_t)model.img_elems * sizeof(float);
    size_t in_bytes = s.direct_bytes;
    for (auto& r : s.staged) {  // one copy per contiguous staged range (normally exactly one)
        chk(cudaMemcpyAsync(s.dev.d_in + (size_t)r.first * model.img_elems, s.h_in + (size_t)r.first * model.img_elems,
                            (size_t)r.second * img_bytes, cudaMemcpyHostToDevice, s.stream), "H2D");
        in_bytes += (size_t)r.second * img_bytes;
    }
    if (s.n_text > 0) {  // raw JSON tuples: ship the text, decode it on the device into d_in
        chk(cudaMemcpyAsync(s.d_text, s.h_text, s.text_used + 16, cudaMemcpyHostToDevice, s.stream), "H2D text");
        chk(cudaMemcpyAsync(s.d_tdesc, s.h_tdesc, (size_t)s.n_text * 4 * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream), "H2D text table");
        in_bytes += s.text_used + (size_t)s.n_text * 16;
        if (rc == SIB_OK) {
            int ke = launch_json_decode(s.d_text, s.d_tdesc, s.n_text, s.dev.d_in, s.d_tflags, (int)model.in_h, (int)model.in_w, (int)model.in_c, s.stream);
            launches.fetch_add(1, std::memory_order_relaxed);
            if (ke) { set_error(std::string("json decode kernel: ") + cudaGetErrorString((cudaError_t)ke)); rc = SIB_ECUDA; }
        }
        chk(cudaMemcpyAsync(s.h_tflags, s.d_tflags, (size_t)s.n_text * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream), "D2H text flags");
    }
    // Results of this launch are packed behind one another on the device -- probs[n][nc] | argmax[n] | guard count --
    // so that they come home in one copy (three small D2H copies were a fifth of a single tuple's round trip).
    const size_t n_img = (size_t)s.images;
    int32_t* d_arg = reinterpret_cast<int32_t*>(s.dev.d_probs + n_img * nc);
    GuardScratch g = s.dev.guard;
    g.d_out_count = d_arg + n_img;  // the guard kernel leaves the list length here and clears its own counter
    const bool guarded = cfg.precision != SIB_PREC_FP32 && cfg.guard;
    if (rc == SIB_OK) rc = run_forward(s.dev.d_in, s.images, s.dev.d_probs, nullptr, d_arg, &g, s.stream);
    const size_t d2h = (n_img * (nc + 1) + (guarded ? 1 : 0)) * sizeof(float);
    chk(cudaMemcpyAsync(s.h_probs, s.dev.d_probs, d2h, cudaMemcpyDeviceToHost, s.stream), "D2H results");
    s.r_argmax = reinterpret_cast<int32_t*>(s.h_probs + n_img * nc);
    s.r_guard_n = s.r_argmax + n_img;
    chk(cudaEventRecord(s.done, s.stream), "event record");
    s.status = rc;
    s.state = Slot::INFLIGHT;
    stats.batches++;
    stats.h2d_bytes += in_bytes;
    stats.d2h_bytes += d2h;
    if (reason == 0) stats.flush_by_size++;
    else if (reason == 1) stats.flush_by_deadline++;
    else stats.flush_explicit++;
    int b = 0;
    while (b < 15 && (1 << (b + 1)) <= s.images) b++;
    stats.batch_hist[b]++;
    cv.notify_all();
    return rc;
}

extern "C" int sib_forward_sync(sib_engine* e, const float* x, int n, float* probs, float* logits, int32_t* argmax);

// The decode kernel only takes one-instance tuples in canonical form; whatever it flagged is read again here with
// the host decoder, which owns the complete Jackson semantics (leniencies, N > 1, error classes). Rare path, but a
// producer that writes 17-digit doubles sends every tuple down it: the slot is parked in state REDO and the engine
// lock is NOT held here, so submitters and the deadline thread go on, and all flagged tuples of the slot share one
// forward pass instead of a synchronous round trip each.
void sib_engine::redo_flagged(Slot& s) {
    const size_t nc = model.n_classes, img = model.img_elems;
    std::vector<float> tmp, all;
    struct Todo { Entry* en; int n; size_t at; };
    std::vector<Todo> todo;
    uint64_t gpu_ok = 0, redone = 0;
    for (auto& en : s.entries) {
        if (en.text < 0) continue;
        if (!s.h_tflags[en.text]) { gpu_ok++; continue; }
        redone++;
        const char* txt = s.h_text + s.h_tdesc[4 * en.text];
        const size_t len = (size_t)s.h_tdesc[4 * en.text + 1];
        const long cap = (long)(len / 2 + 8);
        if ((long)tmp.size() < cap) tmp.resize((size_t)cap);
        int64_t dims[4] = {0, 0, 0, 0};
        const long nf = json_decode(txt, len, tmp.data(), cap, dims);
        if (nf == SIB_EJSON) { en.n = 0; en.status = SIB_EJSON; continue; }
        if (nf < 0 || dims[1] != model.in_h || dims[2] != model.in_w || dims[3] != model.in_c || dims[0] > cfg.max_batch) {
            en.n = 0; en.status = SIB_ESHAPE; continue;
        }
        todo.push_back(Todo{&en, (int)dims[0], all.size()});
        all.insert(all.end(), tmp.begin(), tmp.begin() + (size_t)dims[0] * img);
    }
    int total = 0;
    for (auto& t : todo) total += t.n;
    std::vector<float> probs((size_t)total * nc);
    std::vector<int32_t> arg((size_t)total);
    const bool ok = total == 0 || sib_forward_sync(this, all.data(), total, probs.data(), nullptr, arg.data()) == SIB_OK;
    for (auto& t : todo) {
        if (!ok) { t.en->n = 0; t.en->status = SIB_ECUDA; continue; }
        const size_t first = t.at / img;
        SideResult sr;
        sr.probs.assign(probs.begin() + first * nc, probs.begin() + (first + (size_t)t.n) * nc);
        sr.argmax.assign(arg.begin() + first, arg.begin() + first + (size_t)t.n);
        t.en->n = t.n;
        t.en->side = (int)s.side.size();
        s.side.push_back(std::move(sr));
    }
    std::lock_guard<std::mutex> lk(mu);
    stats.json_gpu_decoded += gpu_ok;
    stats.json_host_redone += redone;
}

void sib_engine::deadline_main() {
    cudaSetDevice(device);
    std::unique_lock<std::mutex> lk(mu);
    while (!closing) {
        if (filling < 0 || slots[filling].images == 0) {
            cv.wait(lk);
            continue;
        }
        const int si = filling;
        const auto due = slots[si].first_ts + std::chrono::microseconds(cfg.deadline_us);
        if (Clock::now() >= due) {
            seal_locked(si, 1);
            continue;
        }
        cv.wait_until(lk, due);
    }
}

// ------------------------------------------------------------------------------------------------ C ABI

extern "C" int sib_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int sib_model_describe(const void* blob, size_t n, sib_model_info* out) {
    ModelDesc m;
    std::string err;
    int r = parse_model(blob, n, &m, &err);
    if (r) { set_error(err); return r; }
    if (out) {
        out->in_h = m.in_h; out->in_w = m.in_w; out->in_c = m.in_c; out->n_layers = m.n_layers;
        out->n_classes = m.n_classes; out->n_weights = m.n_weights;
        out->flops_per_image = m.flops_per_image;
        out->bytes_per_image = (int64_t)m.img_elems * 4 + (int64_t)m.n_classes * 4;
    }
    return SIB_OK;
}

extern "C" long sib_model_import_pb(const void* pb, size_t n, void* out, size_t cap) {
    std::vector<uint8_t> blob;
    std::string err;
    const int r = import_pb(pb, n, &blob, &err);
    if (r) { set_error(err); return r; }
    if (!out) return (long)blob.size();
    if (cap < blob.size()) { set_error("sib_model_import_pb: output buffer too small"); return SIB_ECAP; }
    std::memcpy(out, blob.data(), blob.size());
    return (long)blob.size();
}

extern "C" int sib_engine_open(const void* blob, size_t n, const sib_config* ucfg, sib_engine** out) {
    if (!out) return SIB_EINVAL;
    *out = nullptr;
    sib_engine* e = new sib_engine();
    std::string err;
    int r = parse_model(blob, n, &e->model, &err);
    if (r) { set_error(err); delete e; return r; }
    sib_config c{};
    if (ucfg) c = *ucfg;
    else { c.device = -1; c.guard = 1; }
    if (c.max_batch <= 0) c.max_batch = 4096;
    if (c.deadline_us <= 0) c.deadline_us = 200;
    if (c.n_slots < 2) c.n_slots = 4;
    c.json_on_gpu = c.json_on_gpu == 2 ? 2 : c.json_on_gpu ? 1 : 0;
    c.strict_batching = c.strict_batching ? 1 : 0;
    if (c.precision != SIB_PREC_FP16 && c.precision != SIB_PREC_BF16 && c.precision != SIB_PREC_FP32) {
        set_error("precision must be SIB_PREC_FP16, SIB_PREC_BF16 or SIB_PREC_FP32");
        delete e;
        return SIB_EINVAL;
    }
    // guard threshold: a bit over twice the largest logit error seen for the operand type (DESIGN.md "margin guard")
    if (c.guard_tau <= 0.0f) c.guard_tau = c.precision == SIB_PREC_BF16 ? 0.30f : 0.05f;
    int ndev = sib_device_count();
    if (ndev <= 0) {
        set_error("no CUDA device visible: libsib has no CPU fallback");
        delete e;
        return SIB_ECUDA;
    }
    if (c.device < 0) c.device = (c.task_index < 0 ? 0 : c.task_index) % ndev;  // executor -> GPU, MainTopology.java:62
    if (c.device >= ndev) { set_error("device ordinal out of range"); delete e; return SIB_EINVAL; }
    if (c.precision != SIB_PREC_FP32 && !tc_supported(e->model)) {
        set_error("this model's shapes do not fit the tensor-core kernels (conv output channels even and <= 64, dense input width even, <= 32 classes, 8 images of activations within shared memory); open it with SIB_PREC_FP32");
        delete e;
        return SIB_EINVAL;
    }
    e->cfg = c;
    e->device = c.device;
    auto fail = [&](int code) { sib_engine_close(e); return code; };
    if (cudaSetDevice(e->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return fail(SIB_ECUDA); }
    const ModelDesc& m = e->model;
    if (cudaMalloc((void**)&e->d_weights, m.weights.size() * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(e->d_weights, m.weights.data(), m.weights.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("weight upload failed");
        return fail(SIB_ECUDA);
    }
    DeviceModel& dm = e->dm;
    dm.in_h = m.in_h; dm.in_w = m.in_w; dm.in_c = m.in_c; dm.n_layers = m.n_layers; dm.n_classes = m.n_classes;
    dm.img_elems = m.img_elems; dm.max_store_act = m.max_store_act;
    std::memcpy(dm.layer, m.layer, sizeof(LayerDesc) * kMaxLayers);
    dm.weights = e->d_weights;
    if (fp32_smem_bytes(dm) > 227 * 1024) { set_error("model activations exceed shared memory"); return fail(SIB_EINVAL); }
    if (c.precision != SIB_PREC_FP32) {
        r = tc_prepare(m, c.precision, &e->tc);
        if (r) return fail(r);
    }
    e->slots.assign((size_t)c.n_slots, Slot());
    for (auto& s : e->slots)
        if ((r = e->alloc_slot(s)) != SIB_OK) return fail(r);
    if ((r = e->alloc_slot(e->sync_slot)) != SIB_OK) return fail(r);
    if ((r = e->alloc_dev(e->dev_scratch, false)) != SIB_OK) return fail(r);
    e->deadline_thread = std::thread([e] { e->deadline_main(); });
    *out = e;
    return SIB_OK;
}

extern "C" int sib_engine_close(sib_engine* e) {
    if (!e) return SIB_EINVAL;
    {
        std::unique_lock<std::mutex> lk(e->mu);
        if (e->filling >= 0 && e->slots[e->filling].images > 0) e->seal_locked(e->filling, 2);
        e->closing = true;
        e->cv.notify_all();
        // submitters still copying into sealed slots finish first (they launch the slot themselves)
        e->cv.wait(lk, [e] {
            for (auto& s : e->slots)
                if (s.writers > 0) return false;
            return true;
        });
    }
    if (e->deadline_thread.joinable()) e->deadline_thread.join();
    cudaSetDevice(e->device);
    for (auto& s : e->slots) e->free_slot(s);
    e->free_slot(e->sync_slot);
    e->free_dev(e->dev_scratch);
    tc_release(&e->tc);
    cudaFree(e->d_weights);
    delete e;
    return SIB_OK;
}

extern "C" int sib_engine_info(sib_engine* e, sib_model_info* out, sib_config* eff) {
    if (!e) return SIB_EINVAL;
    if (out) {
        const ModelDesc& m = e->model;
        out->in_h = m.in_h; out->in_w = m.in_w; out->in_c = m.in_c; out->n_layers = m.n_layers;
        out->n_classes = m.n_classes; out->n_weights = m.n_weights;
        out->flops_per_image = m.flops_per_image;
        out->bytes_per_image = (int64_t)m.img_elems * 4 + (int64_t)m.n_classes * 4;
    }
    if (eff) *eff = e->cfg;
    return SIB_OK;
}

// Finds (or starts) the slot to append n images to. Returns slot index or a negative code. Caller holds mu.
static int slot_for(sib_engine* e, int n) {
    if (n > e->cfg.max_batch) { set_error("tuple holds more images than max_batch"); return SIB_ECAP; }
    if (e->filling >= 0 && e->slots[e->filling].images + n > e->cfg.max_batch) e->seal_locked(e->filling, 0);
    if (e->filling < 0) {
        for (size_t i = 0; i < e->slots.size(); i++) {
            if (e->slots[i].state == Slot::FREE) {
                Slot& s = e->slots[i];
                s.state = Slot::FILLING;
                s.staged.clear();
                s.direct_bytes = 0;
                s.text_used = 0;
                s.n_text = 0;
                s.side.clear();
                s.writers = 0;
                s.images = 0;
                s.entries.clear();
                s.next_entry = 0;
                s.status = SIB_OK;
                e->filling = (int)i;
                break;
            }
        }
        if (e->filling < 0) { set_error("all batch slots busy: poll first"); return SIB_EAGAIN; }
    }
    return e->filling;
}

// Reserves room for n images in slot si and fixes the tuple's place in the result order. The caller copies its
// images to the returned offset WITHOUT the lock (so several executor threads fill one batch in parallel), then
// calls written(). Caller holds mu.
static int reserve(sib_engine* e, int si, uint64_t tag, int n, bool staged = true) {
    Slot& s = e->slots[si];
    const int off = s.images;
    if (staged) {
        if (!s.staged.empty() && s.staged.back().first + s.staged.back().second == off) s.staged.back().second += n;
        else s.staged.emplace_back(off, n);
    }
    if (s.images == 0) {
        s.first_ts = Clock::now();
        e->cv.notify_all();  // arm the deadline thread
    }
    s.entries.push_back(Entry{tag, n, off});
    s.images += n;
    s.writers.v.fetch_add(1);
    e->stats.tuples++;
    e->stats.images += (uint64_t)n;
    if (s.images >= e->cfg.max_batch) e->seal_locked(si, 0);  // full: sealed now, launched when the copies are in
    else if (!e->cfg.strict_batching && e->fifo.empty()) e->seal_locked(si, 3);  // idle GPU: do not sit out the deadline
    return off;
}

// A submitter is done copying. No lock on the common path: the count drops, and only the writer that takes it to
// zero looks further. Sealing (under the lock) stores SEALED and then reads the count; this side decrements and then
// reads the state; both sequentially consistent, so at least one of the two sees the other and launches, and the
// state check under the lock makes sure only one does.
static void written(sib_engine* e, int si) {
    Slot& s = e->slots[si];
    if (s.writers.v.fetch_sub(1) != 1) return;
    if (s.state.v.load() != Slot::SEALED && !e->closing.v.load()) return;
    std::unique_lock<std::mutex> lk(e->mu);
    if (s.state == Slot::SEALED && s.writers == 0) e->launch_locked(si);  // notifies the pollers itself
    else if (e->closing) e->cv.notify_all();                             // sib_engine_close waits for the writers to drain
}

extern "C" int sib_submit_f32(sib_engine* e, uint64_t tag, const float* nhwc, int n) {
    if (!e || !nhwc || n <= 0) return SIB_EINVAL;
    int si, off;
    {
        std::unique_lock<std::mutex> lk(e->mu);
        if (e->closing) return SIB_ECLOSED;
        si = slot_for(e, n);
        if (si < 0) return si;
        off = reserve(e, si, tag, n);
    }
    const size_t img = e->model.img_elems;
    std::memcpy(e->slots[si].h_in + (size_t)off * img, nhwc, (size_t)n * img * sizeof(float));
    written(e, si);
    return n;
}

extern "C" int sib_submit_pinned(sib_engine* e, uint64_t tag, const float* pinned_nhwc, int n) {
    if (!e || !pinned_nhwc || n <= 0) return SIB_EINVAL;
    std::unique_lock<std::mutex> lk(e->mu);
    if (e->closing) return SIB_ECLOSED;
    const int si = slot_for(e, n);
    if (si < 0) return si;
    Slot& s = e->slots[si];
    const int off = s.images;
    // No host staging: the DMA reads the caller's page-locked memory and lands in the slot's device buffer. It is
    // enqueued before the reservation can seal the slot, so the batch's kernel is ordered after it on the stream.
    cudaSetDevice(e->device);
    const size_t bytes = (size_t)n * e->model.img_elems * sizeof(float);
    cudaError_t ce = cudaMemcpyAsync(s.dev.d_in + (size_t)off * e->model.img_elems, pinned_nhwc, bytes,
                                     cudaMemcpyHostToDevice, s.stream);
    if (ce != cudaSuccess) { set_error(std::string("sib_submit_pinned H2D: ") + cudaGetErrorString(ce)); return SIB_ECUDA; }
    s.direct_bytes += bytes;
    reserve(e, si, tag, n, false);  // counts this call as a writer, so a slot that just became full is only sealed
    if (s.writers.v.fetch_sub(1) == 1 && s.state == Slot::SEALED) e->launch_locked(si);
    return n;
}

extern "C" int sib_submit_json(sib_engine* e, uint64_t tag, const char* utf8, size_t len) {
    if (!e || !utf8) return SIB_EINVAL;
    const ModelDesc& m = e->model;
    const size_t img = m.img_elems;
    bool raw = e->cfg.json_on_gpu == 1;
    if (e->cfg.json_on_gpu == 2) {
        // auto: an idle engine decodes on the submitting thread and launches at once (a single tuple is back sooner
        // that way: 108 us against 123 us round trip); a busy one ships the text, which is what scales
        std::unique_lock<std::mutex> lk(e->mu);
        raw = !e->fifo.empty() || (e->filling >= 0 && e->slots[e->filling].images > 0);
    }
    // (a tuple longer than a whole slot's text area -- small max_batch, long number text -- goes to the host decoder:
    // after the area of a non-empty slot is closed below, a fresh, empty one always has room for it)
    if (raw && len >= 16 && len <= json_gpu_max_text() && len + 32 <= e->slots[0].text_cap) {
        // Raw path: the text itself goes into the slot's pinned text area and is decoded by a kernel after the H2D.
        // The engine learns only at completion whether the tuple was well-formed (sib_result.status).
        Slot* s;
        int si;
        size_t toff;
        {
            std::unique_lock<std::mutex> lk(e->mu);
            if (e->closing) return SIB_ECLOSED;
            si = slot_for(e, 1);
            if (si < 0) return si;
            s = &e->slots[si];
            toff = (s->text_used + 15) & ~(size_t)15;
            if (toff + len + 32 > s->text_cap) {  // text area full before the image count: close this batch
                e->seal_locked(si, 0);
                si = slot_for(e, 1);
                if (si < 0) return si;
                s = &e->slots[si];
                toff = 0;
            }
            const int off = reserve(e, si, tag, 1, false);
            int32_t* d = s->h_tdesc + 4 * s->n_text;
            d[0] = (int32_t)toff; d[1] = (int32_t)len; d[2] = off; d[3] = 0;
            for (auto it = s->entries.rbegin(); it != s->entries.rend(); ++it)
                if (it->offset == off) { it->text = s->n_text; break; }
            s->n_text++;
            s->text_used = toff + len;
        }
        std::memcpy(s->h_text + toff, utf8, len);
        written(e, si);
        return 1;
    }
    // The usual tuple carries one instance (InferenceBolt.java:86 hard-wires [1][10]): reserve one image and decode
    // the text straight into the pinned slot, outside the lock. A tuple with more instances does not fit that
    // reservation; it is decoded into a thread-local buffer and copied in like sib_submit_f32 does.
    int si, off;
    {
        std::unique_lock<std::mutex> lk(e->mu);
        if (e->closing) return SIB_ECLOSED;
        si = slot_for(e, 1);
        if (si < 0) return si;
        off = reserve(e, si, tag, 1);
    }
    Slot& s = e->slots[si];
    int64_t dims[4] = {0, 0, 0, 0};
    long nf = json_decode(utf8, len, s.h_in + (size_t)off * img, (long)img, dims);
    int rc = 1;
    if (nf == SIB_EJSON) { set_error("malformed instances JSON"); rc = SIB_EJSON; }
    else if (nf == SIB_ESHAPE) { set_error("instances is not a rectangular rank-4 array"); rc = SIB_ESHAPE; }
    else if (nf == SIB_ECAP || nf > 0) {
        if (dims[1] != m.in_h || dims[2] != m.in_w || dims[3] != m.in_c) {
            set_error("instances HxWxC does not match the model input");
            rc = SIB_ESHAPE;
        } else if (dims[0] > e->cfg.max_batch) {
            set_error("tuple holds more images than max_batch");
            rc = SIB_ECAP;
        } else if (dims[0] != 1) {
            rc = 0;  // N > 1: second path below
        }
    } else {
        rc = (int)nf;
    }
    if (rc == 1) {
        written(e, si);
        return 1;
    }
    // The reservation is not used after all: turn it into an empty entry (its image row is computed and dropped).
    {
        std::unique_lock<std::mutex> lk(e->mu);
        for (auto& en : s.entries)
            if (en.offset == off && en.tag == tag && en.n == 1) { en.n = 0; break; }
        std::memset(s.h_in + (size_t)off * img, 0, img * sizeof(float));
        e->stats.tuples--;
        e->stats.images--;
    }
    written(e, si);
    if (rc < 0) return rc;
    static thread_local std::vector<float> tmp;
    const size_t need = (size_t)dims[0] * img;
    if (tmp.size() < need) tmp.resize(need);
    nf = json_decode(utf8, len, tmp.data(), (long)need, dims);
    if (nf < 0) return (int)nf;
    int r = sib_submit_f32(e, tag, tmp.data(), (int)dims[0]);
    return r;
}

extern "C" int sib_flush(sib_engine* e) {
    if (!e) return SIB_EINVAL;
    std::unique_lock<std::mutex> lk(e->mu);
    if (e->filling >= 0 && e->slots[e->filling].images > 0) return e->seal_locked(e->filling, 2);
    return SIB_OK;
}

// sib_poll and sib_poll_json_batch. `budget` (may be null): bytes of message text the caller can still take; a tuple whose
// upper bound does not fit stays queued for the next poll (never the first tuple of a poll, so progress is guaranteed
// for any arena that holds one message).
static size_t message_bound(int n, int nc) { return 32 + (size_t)n * ((size_t)nc * 18 + 4); }  // json_encode's own bound

static int poll_core(sib_engine* e, sib_result* out, int max, int timeout_us, size_t* budget) {
    std::unique_lock<std::mutex> lk(e->mu);
    // results handed out by the previous poll are dead now: recycle their slots
    for (auto& s : e->slots)
        if (s.state == Slot::HANDED) s.state = Slot::FREE;
    const auto t_end = Clock::now() + std::chrono::microseconds(timeout_us > 0 ? timeout_us : 0);
    int got = 0;
    const int nc = (int)e->model.n_classes;
    for (;;) {
        while (got < max && !e->fifo.empty()) {
            Slot& s = e->slots[e->fifo.front()];
            if (s.state == Slot::SEALED || s.state == Slot::REDO) break;  // a submitter is still copying into it | another poller is re-reading it
            if (s.state == Slot::INFLIGHT) {
                cudaError_t q = cudaEventQuery(s.done);
                if (q == cudaErrorNotReady) break;
                if (q != cudaSuccess) {
                    s.status = SIB_ECUDA;
                    set_error(std::string("batch failed: ") + cudaGetErrorString(q));
                }
                s.state = Slot::READY;
                if (e->cfg.precision != SIB_PREC_FP32 && e->cfg.guard) e->stats.guard_recomputed += (uint64_t)s.r_guard_n[0];
                if (s.n_text > 0) {  // host decoder for what the decode kernel declined: outside the engine lock
                    s.state = Slot::REDO;
                    lk.unlock();
                    e->redo_flagged(s);
                    lk.lock();
                    s.state = Slot::READY;
                    e->cv.notify_all();
                }
            }
            bool arena_full = false;
            while (got < max && s.next_entry < s.entries.size()) {
                const Entry& en = s.entries[s.next_entry++];
                if (en.n == 0 && en.status == SIB_OK) continue;  // reservation that was given back (see sib_submit_json)
                if (budget) {
                    const size_t need = message_bound(en.n, nc);
                    if (need > *budget && got > 0) { s.next_entry--; arena_full = true; break; }
                    *budget -= need < *budget ? need : *budget;
                }
                sib_result& r = out[got++];
                r.tag = en.tag; r.n = en.n; r.n_classes = nc; r.status = s.status != SIB_OK ? s.status : en.status; r.reserved = 0;
                if (en.n == 0) { r.probs = nullptr; r.argmax = nullptr; }
                else if (en.side >= 0) { r.probs = s.side[(size_t)en.side].probs.data(); r.argmax = s.side[(size_t)en.side].argmax.data(); }
                else { r.probs = s.h_probs + (size_t)en.offset * nc; r.argmax = s.r_argmax + en.offset; }
            }
            if (s.next_entry == s.entries.size() && !arena_full) {
                s.state = Slot::HANDED;
                e->fifo.pop_front();
            } else {
                break;
            }
        }
        // the GPU just ran dry: whatever queued up behind the batch that was handed out leaves now, not at its deadline
        if (!e->cfg.strict_batching && e->fifo.empty() && e->filling >= 0 && e->slots[e->filling].images > 0)
            e->seal_locked(e->filling, 3);
        if (got > 0 || timeout_us == 0) break;
        const bool pending = !e->fifo.empty() || (e->filling >= 0 && e->slots[e->filling].images > 0);
        if (!pending) break;
        if (timeout_us > 0 && Clock::now() >= t_end) break;
        if (!e->fifo.empty() && (e->slots[e->fifo.front()].state == Slot::SEALED || e->slots[e->fifo.front()].state == Slot::REDO)) {
            if (timeout_us < 0) e->cv.wait(lk);
            else e->cv.wait_until(lk, t_end);
        } else if (!e->fifo.empty()) {
            // wait for the oldest batch without holding the lock
            cudaEvent_t ev = e->slots[e->fifo.front()].done;
            lk.unlock();
            if (timeout_us < 0) {
                cudaEventSynchronize(ev);
            } else {
                while (cudaEventQuery(ev) == cudaErrorNotReady && Clock::now() < t_end) std::this_thread::yield();
            }
            lk.lock();
        } else if (timeout_us < 0) {
            e->cv.wait(lk);  // the deadline thread will launch the filling slot
        } else {
            e->cv.wait_until(lk, t_end);
        }
    }
    return got;
}

extern "C" int sib_poll(sib_engine* e, sib_result* out, int max, int timeout_us) {
    if (!e || !out || max <= 0) return SIB_EINVAL;
    return poll_core(e, out, max, timeout_us, nullptr);
}

// Batched egress: every message of a poll formatted in one native pass, outside the engine lock.
extern "C" int sib_poll_json_batch(sib_engine* e, char* arena, size_t cap, uint64_t* tags, uint32_t* offsets, int32_t* status,
                                   int max, int timeout_us) {
    if (!e || !arena || !tags || !offsets || !status || max <= 0) return SIB_EINVAL;
    if (cap < message_bound(1, (int)e->model.n_classes)) { set_error("sib_poll_json_batch: arena smaller than one message"); return SIB_ECAP; }
    static thread_local std::vector<sib_result> res;
    if ((int)res.size() < max) res.resize((size_t)max);
    size_t budget = cap;
    const int n = poll_core(e, res.data(), max, timeout_us, &budget);
    size_t at = 0;
    for (int i = 0; i < n; i++) {
        tags[i] = res[i].tag;
        status[i] = res[i].status;
        offsets[i] = (uint32_t)at;
        if (res[i].status == SIB_OK && res[i].probs) {
            const long len = json_encode(res[i].probs, res[i].n, res[i].n_classes, arena + at, cap - at);
            if (len > 0) at += (size_t)len;          // written without a terminating NUL: the offsets delimit it
            else status[i] = (int32_t)(len < 0 ? len : SIB_ECAP);  // only the first tuple of a poll can get here, when its message alone
                                                                   // (hundreds of images in one tuple) is larger than the arena: it is
                                                                   // reported, not queued for ever
        }
    }
    if (n >= 0) offsets[n] = (uint32_t)at;
    return n;
}

extern "C" long sib_format_json(const sib_result* r, char* buf, size_t cap) {
    if (!r || !buf || !r->probs) return SIB_EINVAL;
    return json_encode(r->probs, r->n, r->n_classes, buf, cap);
}

extern "C" int sib_stats_get(sib_engine* e, sib_stats* out) {
    if (!e || !out) return SIB_EINVAL;
    std::unique_lock<std::mutex> lk(e->mu);
    *out = e->stats;
    out->kernel_launches = e->launches.load(std::memory_order_relaxed);
    return SIB_OK;
}

extern "C" int sib_forward_sync(sib_engine* e, const float* x, int n, float* probs, float* logits, int32_t* argmax) {
    if (!e || !x || !probs || n < 0) return SIB_EINVAL;
    std::unique_lock<std::mutex> lk(e->sync_mu);
    cudaSetDevice(e->device);
    Slot& s = e->sync_slot;
    const size_t img = e->model.img_elems, nc = e->model.n_classes;
    std::vector<float> lg_host;
    for (int done = 0; done < n;) {
        const int b = std::min(n - done, e->cfg.max_batch);
        std::memcpy(s.h_in, x + (size_t)done * img, (size_t)b * img * sizeof(float));
        CU_TRY(cudaMemcpyAsync(s.dev.d_in, s.h_in, (size_t)b * img * sizeof(float), cudaMemcpyHostToDevice, s.stream));
        const bool guarded = e->cfg.precision != SIB_PREC_FP32 && e->cfg.guard;
        GuardScratch g = s.dev.guard;
        int32_t* d_listed = reinterpret_cast<int32_t*>(s.dev.d_probs + (size_t)e->cfg.max_batch * (nc + 1));  // spare words behind the result block
        g.d_out_count = d_listed;
        int r = e->run_forward(s.dev.d_in, b, s.dev.d_probs, logits ? s.dev.d_logits : nullptr, s.dev.d_argmax, &g, s.stream);
        if (r) return r;
        int32_t listed = 0;
        if (guarded) CU_TRY(cudaMemcpyAsync(&listed, d_listed, sizeof listed, cudaMemcpyDeviceToHost, s.stream));
        CU_TRY(cudaMemcpyAsync(s.h_probs, s.dev.d_probs, (size_t)b * nc * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CU_TRY(cudaMemcpyAsync(s.h_argmax, s.dev.d_argmax, (size_t)b * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        if (logits) CU_TRY(cudaMemcpyAsync(logits + (size_t)done * nc, s.dev.d_logits, (size_t)b * nc * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CU_TRY(cudaStreamSynchronize(s.stream));
        if (guarded) { std::lock_guard<std::mutex> sl(e->mu); e->stats.guard_recomputed += (uint64_t)listed; }
        std::memcpy(probs + (size_t)done * nc, s.h_probs, (size_t)b * nc * sizeof(float));
        if (argmax) std::memcpy(argmax + done, s.h_argmax, (size_t)b * sizeof(int32_t));
        done += b;
    }
    return SIB_OK;
}

extern "C" int sib_forward_device(sib_engine* e, const float* d_nhwc, int n, float* d_probs, float* d_logits,
                                  int32_t* d_argmax, void* stream) {
    if (!e || !d_nhwc || !d_probs || n < 0) return SIB_EINVAL;
    if (n > e->cfg.max_batch) { set_error("n exceeds max_batch"); return SIB_ECAP; }
    if (e->cfg.precision != SIB_PREC_FP32 && e->cfg.guard && !d_argmax) {
        set_error("the guarded tensor-core path needs d_argmax");
        return SIB_EINVAL;
    }
    cudaSetDevice(e->device);
    cudaStream_t st = (cudaStream_t)stream;  // NULL is the CUDA default stream, as everywhere in CUDA
    return e->run_forward(d_nhwc, n, d_probs, d_logits, d_argmax, &e->dev_scratch.guard, st);
}

// Test hook: runs the tensor-core kernel on up to 16 host images and returns the intermediate activations of the
// first 16-image tile (pooled conv1 map of both 8-image slots, pooled conv2 features, fc1 output), as fp32.
// TF exposes the same thing by fetching inner tensors (Session.runner().fetch(name), InferenceBolt.java:81-85).
extern "C" long sib_debug_stage_dump(sib_engine* e, const float* x, int n, float* out, long cap) {
    if (!e || !x || !out || n <= 0 || n > 16) return SIB_EINVAL;
    if (e->cfg.precision == SIB_PREC_FP32) { set_error("stage dump needs a tensor-core precision"); return SIB_EINVAL; }
    if (e->tc.generic) { set_error("stage dump exists for the fused LeNet-5 kernel only"); return SIB_EINVAL; }
    const long need = (long)tc_debug_floats(e->model.preset);
    if (cap < need) return SIB_ECAP;
    std::unique_lock<std::mutex> lk(e->sync_mu);
    cudaSetDevice(e->device);
    Slot& s = e->sync_slot;
    float* d_dbg = nullptr;
    CU_TRY(cudaMalloc((void**)&d_dbg, need * sizeof(float)));
    CU_TRY(cudaMemset(d_dbg, 0, need * sizeof(float)));
    CU_TRY(cudaMemcpy(s.dev.d_in, x, (size_t)n * e->model.img_elems * sizeof(float), cudaMemcpyHostToDevice));
    int r = launch_forward_tc(e->tc, s.dev.d_in, n, s.dev.d_probs, nullptr, s.dev.d_argmax, nullptr, 0.f, d_dbg, s.stream);
    cudaError_t ce = cudaStreamSynchronize(s.stream);
    if (!r && ce == cudaSuccess) ce = cudaMemcpy(out, d_dbg, need * sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(d_dbg);
    if (r || ce != cudaSuccess) { set_error(std::string("stage dump: ") + cudaGetErrorString(r ? (cudaError_t)r : ce)); return SIB_ECUDA; }
    return need;
}

namespace sib {
bool engine_view(sib_engine* e, EngineView* out) {
    if (!e || !out) return false;
    out->device = e->device; out->precision = e->cfg.precision; out->guard = e->cfg.guard; out->tau = e->cfg.guard_tau;
    out->preset = e->model.preset; out->tc = &e->tc; out->dm = &e->dm;
    return true;
}
}  // namespace sib
