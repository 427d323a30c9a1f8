// sib_loadgen.cpp -- open-loop arrival generator + result sink (include/sib_loadgen.h).
#include "../../include/sib_loadgen.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <thread>
#include <vector>

#include "sib_internal.hpp"

namespace {
using Clock = std::chrono::steady_clock;
inline int64_t now_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now().time_since_epoch()).count();
}
inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
}  // namespace

extern "C" int sib_loadgen_run(sib_engine* e, const sib_loadgen_cfg* cfg, const void* const* payloads,
                               const size_t* lens, int n_payloads, int64_t* lat_out, sib_loadgen_report* out) {
    if (!e || !cfg || !payloads || !lens || n_payloads <= 0 || !out || cfg->n_tuples <= 0 || cfg->rate_per_s <= 0) return SIB_EINVAL;
    sib_model_info info;
    sib_engine_info(e, &info, nullptr);
    const size_t img_bytes = (size_t)info.in_h * info.in_w * info.in_c * sizeof(float);
    const int64_t N = cfg->n_tuples;
    const int P = std::max(1, cfg->producers);
    // arrival schedule (ns from start), shared by all producers: tuple i belongs to producer i % P
    std::vector<int64_t> arrive((size_t)N);
    {
        const double gap = 1e9 / cfg->rate_per_s;
        double t = 0;
        for (int64_t i = 0; i < N; i++) {
            arrive[(size_t)i] = (int64_t)t;
            if (cfg->poisson) {
                double u = ((mix64(cfg->seed + (uint64_t)i) >> 11) + 1.0) * (1.0 / 9007199254740993.0);
                t += -std::log(u) * gap;
            } else {
                t += gap;
            }
        }
    }
    std::vector<int64_t> lat((size_t)N, -1);
    std::atomic<int64_t> submitted{0}, rejected{0}, spins{0}, late{0}, images{0};
    std::atomic<bool> producers_done{false};
    const int64_t t0 = now_ns() + 2000000;  // start 2 ms from now so every thread is in place

    auto producer = [&](int p) {
        for (int64_t i = p; i < N; i += P) {
            const int64_t due = t0 + arrive[(size_t)i];
            while (now_ns() < due) {
                if (due - now_ns() > 200000) std::this_thread::sleep_for(std::chrono::microseconds(50));
            }
            if (now_ns() - due > 100000) late++;
            const int k = (int)(i % n_payloads);
            for (;;) {
                int rc = cfg->mode == SIB_LG_JSON
                             ? sib_submit_json(e, (uint64_t)i, (const char*)payloads[k], lens[k])
                             : sib_submit_f32(e, (uint64_t)i, (const float*)payloads[k], (int)(lens[k] / img_bytes));
                if (rc == SIB_EAGAIN) { spins++; std::this_thread::yield(); continue; }
                if (rc <= 0) rejected++;
                else { submitted++; images += rc; }
                break;
            }
        }
    };
    int64_t completed = 0;
    auto consumer = [&]() {
        std::vector<sib_result> res(1024);
        std::vector<char> buf(1 << 16);
        for (;;) {
            int n = sib_poll(e, res.data(), (int)res.size(), 200);
            for (int i = 0; i < n; i++) {
                if (cfg->mode == SIB_LG_JSON) {
                    size_t need = 64 + (size_t)res[i].n * ((size_t)res[i].n_classes * 18 + 4);
                    if (buf.size() < need) buf.resize(need);
                    sib_format_json(&res[i], buf.data(), buf.size());
                }
                const int64_t t = now_ns();
                const uint64_t tag = res[i].tag;
                if (tag < (uint64_t)N) lat[(size_t)tag] = t - (t0 + arrive[(size_t)tag]);
                completed++;
            }
            if (n == 0 && producers_done.load() && completed >= submitted.load()) break;
        }
    };
    std::vector<std::thread> th;
    std::thread cons(consumer);
    for (int p = 0; p < P; p++) th.emplace_back(producer, p);
    for (auto& t : th) t.join();
    sib_flush(e);
    producers_done = true;
    cons.join();
    const int64_t t1 = now_ns();

    std::vector<int64_t> v;
    v.reserve((size_t)N);
    double sum = 0;
    for (int64_t x : lat)
        if (x >= 0) { v.push_back(x); sum += (double)x; }
    std::sort(v.begin(), v.end());
    auto pct = [&](double q) -> int64_t { return v.empty() ? -1 : v[std::min(v.size() - 1, (size_t)(q * (double)v.size()))]; };
    *out = sib_loadgen_report{};
    out->completed = completed; out->rejected = rejected.load(); out->images = images.load();
    out->wall_s = (double)(t1 - t0) * 1e-9;
    out->achieved_tuples_per_s = out->wall_s > 0 ? (double)completed / out->wall_s : 0;
    out->p50_ns = pct(0.50); out->p90_ns = pct(0.90); out->p99_ns = pct(0.99); out->p999_ns = pct(0.999);
    out->max_ns = v.empty() ? -1 : v.back();
    out->mean_ns = v.empty() ? -1 : (int64_t)(sum / (double)v.size());
    out->eagain_spins = spins.load(); out->late_submits = late.load();
    if (lat_out) std::copy(lat.begin(), lat.end(), lat_out);
    return SIB_OK;
}
