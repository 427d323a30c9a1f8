// sib_shared.cpp -- one engine per (device, model, configuration) for all the bolts of a worker process.
//
// Storm runs several executors of a bolt in one worker JVM (MainTopology.java:27 INFERENCE_BOLT_PARAL = 4 over
// setNumWorkers, :65-66); the reference gives each its own SavedModelBundle (InferenceBolt.java:57-58). Here
// co-located bolts may share ONE micro-batcher, stream set and weight upload: sib_client_open hands out a view of a
// process-wide, reference-counted engine, so their tuples fill the same batches. Results are routed back to the
// client that submitted them, in that client's submission order (the engine keeps global submission order, which
// is each client's order interleaved), and are copied into client-owned storage so that one bolt's polling pace
// never holds the others' slots.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "sib_internal.hpp"

namespace {

using Clock = std::chrono::steady_clock;

struct Owned {  // one tuple's result, owned by the client until its next poll
    uint64_t tag = 0;
    int32_t n = 0, n_classes = 0, status = 0;
    std::vector<float> probs;
    std::vector<int32_t> argmax;
};

struct Shared {
    sib_engine* engine = nullptr;
    std::string key;
    int refs = 0;
    std::mutex reg_mu;   // seq -> owner table
    std::mutex poll_mu;  // one thread at a time drains the engine (its results die at the next sib_poll)
    uint64_t next_seq = 1;
    std::unordered_map<uint64_t, std::pair<sib_client*, uint64_t>> owner;  // seq -> (client, caller's tag)
    std::vector<sib_result> scratch = std::vector<sib_result>(512);
};

std::mutex g_mu;
std::map<std::string, Shared*> g_shared;

}  // namespace

struct sib_client {
    Shared* sh = nullptr;
    std::mutex mu;            // ready queue (filled by whichever client drains the engine)
    std::deque<Owned> ready;
    std::vector<Owned> handed;  // what the last poll returned pointers into
    std::atomic<long> owed{0};  // tuples submitted and not yet routed back
};

using namespace sib;

namespace {

// Moves every finished tuple out of the engine and into its owner's queue. Caller holds sh->poll_mu.
int drain(Shared* sh, int timeout_us) {
    int total = 0;
    for (;;) {
        const int n = sib_poll(sh->engine, sh->scratch.data(), (int)sh->scratch.size(), total ? 0 : timeout_us);
        if (n < 0) return n;
        if (n == 0) break;
        for (int i = 0; i < n; i++) {
            const sib_result& r = sh->scratch[(size_t)i];
            sib_client* c = nullptr;
            uint64_t tag = 0;
            {
                std::lock_guard<std::mutex> lk(sh->reg_mu);
                auto it = sh->owner.find(r.tag);
                if (it == sh->owner.end()) continue;  // cannot happen: every submit registers first
                c = it->second.first;
                tag = it->second.second;
                sh->owner.erase(it);
            }
            Owned o;
            o.tag = tag; o.n = r.n; o.n_classes = r.n_classes; o.status = r.status;
            if (r.n > 0 && r.probs) {
                o.probs.assign(r.probs, r.probs + (size_t)r.n * r.n_classes);
                o.argmax.assign(r.argmax, r.argmax + r.n);
            }
            std::lock_guard<std::mutex> lk(c->mu);
            c->ready.push_back(std::move(o));
            c->owed.fetch_sub(1);
        }
        total += n;
    }
    return total;
}

template <typename Submit>
int submit(sib_client* c, uint64_t tag, Submit call) {
    if (!c) return SIB_EINVAL;
    Shared* sh = c->sh;
    uint64_t seq;
    {
        std::lock_guard<std::mutex> lk(sh->reg_mu);  // registered before the engine can possibly finish it
        seq = sh->next_seq++;
        sh->owner.emplace(seq, std::make_pair(c, tag));
    }
    c->owed.fetch_add(1);
    const int rc = call(sh->engine, seq);
    if (rc <= 0) {  // refused (SIB_EAGAIN, SIB_EJSON, ...): nothing will come back for it
        std::lock_guard<std::mutex> lk(sh->reg_mu);
        sh->owner.erase(seq);
        c->owed.fetch_sub(1);
    }
    return rc;
}

}  // namespace

extern "C" int sib_client_open(const void* model, size_t n, const sib_config* ucfg, sib_client** out) {
    if (!out || !model) return SIB_EINVAL;
    *out = nullptr;
    sib_config cfg{};
    if (ucfg) cfg = *ucfg;
    else { cfg.device = -1; cfg.guard = 1; }
    if (cfg.device < 0) {  // resolve the device now: executors with different task indices land on different GPUs
        const int ndev = sib_device_count();
        if (ndev <= 0) { set_error("no CUDA device"); return SIB_ECUDA; }
        cfg.device = (cfg.task_index < 0 ? 0 : cfg.task_index) % ndev;
    }
    cfg.task_index = 0;
    std::string key(reinterpret_cast<const char*>(&cfg), sizeof(cfg));
    key.append(reinterpret_cast<const char*>(model), n);  // same bytes, same engine (models are ~250 KB)
    std::lock_guard<std::mutex> lk(g_mu);
    Shared*& sh = g_shared[key];
    if (!sh) {
        sib_engine* e = nullptr;
        const int rc = sib_engine_open(model, n, &cfg, &e);
        if (rc != SIB_OK) { g_shared.erase(key); return rc; }
        sh = new Shared;
        sh->engine = e;
        sh->key = key;
    }
    sh->refs++;
    sib_client* c = new sib_client;
    c->sh = sh;
    *out = c;
    return SIB_OK;
}

extern "C" int sib_client_close(sib_client* c) {
    if (!c) return SIB_EINVAL;
    Shared* sh = c->sh;
    {   // results still owed to this client are dropped; forget their routes (no drain may be holding `c` meanwhile)
        std::lock_guard<std::mutex> pk(sh->poll_mu);
        std::lock_guard<std::mutex> lk(sh->reg_mu);
        for (auto it = sh->owner.begin(); it != sh->owner.end();)
            it = it->second.first == c ? sh->owner.erase(it) : std::next(it);
    }
    bool last;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        last = --sh->refs == 0;
        if (last) g_shared.erase(sh->key);
    }
    if (last) {
        sib_engine_close(sh->engine);
        delete sh;
    }
    delete c;
    return SIB_OK;
}

extern "C" int sib_client_submit_json(sib_client* c, uint64_t tag, const char* utf8, size_t len) {
    return submit(c, tag, [&](sib_engine* e, uint64_t seq) { return sib_submit_json(e, seq, utf8, len); });
}

extern "C" int sib_client_submit_f32(sib_client* c, uint64_t tag, const float* nhwc, int n) {
    return submit(c, tag, [&](sib_engine* e, uint64_t seq) { return sib_submit_f32(e, seq, nhwc, n); });
}

extern "C" int sib_client_flush(sib_client* c) { return c ? sib_flush(c->sh->engine) : SIB_EINVAL; }

extern "C" int sib_client_stats(sib_client* c, sib_stats* out) { return c ? sib_stats_get(c->sh->engine, out) : SIB_EINVAL; }

extern "C" int sib_client_poll(sib_client* c, sib_result* out, int max, int timeout_us) {
    if (!c || !out || max <= 0) return SIB_EINVAL;
    Shared* sh = c->sh;
    c->handed.clear();  // the previous poll's results die here
    const auto t_end = Clock::now() + std::chrono::microseconds(timeout_us > 0 ? timeout_us : 0);
    for (;;) {
        {
            std::lock_guard<std::mutex> lk(c->mu);
            while (!c->ready.empty() && (int)c->handed.size() < max) {
                c->handed.push_back(std::move(c->ready.front()));
                c->ready.pop_front();
            }
        }
        if (!c->handed.empty()) break;
        if (c->owed.load() == 0) {  // nothing of ours is pending: like sib_poll, do not wait for nothing
            std::lock_guard<std::mutex> lk(c->mu);
            if (c->ready.empty()) break;
            continue;
        }
        // nothing of ours yet: take a turn at draining the engine (someone has to), in short slices so that the
        // other clients get their turn and see what this one routed to them
        int slice = 0;
        if (timeout_us < 0) slice = 200;
        else if (timeout_us > 0) {
            const auto left = std::chrono::duration_cast<std::chrono::microseconds>(t_end - Clock::now()).count();
            if (left <= 0) break;
            slice = (int)(left < 200 ? left : 200);
        }
        int rc;
        {
            std::lock_guard<std::mutex> lk(sh->poll_mu);
            rc = drain(sh, slice);
        }
        if (rc < 0) return rc;
        if (rc == 0 && timeout_us == 0) {
            std::lock_guard<std::mutex> lk(c->mu);
            if (c->ready.empty()) break;
        }
    }
    for (size_t i = 0; i < c->handed.size(); i++) {
        const Owned& o = c->handed[i];
        sib_result& r = out[i];
        r.tag = o.tag; r.n = o.n; r.n_classes = o.n_classes; r.status = o.status; r.reserved = 0;
        r.probs = o.n > 0 ? o.probs.data() : nullptr;
        r.argmax = o.n > 0 ? o.argmax.data() : nullptr;
    }
    return (int)c->handed.size();
}
