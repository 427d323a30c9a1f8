// sib_bolt.cpp -- dke::model::InferenceBolt over the C ABI, plus the bolt-contract harness (include/sib_bolt.h).
//
// Contract kept from the reference (InferenceBolt.java:69-100): one emit and one ack per input tuple, in
// per-executor arrival order; malformed JSON => emit a null message and still ack; a payload whose shape the model
// cannot take => IllegalArgumentException out of execute(). What changes: tuples are batched, so emit/ack happen on
// a completion thread under the collector's monitor (the reference's own KafkaBolt does the same, KafkaBolt.java:130-142).
#include "sib_bolt.hpp"

#include <chrono>
#include <cstring>

#include "../../include/sib_bolt.h"

namespace dke {
namespace model {

static int conf_int(const Conf& c, const char* k, int dflt) {
    auto it = c.find(k);
    if (it == c.end() || it->second.empty()) return dflt;
    return std::atoi(it->second.c_str());
}

void InferenceBolt::prepare(const Conf& conf, const TopologyContext& ctx, OutputCollector* collector) {
    collector_ = collector;
    sib_config cfg{};
    cfg.task_index = ctx.getThisTaskIndex();
    cfg.device = conf_int(conf, "sib.device", -1);  // default: task index mod GPU count
    cfg.max_batch = conf_int(conf, "sib.max.batch", 0);
    cfg.deadline_us = conf_int(conf, "sib.deadline.us", 0);
    cfg.n_slots = conf_int(conf, "sib.slots", 0);
    cfg.guard = conf_int(conf, "sib.guard", 1);
    cfg.json_on_gpu = conf_int(conf, "sib.json.gpu", 0);
    cfg.strict_batching = conf_int(conf, "sib.strict.batching", 0);
    auto tau = conf.find("sib.guard.tau");
    cfg.guard_tau = tau == conf.end() ? 0.0f : (float)std::atof(tau->second.c_str());
    auto pr = conf.find("sib.precision");
    cfg.precision = pr == conf.end() ? SIB_PREC_FP16 : pr->second == "fp32" ? SIB_PREC_FP32 : pr->second == "bf16" ? SIB_PREC_BF16 : SIB_PREC_FP16;
    // sib.shared = 1: the bolts of this worker process that load the same model with the same settings share one engine
    // (one batcher, one set of streams, one weight upload); each still sees only its own tuples, in its own order
    shared_ = conf_int(conf, "sib.shared", 0) != 0;
    int rc = shared_ ? sib_client_open(blob_.data(), blob_.size(), &cfg, &client_)
                     : sib_engine_open(blob_.data(), blob_.size(), &cfg, &engine_);
    // the reference lets SavedModelBundle.load throw out of prepare(); so do we
    if (rc != SIB_OK) throw std::runtime_error(std::string("InferenceBolt.prepare: ") + sib_last_error());
    stop_ = false;
    completion_ = std::thread([this] { completion_main(); });
}

InferenceBolt::~InferenceBolt() {
    try { cleanup(); } catch (...) {}
}

void InferenceBolt::emit_front_nulls_locked() {
    while (!fifo_.empty() && fifo_.front().null_msg) {
        Pending p = fifo_.front();
        fifo_.pop_front();
        std::lock_guard<std::mutex> sync(collector_->monitor);
        collector_->emit(Values{});  // outputJson stayed null
        collector_->ack(p.tuple);
    }
}

void InferenceBolt::completion_main() {
    // One native pass formats a whole poll's messages into the arena (sib_poll_json_batch); they are then emitted and
    // acked under ONE hold of the collector's monitor instead of one lock, one buffer and one formatting call per tuple.
    // (sib.shared = 1 keeps the per-tuple path: results of a shared engine are copied per client, sib_client_poll.)
    constexpr int kMax = 256;
    std::vector<sib_result> res(kMax);
    std::vector<char> buf(1 << 14);
    std::vector<char> arena((size_t)kMax * 256);
    std::vector<uint64_t> tags(kMax);
    std::vector<uint32_t> offs(kMax + 1);
    std::vector<int32_t> status(kMax);
    std::unique_lock<std::mutex> lk(mu_);
    for (;;) {
        emit_front_nulls_locked();
        if (fifo_.empty()) {
            if (stop_) break;
            cv_.notify_all();  // cleanup() waits for "drained"
            cv_.wait(lk);
            continue;
        }
        lk.unlock();
        int n;
        if (shared_) n = sib_client_poll(client_, res.data(), kMax, 2000);
        else n = sib_poll_json_batch(engine_, arena.data(), arena.size(), tags.data(), offs.data(), status.data(), kMax, 2000);
        lk.lock();
        if (n > 0 && !shared_) {
            std::lock_guard<std::mutex> sync(collector_->monitor);  // synchronized (collector) { for (...) { emit; ack; } }
            for (int i = 0; i < n; i++) {
                while (!fifo_.empty() && fifo_.front().null_msg) {  // decode failures keep their place in the order
                    collector_->emit(Values{});
                    collector_->ack(fifo_.front().tuple);
                    fifo_.pop_front();
                }
                if (fifo_.empty()) break;  // cannot happen: every submitted tuple has a FIFO entry
                Values v;
                if (status[i] == SIB_OK && offs[i + 1] > offs[i]) {
                    v.is_null = false;
                    v.message.assign(arena.data() + offs[i], offs[i + 1] - offs[i]);
                }
                collector_->emit(v);
                collector_->ack(fifo_.front().tuple);
                fifo_.pop_front();
            }
        }
        for (int i = 0; shared_ && i < n; i++) {
            emit_front_nulls_locked();
            if (fifo_.empty()) break;
            Pending p = fifo_.front();
            fifo_.pop_front();
            Values v;
            if (res[i].status == SIB_OK) {
                size_t need = 64 + (size_t)res[i].n * ((size_t)res[i].n_classes * 18 + 4);
                if (buf.size() < need) buf.resize(need);
                long len = sib_format_json(&res[i], buf.data(), buf.size());
                if (len > 0) { v.is_null = false; v.message.assign(buf.data(), (size_t)len); }
            }
            std::lock_guard<std::mutex> sync(collector_->monitor);
            collector_->emit(v);
            collector_->ack(p.tuple);
        }
        if (n < 0 && stop_) break;
    }
    cv_.notify_all();
}

void InferenceBolt::execute(const Tuple& tuple) {
    const std::string& inputJson = tuple.getString(0);
    Tuple handle;
    handle.id = tuple.id;
    {
        // reserve the tuple's place in the emit order before the engine can possibly complete it
        std::lock_guard<std::mutex> lk(mu_);
        fifo_.push_back(Pending{handle, false});
        cv_.notify_all();
    }
    for (;;) {
        int rc = shared_ ? sib_client_submit_json(client_, seq_, inputJson.data(), inputJson.size())
                         : sib_submit_json(engine_, seq_, inputJson.data(), inputJson.size());
        if (rc > 0) {
            seq_++;
            return;
        }
        if (rc == SIB_EJSON) {  // catch (JsonProcessingException) -> outputJson = null; emit; ack
            std::lock_guard<std::mutex> lk(mu_);
            fifo_.back().null_msg = true;  // only this thread appends, so back() is still our entry
            cv_.notify_all();
            return;
        }
        if (rc == SIB_EAGAIN) {  // ring full: the completion thread is draining; retry
            std::this_thread::sleep_for(std::chrono::microseconds(20));
            continue;
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            fifo_.pop_back();
            cv_.notify_all();
        }
        if (rc == SIB_ESHAPE || rc == SIB_ECAP) throw IllegalArgumentException(sib_last_error());
        throw std::runtime_error(std::string("InferenceBolt.execute: ") + sib_last_error());
    }
}

void InferenceBolt::cleanup() {
    if (!engine_ && !client_) return;
    if (shared_) sib_client_flush(client_); else sib_flush(engine_);
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.notify_all();
        cv_.wait(lk, [this] { return fifo_.empty(); });
        stop_ = true;
        cv_.notify_all();
    }
    if (completion_.joinable()) completion_.join();
    if (shared_) sib_client_close(client_); else sib_engine_close(engine_);
    engine_ = nullptr;
    client_ = nullptr;
}

}  // namespace model
}  // namespace dke

// ------------------------------------------------------------------------------------------------ harness

using namespace dke::model;

namespace {
struct Event {
    int kind;
    uint64_t id;
    bool is_null;
    std::string msg;
    int64_t t_ns;
};
struct RecordingCollector : OutputCollector {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<Event> events;
    static int64_t now() {
        return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }
    void emit(const Values& v) override {
        std::lock_guard<std::mutex> lk(mu);
        events.push_back(Event{SIB_EV_EMIT, 0, v.is_null, v.message, now()});
        cv.notify_all();
    }
    void ack(const Tuple& t) override {
        std::lock_guard<std::mutex> lk(mu);
        events.push_back(Event{SIB_EV_ACK, t.id, true, std::string(), now()});
        cv.notify_all();
    }
};
}  // namespace

struct sib_harness {
    InferenceBolt bolt;
    RecordingCollector collector;
    bool prepared = false;
};

extern "C" int sib_harness_open(const void* blob, size_t n, const char* conf, int task_index, sib_harness** out) {
    if (!out || !blob) return SIB_EINVAL;
    *out = nullptr;
    Conf c;
    if (conf) {
        std::string s(conf);
        size_t pos = 0;
        while (pos < s.size()) {
            size_t e = s.find('\n', pos);
            if (e == std::string::npos) e = s.size();
            std::string line = s.substr(pos, e - pos);
            size_t eq = line.find('=');
            if (eq != std::string::npos) c[line.substr(0, eq)] = line.substr(eq + 1);
            pos = e + 1;
        }
    }
    sib_harness* h = new sib_harness();
    h->bolt.setModelBlob(blob, n);
    TopologyContext ctx;
    ctx.thisTaskIndex = task_index;
    try {
        h->bolt.prepare(c, ctx, &h->collector);
    } catch (const std::exception&) {
        delete h;
        return SIB_ECUDA;
    }
    h->prepared = true;
    *out = h;
    return SIB_OK;
}

extern "C" int sib_harness_execute(sib_harness* h, uint64_t id, const char* utf8, size_t len) {
    if (!h || !utf8) return SIB_EINVAL;
    Tuple t;
    t.id = id;
    t.field0.assign(utf8, len);
    try {
        h->bolt.execute(t);
    } catch (const IllegalArgumentException&) {
        return SIB_ESHAPE;
    } catch (const std::exception&) {
        return SIB_ECUDA;
    }
    return SIB_OK;
}

extern "C" int sib_harness_cleanup(sib_harness* h) {
    if (!h) return SIB_EINVAL;
    h->bolt.cleanup();
    return SIB_OK;
}

extern "C" int sib_harness_close(sib_harness* h) {
    if (!h) return SIB_EINVAL;
    h->bolt.cleanup();
    delete h;
    return SIB_OK;
}

extern "C" long sib_harness_event_count(sib_harness* h) {
    if (!h) return SIB_EINVAL;
    std::lock_guard<std::mutex> lk(h->collector.mu);
    return (long)h->collector.events.size();
}

extern "C" int sib_harness_event(sib_harness* h, long i, int* kind, uint64_t* id, const char** msg, size_t* len,
                                 int64_t* t_ns) {
    if (!h) return SIB_EINVAL;
    std::lock_guard<std::mutex> lk(h->collector.mu);
    if (i < 0 || (size_t)i >= h->collector.events.size()) return SIB_EINVAL;
    const Event& e = h->collector.events[(size_t)i];
    if (kind) *kind = e.kind;
    if (id) *id = e.id;
    if (msg) *msg = e.is_null ? nullptr : e.msg.data();
    if (len) *len = e.is_null ? 0 : e.msg.size();
    if (t_ns) *t_ns = e.t_ns;
    return SIB_OK;
}

extern "C" int sib_harness_output_fields(sib_harness* h, char* buf, size_t cap) {
    if (!h || !buf) return SIB_EINVAL;
    OutputFieldsDeclarer d;
    h->bolt.declareOutputFields(d);
    std::string s;
    for (size_t i = 0; i < d.fields.size(); i++) s += (i ? "," : "") + d.fields[i];
    if (s.size() + 1 > cap) return SIB_ECAP;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return (int)s.size();
}

extern "C" long sib_harness_wait_events(sib_harness* h, long n, long timeout_us) {
    if (!h) return SIB_EINVAL;
    std::unique_lock<std::mutex> lk(h->collector.mu);
    h->collector.cv.wait_for(lk, std::chrono::microseconds(timeout_us),
                             [&] { return (long)h->collector.events.size() >= n; });
    return (long)h->collector.events.size();
}
