// sib_bolt.hpp -- C++ mirror of the reference's bolt interface for this path.
//
// The reference is Java on Storm 1.2.3 (InferenceBolt.java); neither a JDK nor Storm exists on the build or GPU
// boxes, so the host side above the C ABI is written here with the same names, argument meaning and error
// behaviour, and the Java drop-in (java/dke/model/InferenceBolt.java) is a line-for-line thin equivalent over JNI.
#pragma once
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/sib.h"

namespace dke {
namespace model {

// org.apache.storm.tuple.Tuple, as far as execute() uses it (InferenceBolt.java:71)
struct Tuple {
    uint64_t id = 0;
    std::string field0;
    const std::string& getString(int) const { return field0; }
};

// org.apache.storm.tuple.Values with one String field that may be null (InferenceBolt.java:98)
struct Values {
    bool is_null = true;
    std::string message;
};

// org.apache.storm.task.OutputCollector (1.x: NOT thread-safe; callers off the executor thread synchronise on it,
// as KafkaBolt.java:130-142 does)
struct OutputCollector {
    virtual ~OutputCollector() = default;
    virtual void emit(const Values& v) = 0;  // unanchored, default stream
    virtual void ack(const Tuple& t) = 0;
    std::mutex monitor;                      // "synchronized (collector)"
};

struct TopologyContext {
    int thisTaskIndex = 0;
    int getThisTaskIndex() const { return thisTaskIndex; }
};

struct OutputFieldsDeclarer {
    std::vector<std::string> fields;
    void declare(const std::vector<std::string>& f) { fields = f; }
};

using Conf = std::map<std::string, std::string>;

struct IllegalArgumentException : std::invalid_argument {
    using std::invalid_argument::invalid_argument;
};

class InferenceBolt {
public:
    InferenceBolt() = default;
    ~InferenceBolt();
    // The Java bolt reads the classpath resource model/...; the mirror is handed the same bytes.
    void setModelBlob(const void* blob, size_t n) { blob_.assign((const char*)blob, (const char*)blob + n); }

    void prepare(const Conf& stormConf, const TopologyContext& context, OutputCollector* collector);  // :43-62
    void execute(const Tuple& tuple);                                                                  // :69-100
    void declareOutputFields(OutputFieldsDeclarer& declarer) const { declarer.declare({"message"}); }   // :102-105
    void cleanup();  // added: flush + wait; the reference leaks its session instead

    sib_engine* engine() const { return engine_; }

private:
    struct Pending {
        Tuple tuple;     // only id is kept
        bool null_msg;   // decode failed: emit null, ack (InferenceBolt.java:92-99)
    };
    void completion_main();
    void emit_front_nulls_locked();

    std::string blob_;
    sib_engine* engine_ = nullptr;
    sib_client* client_ = nullptr;  // sib.shared = 1: a view of the worker process's shared engine instead
    bool shared_ = false;
    OutputCollector* collector_ = nullptr;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Pending> fifo_;  // arrival order == emit order
    std::thread completion_;
    bool stop_ = false;
    uint64_t seq_ = 0;
};

}  // namespace model
}  // namespace dke
