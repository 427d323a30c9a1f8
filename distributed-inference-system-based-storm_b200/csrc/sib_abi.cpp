// sib_abi.cpp -- the stateless part of the C ABI (codec entry points, error text).
#include "sib_internal.hpp"

namespace sib {
const std::string& get_error();
}

extern "C" long sib_json_decode(const char* utf8, size_t len, float* out, long cap, int64_t dims[4]) {
    if (!utf8 || !out || !dims) return SIB_EINVAL;
    return sib::json_decode(utf8, len, out, cap, dims);
}

extern "C" long sib_json_encode(const float* probs, long n, int n_classes, char* buf, size_t cap) {
    if (!probs || !buf || n < 0 || n_classes <= 0) return SIB_EINVAL;
    return sib::json_encode(probs, n, n_classes, buf, cap);
}

extern "C" int sib_float_to_string(float x, char* buf24) { return sib::float_to_string(x, buf24); }

extern "C" const char* sib_last_error(void) { return sib::get_error().c_str(); }

extern "C" const char* sib_strerror(int code) {
    switch (code) {
        case SIB_OK: return "ok";
        case SIB_EJSON: return "malformed JSON (reference: emit null message, ack)";
        case SIB_ESHAPE: return "instances shape does not match the model (reference: IllegalArgumentException)";
        case SIB_ECAP: return "buffer too small / tuple larger than max_batch";
        case SIB_EAGAIN: return "all batch slots busy: poll and retry";
        case SIB_EINVAL: return "invalid argument or model blob";
        case SIB_ECUDA: return "CUDA failure or no GPU (no CPU fallback)";
        case SIB_ECLOSED: return "engine closed";
        default: return "unknown error";
    }
}

extern "C" const char* sib_version(void) { return "libsib 0.1 (sm_100a)"; }
