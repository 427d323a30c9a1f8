// fuzz_codec_diff.cpp -- differential fuzzer: the product's codec (csrc/sib_json.cpp) against the oracle's (oracle/sib_oracle.c)
// on mutated tuples. Test infrastructure (built and run by tests/test_fuzz_host.py): the oracle is the checker.
//   same verdict (count, SIB_EJSON, SIB_ESHAPE) for every input, same floats and dims bit for bit when both accept;
//   the same message text for every probability row; the same Float.toString text for every float.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../distributed-inference-system-based-storm_b200/csrc/sib_internal.hpp"

extern "C" {
long sibo_json_decode(const char* text, size_t len, float* out, long cap, long dims[4]);
int sibo_float_to_string(float x, char* buf);
long sibo_json_encode(const float* probs, long n, int ncls, char* buf, size_t cap);
}

namespace {
uint64_t g_s = 0x243F6A8885A308D3ull;
uint64_t rnd() {
    uint64_t z = (g_s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
size_t below(size_t n) { return n ? (size_t)(rnd() % n) : 0; }

const char* kTokens[] = {"[", "]", ",", "{", "}", ":", "\"instances\"", "\"", "e", "E", "-", "+", ".", "0", "1e400", "1e-400", "-0",
                         "null", "true", "NaN", "Infinity", " ", "\n", "\t", "\r", "\\u0041", "\\", "0x1p3", "1.00000005960464477550",
                         "9999999999999999999999999999999999999999", "[[[[", "]]]]", "\"predictions\"", "\xEF\xBB\xBF", "00", "01", "1.",
                         ".5", "-", "--1", "1e", "1e+", "1E-7", "4.9e-324", "1.17549435E-38", "3.4028235e38", "3.4028236e38", "1e39",
                         "0.1e1", "[]", "[[]]", "{}", ",,", "\"a\":1,", "\"instances\":null"};

void mutate(std::string& s) {
    const int rounds = 1 + (int)below(3);
    for (int r = 0; r < rounds; r++) {
        switch (below(7)) {
            case 0: if (!s.empty()) s[below(s.size())] ^= (char)(1u << below(7)); break;
            case 1: if (!s.empty()) s.erase(below(s.size()), 1 + below(4)); break;
            case 2: { const char* t = kTokens[below(sizeof(kTokens) / sizeof(*kTokens))]; s.insert(below(s.size() + 1), t); } break;
            case 3: if (!s.empty()) s.resize(below(s.size())); break;
            case 4: if (!s.empty()) { const size_t a = below(s.size()), n = 1 + below(16); s.insert(below(s.size() + 1), s.substr(a, n)); } break;
            case 5: if (s.size() > 1) { const size_t a = below(s.size()), b = below(s.size()); std::swap(s[a], s[b]); } break;
            case 6: if (!s.empty()) { const char* t = kTokens[below(sizeof(kTokens) / sizeof(*kTokens))]; const size_t a = below(s.size()); s.replace(a, std::min<size_t>(1 + below(6), s.size() - a), t); } break;
        }
    }
}

float random_float() {
    switch (below(8)) {
        case 0: { uint32_t b = (uint32_t)rnd(); float f; std::memcpy(&f, &b, 4); return f; }
        case 1: return (float)std::ldexp(1.0, (int)below(300) - 150);
        case 2: return (float)std::pow(10.0, (double)((int)below(80) - 45));
        case 3: return below(2) ? 0.0f : -0.0f;
        case 4: { uint32_t b = (uint32_t)below(1u << 23); float f; std::memcpy(&f, &b, 4); return f; }
        case 5: return (float)((double)rnd() / 18446744073709551616.0);
        case 6: { float f = (float)((double)rnd() / 18446744073709551616.0); uint32_t b; std::memcpy(&b, &f, 4); b += (uint32_t)below(3) - 1; std::memcpy(&f, &b, 4); return f; }
        default: return (float)((double)(int64_t)rnd() * 1e-12);
    }
}

std::string number_text(float v) {
    char buf[96];
    if (!std::isfinite(v)) return "0.25";
    switch (below(5)) {
        case 0: std::snprintf(buf, sizeof buf, "%.17g", (double)v); break;            // json.dumps of a double
        case 1: std::snprintf(buf, sizeof buf, "%.9g", (double)v); break;
        case 2: std::snprintf(buf, sizeof buf, "%.40f", (double)v); break;            // long plain decimals (slow path)
        case 3: std::snprintf(buf, sizeof buf, "%.25e", (double)v); break;
        default: buf[sib::float_to_string(v, buf)] = 0; break;                        // what a Java producer writes
    }
    std::string s = buf;
    if (s.find("inf") != std::string::npos || s.find("nan") != std::string::npos) return "0.25";
    return s;
}

std::string valid_tuple(int n, int h, int w, int c) {
    std::string s = below(2) ? "{\"instances\":" : " {\n \"instances\" :\t";
    s += '[';
    for (int i = 0; i < n; i++) {
        s += i ? ",[" : "[";
        for (int y = 0; y < h; y++) {
            s += y ? ",[" : "[";
            for (int x = 0; x < w; x++) {
                s += x ? (below(8) ? ",[" : " , [ ") : "[";
                for (int k = 0; k < c; k++) {
                    if (k) s += ',';
                    s += number_text(random_float());
                }
                s += ']';
            }
            s += ']';
        }
        s += ']';
    }
    s += below(4) ? "]}" : "] } trailing";
    return s;
}
}  // namespace

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? std::atof(argv[1]) : 10.0;
    if (argc > 2) g_s = std::strtoull(argv[2], nullptr, 0);
    const auto t0 = std::chrono::steady_clock::now();
    auto elapsed = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    unsigned long long it = 0, both_ok = 0, both_rej = 0;
    const long cap = 4096;
    std::vector<float> a((size_t)cap), b((size_t)cap);
    while (elapsed() < seconds) {
        it++;
        const int n = 1 + (int)below(2), h = 1 + (int)below(3), w = 1 + (int)below(3), c = 1 + (int)below(3);
        std::string s = valid_tuple(n, h, w, c);
        if (below(4)) mutate(s);
        std::vector<char> text(s.begin(), s.end());
        int64_t da[4] = {0, 0, 0, 0};
        long db[4] = {0, 0, 0, 0};
        const long ra = sib::json_decode(text.data(), text.size(), a.data(), cap, da);
        const long rb = sibo_json_decode(text.data(), text.size(), b.data(), cap, db);
        if (ra != rb) {
            std::fprintf(stderr, "VERDICT differs: product %ld, oracle %ld on (%zu bytes): %s\n", ra, rb, s.size(), s.c_str());
            return 1;
        }
        if (ra >= 0) {
            both_ok++;
            if (std::memcmp(a.data(), b.data(), (size_t)ra * 4) != 0 || da[0] != db[0] || da[1] != db[1] || da[2] != db[2] || da[3] != db[3]) {
                std::fprintf(stderr, "VALUES differ on: %s\n", s.c_str());
                return 1;
            }
        } else {
            both_rej++;
        }
        // encoder and printer
        const int nc = 1 + (int)below(12), rows = 1 + (int)below(2);
        std::vector<float> p((size_t)rows * nc);
        for (auto& v : p) v = random_float();
        char ea[1024], eb[1024];
        const long la = sib::json_encode(p.data(), rows, nc, ea, sizeof ea), lb = sibo_json_encode(p.data(), rows, nc, eb, sizeof eb);
        if (la != lb || (la > 0 && std::memcmp(ea, eb, (size_t)la) != 0)) {
            std::fprintf(stderr, "ENCODE differs (%ld vs %ld): %.*s | %.*s\n", la, lb, (int)std::max(0l, la), ea, (int)std::max(0l, lb), eb);
            return 1;
        }
        const float v = random_float();
        char fa[40], fb[40];
        const int na = sib::float_to_string(v, fa), nb = sibo_float_to_string(v, fb);
        if (na != nb || std::memcmp(fa, fb, (size_t)na) != 0) {
            std::fprintf(stderr, "FLOAT differs for %a: %.*s | %.*s\n", v, na, fa, nb, fb);
            return 1;
        }
    }
    std::printf("{\"seconds\": %.1f, \"rounds\": %llu, \"both_accept\": %llu, \"both_reject\": %llu, \"differences\": 0}\n", elapsed(), it, both_ok, both_rej);
    return 0;
}
