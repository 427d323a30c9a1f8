// fuzz_host.cpp -- mutation fuzzer for the host-side parsers of libsib.so under AddressSanitizer + UBSan (no GPU needed).
//
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -I include \
//       tools/fuzz_host.cpp distributed-inference-system-based-storm_b200/csrc/{sib_json,sib_pbimport,sib_model}.cpp \
//       -o tools/fuzz_host && tools/fuzz_host <seconds> [seed.pb ...]
//
// Targets (DESIGN.md 2 "a blob or an error, never a crash"):
//   json_decode     the tuple decoder (InferenceBolt.java:76-80): valid tuples of small shapes, mutated; output capacities
//                   from 0 to exact; the decoder may answer anything but must stay inside `out[cap]` and `utf8[len]`
//   json_encode     the message encoder (:89-91): random bit patterns as probabilities, every capacity from 0 up
//   float_to_string every float bit pattern class (zeros, denormals, powers of two and ten, NaN, Inf)
//   import_pb       SavedModel bytes (:49-58): the seed files given on the command line, mutated
//   parse_model     the model blob: the blobs import_pb produced, mutated
// A finding is a sanitizer report (the process aborts); the exit code is 0 when the time is up.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "../distributed-inference-system-based-storm_b200/csrc/sib_internal.hpp"

namespace {

uint64_t g_s = 0x9E3779B97F4A7C15ull;
uint64_t rnd() {  // splitmix64
    uint64_t z = (g_s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
size_t below(size_t n) { return n ? (size_t)(rnd() % n) : 0; }

const char* kTokens[] = {"[", "]", ",", "{", "}", ":", "\"instances\"", "\"", "e", "E", "-", "+", ".", "0", "1e400", "1e-400", "-0",
                         "null", "true", "NaN", "Infinity", " ", "\n", "\t", "\\u0041", "\\", "0x1p3", "1.00000005960464477550",
                         "9999999999999999999999999999999999999999", "[[[[", "]]]]", "\"predictions\"", "\xEF\xBB\xBF", "\0"};

void mutate(std::string& s) {
    const int rounds = 1 + (int)below(4);
    for (int r = 0; r < rounds; r++) {
        switch (below(8)) {
            case 0: if (!s.empty()) s[below(s.size())] ^= (char)(1u << below(8)); break;                       // bit flip
            case 1: if (!s.empty()) s[below(s.size())] = (char)rnd(); break;                                   // random byte
            case 2: if (!s.empty()) s.erase(below(s.size()), 1 + below(8)); break;                             // delete
            case 3: { const char* t = kTokens[below(sizeof(kTokens) / sizeof(*kTokens))]; s.insert(below(s.size() + 1), t); } break;
            case 4: if (!s.empty()) s.resize(below(s.size()));  break;                                         // truncate
            case 5: if (!s.empty()) { const size_t a = below(s.size()), n = 1 + below(32); s.insert(below(s.size() + 1), s.substr(a, n)); } break;  // duplicate
            case 6: if (s.size() > 1) { const size_t a = below(s.size()), b = below(s.size()); std::swap(s[a], s[b]); } break;
            case 7: if (!s.empty()) { const size_t a = below(s.size()); const size_t n = std::min<size_t>(1 + below(64), s.size() - a); for (size_t i = 0; i < n; i++) s[a + i] = (char)('0' + below(10)); } break;
        }
    }
}

float random_float() {
    switch (below(8)) {
        case 0: { uint32_t b = (uint32_t)rnd(); float f; std::memcpy(&f, &b, 4); return f; }  // any bit pattern
        case 1: return (float)std::ldexp(1.0, (int)below(300) - 150);
        case 2: return (float)std::pow(10.0, (double)((int)below(80) - 45));
        case 3: return 0.0f;
        case 4: return -0.0f;
        case 5: { uint32_t b = (uint32_t)below(1u << 23); float f; std::memcpy(&f, &b, 4); return f; }  // denormal
        case 6: return (float)((double)rnd() / 18446744073709551616.0);
        default: return (float)((double)(int64_t)rnd() * 1e-12);
    }
}

std::string valid_tuple(int n, int h, int w, int c) {
    std::string s = below(2) ? "{\"instances\":" : "{ \"instances\" : ";
    char buf[64];
    s += '[';
    for (int i = 0; i < n; i++) {
        s += i ? ",[" : "[";
        for (int y = 0; y < h; y++) {
            s += y ? ",[" : "[";
            for (int x = 0; x < w; x++) {
                s += x ? ",[" : "[";
                for (int k = 0; k < c; k++) {
                    if (k) s += ',';
                    const float v = random_float();
                    if (std::isfinite(v)) {
                        if (below(3) == 0) std::snprintf(buf, sizeof buf, "%.17g", (double)v);   // json.dumps-style doubles
                        else buf[sib::float_to_string(v, buf)] = 0;
                        s += buf;
                    } else {
                        s += "0.5";
                    }
                }
                s += ']';
            }
            s += ']';
        }
        s += ']';
    }
    s += "]}";
    return s;
}

std::vector<uint8_t> read_file(const char* path) {
    std::ifstream f(path, std::ios::binary);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

}  // namespace

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? std::atof(argv[1]) : 10.0;
    if (const char* sd = std::getenv("FUZZ_SEED")) g_s = std::strtoull(sd, nullptr, 0);  // another stream of mutations
    std::vector<std::vector<uint8_t>> pb_seeds, blob_seeds;
    for (int i = 2; i < argc; i++) {
        auto b = read_file(argv[i]);
        if (b.empty()) { std::fprintf(stderr, "cannot read %s\n", argv[i]); return 2; }
        pb_seeds.push_back(b);
        std::vector<uint8_t> blob;
        std::string err;
        if (sib::import_pb(b.data(), b.size(), &blob, &err) == SIB_OK) blob_seeds.push_back(blob);
    }
    const auto t0 = std::chrono::steady_clock::now();
    auto elapsed = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    uint64_t it = 0, dec_ok = 0, dec_err = 0, pb_ok = 0, pb_err = 0, blob_ok = 0, blob_err = 0, enc = 0, rt = 0;
    while (elapsed() < seconds) {
        it++;
        // ---- json_decode
        {
            const int n = 1 + (int)below(2), h = 1 + (int)below(4), w = 1 + (int)below(4), c = 1 + (int)below(3);
            std::string s = valid_tuple(n, h, w, c);
            const long elems = (long)n * h * w * c;
            if (below(5)) mutate(s);
            // exact-size heap copies: one byte past either end is a sanitizer report
            std::vector<char> text(s.begin(), s.end());
            const long cap = below(4) ? elems : (long)below((size_t)elems + 1);
            std::vector<float> out((size_t)cap);
            int64_t dims[4] = {0, 0, 0, 0};
            const long r = sib::json_decode(text.data(), text.size(), out.data(), cap, dims);
            if (r >= 0) {
                dec_ok++;
                if (r > cap) { std::fprintf(stderr, "json_decode returned %ld > cap %ld\n", r, cap); return 1; }
                if (dims[0] * dims[1] * dims[2] * dims[3] != r) { std::fprintf(stderr, "dims do not multiply to the count\n"); return 1; }
            } else {
                dec_err++;
            }
        }
        // ---- json_encode / float_to_string round trip
        {
            const int nc = 1 + (int)below(12), n = 1 + (int)below(3);
            std::vector<float> p((size_t)n * nc);
            for (auto& v : p) v = random_float();
            const size_t cap = below(3) ? 32 + (size_t)n * ((size_t)nc * 18 + 4) : below(64);
            std::vector<char> buf(cap);
            const long len = sib::json_encode(p.data(), n, nc, buf.data(), cap);
            if (len > (long)cap) { std::fprintf(stderr, "json_encode wrote %ld into %zu\n", len, cap); return 1; }
            enc++;
            char fb[32];
            const float v = random_float();
            const int l = sib::float_to_string(v, fb);
            if (l <= 0 || l > 24) { std::fprintf(stderr, "float_to_string length %d\n", l); return 1; }
            if (std::isfinite(v)) {  // Float.toString's contract: the text reads back as the same float
                fb[l] = 0;
                const float back = std::strtof(fb, nullptr);
                if (std::memcmp(&back, &v, 4) != 0 && !(back == 0.f && v == 0.f)) { std::fprintf(stderr, "float_to_string(%a) = %s reads back %a\n", v, fb, back); return 1; }
                rt++;
            }
        }
        // ---- import_pb / parse_model (slower: every 16th round)
        if ((it & 15) == 0 && !pb_seeds.empty()) {
            const auto& seed = pb_seeds[below(pb_seeds.size())];
            std::string s(seed.begin(), seed.end());
            // most mutations should land in the structure (the first bytes of messages), not in the weight payload
            const int rounds = 1 + (int)below(3);
            for (int r = 0; r < rounds; r++) {
                const size_t at = below(4) ? below(std::min<size_t>(s.size(), 4096)) : below(s.size());
                switch (below(4)) {
                    case 0: s[at] ^= (char)(1u << below(8)); break;
                    case 1: s[at] = (char)rnd(); break;
                    case 2: s.erase(at, 1 + below(16)); break;
                    case 3: s.insert(at, 1 + below(8), (char)rnd()); break;
                }
                if (s.empty()) s = "\x0a";
            }
            std::vector<uint8_t> bytes(s.begin(), s.end());
            std::vector<uint8_t> blob;
            std::string err;
            if (sib::import_pb(bytes.data(), bytes.size(), &blob, &err) == SIB_OK) {
                pb_ok++;
                sib::ModelDesc m;
                if (sib::parse_model(blob.data(), blob.size(), &m, &err) != SIB_OK) { std::fprintf(stderr, "import_pb produced a blob parse_model rejects: %s\n", err.c_str()); return 1; }
            } else {
                pb_err++;
            }
        }
        if ((it & 15) == 8 && !blob_seeds.empty()) {
            const auto& seed = blob_seeds[below(blob_seeds.size())];
            std::vector<uint8_t> b(seed);
            const int rounds = 1 + (int)below(3);
            for (int r = 0; r < rounds; r++) {
                const size_t at = below(4) ? below(std::min<size_t>(b.size(), 32 + 16 * 32)) : below(b.size());
                if (below(3) == 0) { uint32_t v = (uint32_t)rnd() >> below(32); std::memcpy(&b[at & ~(size_t)3], &v, std::min<size_t>(4, b.size() - (at & ~(size_t)3))); }
                else if (below(2)) b[at] ^= (uint8_t)(1u << below(8));
                else b.resize(below(b.size()) + 1);
            }
            sib::ModelDesc m;
            std::string err;
            if (sib::parse_model(b.data(), b.size(), &m, &err) == SIB_OK) blob_ok++; else blob_err++;
        }
    }
    std::printf("{\"seconds\": %.1f, \"rounds\": %llu, \"json_decode_ok\": %llu, \"json_decode_rejected\": %llu, \"json_encode\": %llu, "
                "\"float_round_trips\": %llu, \"import_pb_ok\": %llu, \"import_pb_rejected\": %llu, \"parse_model_ok\": %llu, "
                "\"parse_model_rejected\": %llu, \"findings\": 0}\n",
                elapsed(), (unsigned long long)it, (unsigned long long)dec_ok, (unsigned long long)dec_err, (unsigned long long)enc,
                (unsigned long long)rt, (unsigned long long)pb_ok, (unsigned long long)pb_err, (unsigned long long)blob_ok,
                (unsigned long long)blob_err);
    return 0;
}
