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

// ---- the GPU decode kernel's acceptance logic (csrc/sib_kernels_json.cu), restated sequentially: token starts, the
// number grammar with the exact fast path's limits, the punctuation between one token's end and the next one's start.
// The CUDA kernel itself is checked on the GPU (tests/test_gpu_json.py); this checks the ALGORITHM on far more inputs:
// whatever it accepts, the host decoder must accept as one image of this shape with the same bits.
namespace gpu_logic {
bool is_ws(unsigned char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r'; }
bool is_digit(unsigned char c) { return (unsigned)(c - '0') <= 9u; }
bool is_num(unsigned char c) { return is_digit(c) || c == '-' || c == '+' || c == '.' || c == 'e' || c == 'E'; }
bool is_start(const unsigned char* s, int i) { return (is_digit(s[i]) || s[i] == '-') && !(i > 0 && is_num(s[i - 1])); }
bool back_expect(const unsigned char* s, int& q, unsigned char ch) {
    while (q >= 0 && is_ws(s[q])) q--;
    if (q < 0 || s[q] != ch) return false;
    q--;
    return true;
}
bool fwd_expect(const unsigned char* s, int len, int& q, unsigned char ch) {
    while (q < len && is_ws(s[q])) q++;
    if (q >= len || s[q] != ch) return false;
    q++;
    return true;
}
const double pow10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// returns true when the kernel would leave the tuple's flag at 0 (decoded on the GPU)
bool decode(const unsigned char* s, int len, int H, int W, int Cc, float* dst) {
    const int elems = H * W * Cc;
    if (len > 56 * 1024 || len < 16 || elems > 4096) return false;
    std::vector<int> pos, endp;
    for (int i = 0; i < len; i++) if (is_start(s, i)) pos.push_back(i);
    if ((int)pos.size() != elems) return false;
    endp.resize((size_t)elems);
    bool bad = false;
    for (int k = 0; k < elems; k++) {
        int p = pos[(size_t)k];
        bool neg = false;
        if (s[p] == '-') { neg = true; p++; }
        unsigned long long mant = 0;
        int nd = 0, e10 = 0;
        if (p < len && s[p] == '0') { p++; if (p < len && is_digit(s[p])) bad = true; }
        else if (p < len && is_digit(s[p])) { while (p < len && is_digit(s[p])) { mant = mant * 10ull + (unsigned)(s[p] - '0'); nd++; p++; if (nd > 18) break; } }
        else bad = true;
        if (p < len && s[p] == '.') {
            p++;
            if (p >= len || !is_digit(s[p])) bad = true;
            while (p < len && is_digit(s[p]) && nd <= 18) {
                const unsigned d = (unsigned)(s[p] - '0');
                if (mant != 0 || d != 0) nd++;
                mant = mant * 10ull + d;
                e10--;
                p++;
            }
        }
        if (p < len && (s[p] == 'e' || s[p] == 'E')) {
            p++;
            bool eneg = false;
            if (p < len && (s[p] == '+' || s[p] == '-')) { eneg = s[p] == '-'; p++; }
            if (p >= len || !is_digit(s[p])) bad = true;
            int ex = 0;
            while (p < len && is_digit(s[p])) { if (ex < 10000) ex = ex * 10 + (s[p] - '0'); p++; }
            e10 += eneg ? -ex : ex;
        }
        if (p < len && is_num(s[p])) bad = true;
        if (nd > 15 || mant > (1ull << 53) || e10 < -22 || e10 > 22) bad = true;
        double d = (double)mant;
        const int ae = e10 < 0 ? -e10 : e10;
        const double pw = pow10[ae > 22 ? 22 : ae];
        d = e10 < 0 ? d / pw : d * pw;
        dst[k] = (float)(neg ? -d : d);
        endp[(size_t)k] = p;
    }
    for (int k = 0; k < elems; k++) {
        const int i = pos[(size_t)k], p = endp[(size_t)k];
        int q = i - 1;
        if (k == 0) {
            bool ok = back_expect(s, q, '[') && back_expect(s, q, '[') && back_expect(s, q, '[') && back_expect(s, q, '[') &&
                      back_expect(s, q, ':') && back_expect(s, q, '"');
            const char key[] = "instances";
            for (int c = 8; ok && c >= 0; c--) { if (q < 0 || s[q] != (unsigned char)key[c]) ok = false; q--; }
            ok = ok && q >= 0 && s[q] == '"';
            q--;
            ok = ok && back_expect(s, q, '{');
            while (ok && q >= 0) { if (!is_ws(s[q])) ok = false; q--; }
            bad |= !ok;
        } else {
            const int c_i = k % Cc, w_i = (k / Cc) % W;
            bool ok;
            if (c_i != 0) ok = back_expect(s, q, ',');
            else if (w_i != 0) ok = back_expect(s, q, '[') && back_expect(s, q, ',') && back_expect(s, q, ']');
            else ok = back_expect(s, q, '[') && back_expect(s, q, '[') && back_expect(s, q, ',') && back_expect(s, q, ']') && back_expect(s, q, ']');
            while (ok && q >= 0 && is_ws(s[q])) q--;
            ok = ok && q + 1 == endp[(size_t)k - 1];
            bad |= !ok;
        }
        if (k == elems - 1) {
            int f = p;
            bool ok = fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, '}');
            bad |= !ok;
        }
    }
    return !bad;
}
}  // namespace gpu_logic

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? std::atof(argv[1]) : 10.0;
    if (argc > 2) g_s = std::strtoull(argv[2], nullptr, 0);
    const auto t0 = std::chrono::steady_clock::now();
    auto elapsed = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    unsigned long long it = 0, both_ok = 0, both_rej = 0, gpu_ok = 0, gpu_declined_valid = 0;
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
        // the GPU kernel's acceptance logic against the host decoder, for the shape the tuple was generated with and
        // for a transposed one of the same element count
        for (int variant = 0; variant < 2; variant++) {
            const int H = variant ? w : h, W = variant ? h : w;
            if (variant && h == w) break;
            std::vector<float> g((size_t)h * w * c + 1);
            if (gpu_logic::decode(reinterpret_cast<const unsigned char*>(text.data()), (int)text.size(), H, W, c, g.data())) {
                gpu_ok++;
                const long want = (long)H * W * c;
                if (ra != want || da[0] != 1 || da[1] != H || da[2] != W || da[3] != c || std::memcmp(g.data(), a.data(), (size_t)want * 4) != 0) {
                    std::fprintf(stderr, "GPU LOGIC accepts as 1x%dx%dx%d what the host decoder reads as %ld [%lld,%lld,%lld,%lld]%s: %s\n", H, W, c, ra,
                                 (long long)da[0], (long long)da[1], (long long)da[2], (long long)da[3], ra == want ? " with other bits" : "", s.c_str());
                    return 1;
                }
            }
        }
        // ... and tuples made for that logic: one image, shortest-form numbers, one light mutation, so that most are
        // accepted or fail on exactly one token
        {
            const int H = 1 + (int)below(4), W = 1 + (int)below(4), C = 1 + (int)below(3);
            std::string g = "{\"instances\":[";
            g += '[';
            for (int y = 0; y < H; y++) {
                g += y ? ",[" : "[";
                for (int x = 0; x < W; x++) {
                    g += x ? ",[" : "[";
                    for (int k = 0; k < C; k++) {
                        if (k) g += below(16) ? "," : " , ";
                        char nb[64];
                        const float v = random_float();
                        if (!std::isfinite(v) || below(4) == 0) std::snprintf(nb, sizeof nb, "%u.%u", (unsigned)below(256), (unsigned)below(1000));
                        else nb[sib::float_to_string(v, nb)] = 0;
                        g += nb;
                    }
                    g += ']';
                }
                g += ']';
            }
            g += "]]}";
            if (below(10) < 7) {
                const char* t = kTokens[below(sizeof(kTokens) / sizeof(*kTokens))];
                switch (below(4)) {
                    case 0: g.insert(below(g.size() + 1), t); break;
                    case 1: g.erase(below(g.size()), 1); break;
                    case 2: { const size_t at = below(g.size()); g.replace(at, 1, t); } break;
                    case 3: { const size_t x = below(g.size()), y = below(g.size()); std::swap(g[x], g[y]); } break;
                }
            }
            std::vector<char> gt(g.begin(), g.end());
            std::vector<float> gv((size_t)H * W * C + 1);
            int64_t dg[4] = {0, 0, 0, 0};
            const long rh = sib::json_decode(gt.data(), gt.size(), a.data(), cap, dg);
            if (gpu_logic::decode(reinterpret_cast<const unsigned char*>(gt.data()), (int)gt.size(), H, W, C, gv.data())) {
                gpu_ok++;
                const long want = (long)H * W * C;
                if (rh != want || dg[0] != 1 || dg[1] != H || dg[2] != W || dg[3] != C || std::memcmp(gv.data(), a.data(), (size_t)want * 4) != 0) {
                    std::fprintf(stderr, "GPU LOGIC accepts as 1x%dx%dx%d what the host decoder reads as %ld [%lld,%lld,%lld,%lld]%s: %s\n", H, W, C, rh,
                                 (long long)dg[0], (long long)dg[1], (long long)dg[2], (long long)dg[3], rh == want ? " with other bits" : "", g.c_str());
                    return 1;
                }
            } else if (rh == (long)H * W * C && dg[0] == 1 && dg[1] == H && dg[2] == W && dg[3] == C) {
                gpu_declined_valid++;   // allowed (the host decoder takes over), counted to see the fast path's reach
            }
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
    std::printf("{\"seconds\": %.1f, \"rounds\": %llu, \"both_accept\": %llu, \"both_reject\": %llu, \"gpu_logic_accepts\": %llu, \"gpu_logic_declines_valid\": %llu, \"differences\": 0}\n", elapsed(), it, both_ok, both_rej, gpu_ok, gpu_declined_valid);
    return 0;
}
