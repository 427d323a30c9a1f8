// sib_json.cpp -- native codec for the bolt's two DTOs.
//   decode: {"instances": float[N][H][W][C]}  (InstObj.java:7-16, read at InferenceBolt.java:76-77), written straight
//           into a caller buffer (the engine passes its pinned batch slot, replacing Tensor.create's copy, :80)
//   encode: {"predictions": float[N][10]}     (PredObj.java:8-17, written at InferenceBolt.java:89-91)
// Number semantics follow the libraries the reference calls:
//   Jackson 2.10 ParserBase.getFloatValue(): text -> double (correctly rounded) -> (float)   [double rounding kept]
//   java.lang.Float.toString: shortest decimal that round-trips, >= 2 digits, plain for 1e-3 <= |x| < 1e7 else d.dE[-]n
#include <charconv>
#include <cmath>
#include <cstring>

#include "sib_internal.hpp"

namespace sib {
namespace {

struct Cur {
    const char* p;
    const char* end;
};

inline void skip_ws(Cur& c) {
    while (c.p < c.end) {
        unsigned char ch = (unsigned char)*c.p;
        if (ch == ' ' || ch == '\n' || ch == '\t' || ch == '\r') c.p++;
        else break;
    }
}

const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                           1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// RFC 8259 number -> float via double. Fast path (Clinger): <= 15 significant digits and |e10| <= 22 gives the
// correctly rounded double with one exact multiply/divide; everything else goes to std::from_chars<double>.
inline int parse_number(Cur& c, float* out) {
    const char* s = c.p;
    const char* q = s;
    const char* end = c.end;
    bool neg = false;
    if (q < end && *q == '-') { neg = true; q++; }
    if (q >= end) return SIB_EJSON;
    uint64_t mant = 0;
    int nd = 0;        // significant digits accumulated in mant
    int dropped = 0;   // digits not accumulated (precision overflow) -> slow path
    int e10 = 0;
    if (*q == '0') {
        q++;
    } else if (*q >= '1' && *q <= '9') {
        while (q < end && (unsigned)(*q - '0') <= 9u) {
            if (nd < 18) { mant = mant * 10 + (unsigned)(*q - '0'); nd++; } else { dropped++; e10++; }
            q++;
        }
    } else {
        return SIB_EJSON;
    }
    if (q < end && *q == '.') {
        q++;
        if (q >= end || (unsigned)(*q - '0') > 9u) return SIB_EJSON;
        // Eight fraction digits at a time (a pixel is typically 0.dddddddd): one 64-bit load, a mask test for "all
        // digits", three multiplies. nd may over-count leading zeros here; that can only send a number to the slow
        // path, which is correctly rounded as well, so the result is the same either way.
        while (end - q >= 8 && nd + 8 <= 18) {
            uint64_t v;
            std::memcpy(&v, q, 8);
            if ((((v & 0xF0F0F0F0F0F0F0F0ull) | (((v + 0x0606060606060606ull) & 0xF0F0F0F0F0F0F0F0ull) >> 4)) != 0x3333333333333333ull)) break;
            v -= 0x3030303030303030ull;
            v = (v * 10) + (v >> 8);
            v = (((v & 0x000000FF000000FFull) * (100ull + (1000000ull << 32))) + (((v >> 16) & 0x000000FF000000FFull) * (1ull + (10000ull << 32)))) >> 32;
            if (mant != 0 || v != 0) nd += 8;
            mant = mant * 100000000ull + v;
            e10 -= 8;
            q += 8;
        }
        while (q < end && (unsigned)(*q - '0') <= 9u) {
            if (nd < 18) {
                unsigned d = (unsigned)(*q - '0');
                if (mant != 0 || d != 0) nd++;  // leading zeros of 0.000x do not consume precision
                mant = mant * 10 + d;
                e10--;
            } else {
                dropped++;
            }
            q++;
        }
    }
    if (q < end && (*q == 'e' || *q == 'E')) {
        q++;
        bool eneg = false;
        if (q < end && (*q == '+' || *q == '-')) { eneg = (*q == '-'); q++; }
        if (q >= end || (unsigned)(*q - '0') > 9u) return SIB_EJSON;
        int ex = 0;
        while (q < end && (unsigned)(*q - '0') <= 9u) {
            if (ex < 100000) ex = ex * 10 + (*q - '0');
            q++;
        }
        e10 += eneg ? -ex : ex;
    }
    double d;
    if (dropped == 0 && mant <= (1ull << 53) && e10 >= -22 && e10 <= 22) {
        d = (double)mant;
        if (e10 < 0) d /= kPow10[-e10];
        else d *= kPow10[e10];
    } else {
        const char* b = neg ? s + 1 : s;
        auto r = std::from_chars(b, q, d, std::chars_format::general);
        if (r.ec == std::errc::result_out_of_range) {
            // Double.parseDouble saturates: overflow -> Infinity, underflow -> 0
            d = (e10 + nd > 0) ? HUGE_VAL : 0.0;
        } else if (r.ec != std::errc()) {
            return SIB_EJSON;
        }
    }
    *out = (float)(neg ? -d : d);
    c.p = q;
    return 0;
}

inline bool match_lit(Cur& c, const char* lit, size_t n) {
    if ((size_t)(c.end - c.p) >= n && std::memcmp(c.p, lit, n) == 0) { c.p += n; return true; }
    return false;
}

struct ArrState {
    int64_t dims[4];
    bool first[4];
    bool ragged;
    float* out;
    long cap, count;
    bool overflow;
};

// depth 0..3. Syntax errors win over shape errors (Jackson finishes parsing before Tensor.create sees the arrays).
int parse_array(Cur& c, int depth, ArrState& st) {
    skip_ws(c);
    if (c.p >= c.end) return SIB_EJSON;
    if (*c.p != '[') {
        if (match_lit(c, "null", 4)) { st.ragged = true; return 0; }  // null row -> NPE in Tensor.create
        return SIB_EJSON;                                            // MismatchedInputException
    }
    c.p++;
    int64_t n = 0;
    skip_ws(c);
    if (c.p < c.end && *c.p == ']') {
        c.p++;
    } else {
        for (;;) {
            if (depth == 3) {
                skip_ws(c);
                float v;
                if (c.p < c.end && *c.p == 'n') {
                    if (!match_lit(c, "null", 4)) return SIB_EJSON;
                    v = 0.0f;  // Jackson: null in a primitive float array reads as 0.0f
                } else {
                    int r = parse_number(c, &v);
                    if (r) return r;
                }
                if (st.count < st.cap) st.out[st.count] = v;
                else st.overflow = true;
                st.count++;
            } else {
                int r = parse_array(c, depth + 1, st);
                if (r) return r;
            }
            n++;
            skip_ws(c);
            if (c.p >= c.end) return SIB_EJSON;
            if (*c.p == ',') { c.p++; continue; }
            if (*c.p == ']') { c.p++; break; }
            return SIB_EJSON;
        }
    }
    if (st.first[depth]) { st.dims[depth] = n; st.first[depth] = false; }
    else if (st.dims[depth] != n) st.ragged = true;
    return 0;
}

}  // namespace

long json_decode(const char* text, size_t len, float* out, long cap, int64_t dims[4]) {
    Cur c{text, text + len};
    ArrState st{};
    bool seen = false;
    skip_ws(c);
    if (c.p >= c.end || *c.p != '{') return SIB_EJSON;
    c.p++;
    skip_ws(c);
    if (c.p < c.end && *c.p == '}') {
        c.p++;
    } else {
        for (;;) {
            skip_ws(c);
            if (c.p >= c.end || *c.p != '"') return SIB_EJSON;
            const char* k = ++c.p;
            while (c.p < c.end && *c.p != '"') {
                if (*c.p == '\\') return SIB_EJSON;
                c.p++;
            }
            if (c.p >= c.end) return SIB_EJSON;
            size_t kl = (size_t)(c.p - k);
            c.p++;
            // any other property: UnrecognizedPropertyException (FAIL_ON_UNKNOWN_PROPERTIES defaults to true)
            if (kl != 9 || std::memcmp(k, "instances", 9) != 0) return SIB_EJSON;
            skip_ws(c);
            if (c.p >= c.end || *c.p != ':') return SIB_EJSON;
            c.p++;
            st = ArrState{};
            for (int i = 0; i < 4; i++) st.first[i] = true;
            st.out = out;
            st.cap = cap;
            int r = parse_array(c, 0, st);
            if (r) return r;
            seen = true;
            skip_ws(c);
            if (c.p >= c.end) return SIB_EJSON;
            if (*c.p == ',') { c.p++; continue; }
            if (*c.p == '}') { c.p++; break; }
            return SIB_EJSON;
        }
    }
    // Trailing bytes after the root value are ignored (Jackson FAIL_ON_TRAILING_TOKENS defaults to false).
    if (!seen) return SIB_ESHAPE;  // instances stays null -> NullPointerException in Tensor.create
    if (st.ragged || st.first[3] || st.count == 0) return SIB_ESHAPE;
    if (st.dims[0] * st.dims[1] * st.dims[2] * st.dims[3] != st.count) return SIB_ESHAPE;
    for (int i = 0; i < 4; i++) dims[i] = st.dims[i];
    if (st.overflow) return SIB_ECAP;
    return st.count;
}

int float_to_string(float x, char* buf) {
    if (std::isnan(x)) { std::memcpy(buf, "\"NaN\"", 5); return 5; }
    if (std::isinf(x)) {
        if (x > 0) { std::memcpy(buf, "\"Infinity\"", 10); return 10; }
        std::memcpy(buf, "\"-Infinity\"", 11);
        return 11;
    }
    char* o = buf;
    if (std::signbit(x)) { *o++ = '-'; x = -x; }
    if (x == 0.0f) { std::memcpy(o, "0.0", 3); return (int)(o + 3 - buf); }
    char t[32];
    auto r = std::to_chars(t, t + sizeof t, x, std::chars_format::scientific);  // shortest round-trip
    // t = d[.ddd]e[+-]XX
    char dig[16];
    int nd = 0;
    const char* s = t;
    for (; s < r.ptr && *s != 'e'; s++)
        if (*s != '.') dig[nd++] = *s;
    if (nd == 1) {
        // Java shows at least two digits and picks the 2-digit decimal closest to x (matters for e.g. 1.4E-45)
        r = std::to_chars(t, t + sizeof t, x, std::chars_format::scientific, 1);
        nd = 0;
        for (s = t; s < r.ptr && *s != 'e'; s++)
            if (*s != '.') dig[nd++] = *s;
        if (nd == 2 && dig[1] == '0') nd = 1;
    }
    int e10 = 0;
    {
        const char* e = s + 1;
        bool eneg = false;
        if (*e == '+') e++;
        else if (*e == '-') { eneg = true; e++; }
        for (; e < r.ptr; e++) e10 = e10 * 10 + (*e - '0');
        if (eneg) e10 = -e10;
    }
    if (x >= 1e-3f && x < 1e7f) {
        if (e10 >= 0) {
            for (int i = 0; i <= e10; i++) *o++ = i < nd ? dig[i] : '0';
            *o++ = '.';
            if (nd > e10 + 1) for (int i = e10 + 1; i < nd; i++) *o++ = dig[i];
            else *o++ = '0';
        } else {
            *o++ = '0';
            *o++ = '.';
            for (int i = -1; i > e10; i--) *o++ = '0';
            for (int i = 0; i < nd; i++) *o++ = dig[i];
        }
    } else {
        *o++ = dig[0];
        *o++ = '.';
        if (nd > 1) for (int i = 1; i < nd; i++) *o++ = dig[i];
        else *o++ = '0';
        *o++ = 'E';
        if (e10 < 0) { *o++ = '-'; e10 = -e10; }
        char eb[8];
        int ne = 0;
        do { eb[ne++] = (char)('0' + e10 % 10); e10 /= 10; } while (e10);
        while (ne) *o++ = eb[--ne];
    }
    return (int)(o - buf);
}

long json_encode(const float* probs, long n, int ncls, char* buf, size_t cap) {
    size_t need = 32 + (size_t)n * ((size_t)ncls * 18 + 4);
    if (cap < need) return SIB_ECAP;
    char* o = buf;
    static const char head[] = "{\"predictions\":[";
    std::memcpy(o, head, sizeof head - 1);
    o += sizeof head - 1;
    for (long i = 0; i < n; i++) {
        if (i) *o++ = ',';
        *o++ = '[';
        for (int k = 0; k < ncls; k++) {
            if (k) *o++ = ',';
            o += float_to_string(probs[i * ncls + k], o);
        }
        *o++ = ']';
    }
    *o++ = ']';
    *o++ = '}';
    *o = 0;
    return (long)(o - buf);
}

}  // namespace sib
