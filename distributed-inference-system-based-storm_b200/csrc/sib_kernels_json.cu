// sib_kernels_json.cu -- GPU-side decode of {"instances":[[[[...]]]]} tuples.
//
// The reference decodes each tuple on the executor thread with Jackson (InferenceBolt.java:76-77) and copies the
// nested arrays into a tensor (:80); on the host that costs more than the model (tens of microseconds per 37 KB
// CIFAR tuple per core). Here the raw UTF-8 of a whole micro-batch crosses PCIe once and one CTA per tuple turns it
// into fp32 NHWC directly in the batch's device buffer.
//
// Scope of the fast path: ONE instance per tuple (what the reference's copyTo(new float[1][10]) admits,
// InferenceBolt.java:86), in canonical JSON nesting for the model's HxWxC, optional whitespace between tokens,
// numbers that the exact fast conversion covers (<= 15 significant digits, |decimal exponent| <= 22 -- every
// shortest-form fp32 literal). Anything else -- other keys, nulls, N != 1, ragged rows, syntax errors, long
// mantissas -- only sets the tuple's flag; the engine then re-reads that tuple with the host decoder
// (csrc/sib_json.cpp), which owns the full Jackson semantics and error classes. So results never differ from the
// host path, the GPU merely handles the common case.
//
// Number semantics are the host fast path's, operation for operation: text -> uint64 significand and decimal
// exponent -> double (one exact multiply or divide by a power of ten: correctly rounded) -> (float). This is
// Jackson's text -> double -> float, double rounding included.
#include <cuda_runtime.h>

#include "sib_internal.hpp"

namespace sib {
namespace {

// Threads per tuple (one CTA per tuple). 256 gives the highest rate when a batch has many tuples (1.30 M tuples/s);
// 1024 gets a lone tuple through soonest (round trip 81 us against 116 us; 6% lower rate): the launcher picks by the
// number of tuples in the batch.
constexpr int kJTWide = 1024, kJTNarrow = 256;
constexpr int kMaxText = 56 * 1024;  // bytes of one tuple staged in shared memory; longer tuples go to the host
constexpr int kMaxElems = 4096;      // token positions kept behind the text (2 bytes each)

__constant__ double c_pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

__device__ __forceinline__ bool is_ws(unsigned char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r'; }
__device__ __forceinline__ bool is_digit(unsigned char c) { return (unsigned)(c - '0') <= 9u; }
__device__ __forceinline__ bool is_num(unsigned char c) {
    return is_digit(c) || c == '-' || c == '+' || c == '.' || c == 'e' || c == 'E';
}
// A number token starts at a digit or '-' that does not continue a number ('e' alone is not a start: the key
// "instances" contains one).
__device__ __forceinline__ bool is_start(const unsigned char* s, int i) {
    return (is_digit(s[i]) || s[i] == '-') && !(i > 0 && is_num(s[i - 1]));
}

// walks backwards over optional whitespace and expects `ch`; returns false if something else is there
__device__ __forceinline__ bool back_expect(const unsigned char* s, int& q, unsigned char ch) {
    while (q >= 0 && is_ws(s[q])) q--;
    if (q < 0 || s[q] != ch) return false;
    q--;
    return true;
}
__device__ __forceinline__ bool fwd_expect(const unsigned char* s, int len, int& q, unsigned char ch) {
    while (q < len && is_ws(s[q])) q++;
    if (q >= len || s[q] != ch) return false;
    q++;
    return true;
}

// desc[4*t + {0,1,2}] = text offset, text length, destination image index of tuple t
template <int kJT>
__global__ void __launch_bounds__(kJT) json_decode_kernel(const unsigned char* __restrict__ text,
                                                         const int32_t* __restrict__ desc, float* __restrict__ out,
                                                         int32_t* __restrict__ flags, int H, int W, int Cc) {
    extern __shared__ __align__(16) unsigned char s[];
    __shared__ int s_cnt[kJT];
    __shared__ int s_bad;
    const int t = blockIdx.x, tid = threadIdx.x;
    const int off = desc[4 * t], len = desc[4 * t + 1], img = desc[4 * t + 2];
    const int elems = H * W * Cc;
    if (tid == 0) s_bad = 0;
    if (len > kMaxText || len < 16 || elems > kMaxElems) {
        if (tid == 0) flags[t] = 1;
        return;
    }
    // ---- stage the tuple's bytes (text offsets are 16-byte aligned by the engine)
    {
        const uint4* src = reinterpret_cast<const uint4*>(text + off);
        uint4* dst = reinterpret_cast<uint4*>(s);
        for (int i = tid; i < (len + 15) / 16; i += kJT) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    // ---- pass A: number tokens starting in this thread's chunk
    const int chunk = (len + kJT - 1) / kJT;
    const int lo = tid * chunk, hi = min(len, lo + chunk);
    int cnt = 0;
    for (int i = lo; i < hi; i++) {
        cnt += is_start(s, i) ? 1 : 0;
    }
    s_cnt[tid] = cnt;
    __syncthreads();
    // exclusive scan (kJT entries: one warp-level pass per 32, then the kJT/32 partials)
    {
        const int lane = tid & 31, wrp = tid >> 5;
        int v = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        __shared__ int s_w[kJT / 32];
        if (lane == 31) s_w[wrp] = v;
        __syncthreads();
        int base = 0;
        for (int w = 0; w < wrp; w++) base += s_w[w];
        s_cnt[tid] = base + v - cnt;  // exclusive prefix
        __syncthreads();
        if (tid == kJT - 1 && base + v != elems) s_bad = 1;  // N != 1, ragged, or not this model's shape
    }
    __syncthreads();
    if (s_bad) {
        if (tid == 0) flags[t] = 1;
        return;
    }
    // ---- pass B: record where every token starts (thread-sequential over its chunk, but cheap) ...
    unsigned short* pos = reinterpret_cast<unsigned short*>(s + ((len + 15) & ~15));  // elems x 2 bytes behind the text
    unsigned short* endp = pos + kMaxElems;                                           // ... and where every token ends
    {
        int k = s_cnt[tid];
        for (int i = lo; i < hi; i++)
            if (is_start(s, i)) pos[k++] = (unsigned short)i;
    }
    __syncthreads();
    // ---- ... pass C: one token per lane at a time, so a warp parses 32 numbers in lock-step instead of taking the
    // parse body once per byte position (the byte-driven form ran 10x longer: every position starts a token in
    // SOME lane). Parse, check the punctuation in front of it, store.
    bool bad = false;
    float* dst = out + (size_t)img * elems;
    for (int k = tid; k < elems; k += kJT) {
        const int i = pos[k];
        // -- number, RFC 8259 grammar
        int p = i;
        bool neg = false;
        if (s[p] == '-') { neg = true; p++; }
        unsigned long long mant = 0;
        int nd = 0, e10 = 0;
        if (p < len && s[p] == '0') {
            p++;
            if (p < len && is_digit(s[p])) bad = true;  // leading zero
        } else if (p < len && is_digit(s[p])) {
            while (p < len && is_digit(s[p])) { mant = mant * 10ull + (unsigned)(s[p] - '0'); nd++; p++; if (nd > 18) break; }
        } else {
            bad = true;
        }
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
        if (p < len && is_num(s[p])) bad = true;                                  // "1.2.3", "1e5e", "1-2"
        if (nd > 15 || mant > (1ull << 53) || e10 < -22 || e10 > 22) bad = true;  // outside the exact fast path
        double d = (double)mant;
        const int ae = e10 < 0 ? -e10 : e10;
        const double pw = c_pow10[ae > 22 ? 22 : ae];
        d = e10 < 0 ? d / pw : d * pw;
        if (k < elems) dst[k] = (float)(neg ? -d : d);
        endp[k] = (unsigned short)p;
    }
    __syncthreads();
    // ---- pass D: the punctuation between the previous token's END and this token must be exactly what the canonical
    // nesting puts there. (Walking back only as far as "some digit" let a stray run that holds no token start but
    // ends in a digit -- "7,e5,8", "7,.5,8", "7,+5,8" -- pass; the host decoder and Jackson reject those.)
    for (int k = tid; k < elems; k += kJT) {
        const int i = pos[k], p = endp[k];
        int q = i - 1;
        if (k == 0) {
            bool ok = back_expect(s, q, '[') && back_expect(s, q, '[') && back_expect(s, q, '[') && back_expect(s, q, '[') &&
                      back_expect(s, q, ':') && back_expect(s, q, '"');
            const char key[] = "instances";
            for (int c = 8; ok && c >= 0; c--) {  // the key itself admits no whitespace
                if (q < 0 || s[q] != (unsigned char)key[c]) ok = false;
                q--;
            }
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
            ok = ok && q + 1 == (int)endp[k - 1];  // ... and nothing else since the previous number ended
            bad |= !ok;
        }
        if (k == elems - 1) {  // closing: ]]]] }  (bytes after the root object are ignored, like Jackson)
            int f = p;
            bool ok = fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, ']') &&
                      fwd_expect(s, len, f, ']') && fwd_expect(s, len, f, '}');
            bad |= !ok;
        }
    }
    if (bad) s_bad = 1;
    __syncthreads();
    if (tid == 0) flags[t] = s_bad;
}

}  // namespace

size_t json_gpu_max_text() { return kMaxText; }

int launch_json_decode(const char* d_text, const int32_t* d_desc, int n_tuples, float* d_out, int32_t* d_flags, int H,
                       int W, int C, void* stream) {
    if (n_tuples <= 0) return 0;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    constexpr int smem = kMaxText + 16 + 4 * kMaxElems;
    const unsigned char* text = reinterpret_cast<const unsigned char*>(d_text);
    if (n_tuples <= sms) {  // fewer tuples than SMs: latency matters, every tuple has an SM to itself anyway
        static SmemLimit limit;
        if (int rc = limit.ensure(dev, smem, [](int b) {
                return (int)cudaFuncSetAttribute(json_decode_kernel<kJTWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
            }))
            return rc;
        json_decode_kernel<kJTWide><<<n_tuples, kJTWide, smem, (cudaStream_t)stream>>>(text, d_desc, d_out, d_flags, H, W, C);
    } else {
        static SmemLimit limit;
        if (int rc = limit.ensure(dev, smem, [](int b) {
                return (int)cudaFuncSetAttribute(json_decode_kernel<kJTNarrow>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
            }))
            return rc;
        json_decode_kernel<kJTNarrow><<<n_tuples, kJTNarrow, smem, (cudaStream_t)stream>>>(text, d_desc, d_out, d_flags, H, W, C);
    }
    return (int)cudaGetLastError();
}

}  // namespace sib
