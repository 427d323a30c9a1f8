/*
 * sib_oracle.c -- CPU ORACLE for the InferenceBolt hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the checker, not the product: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call it.  Nothing under
 * distributed-inference-system-based-storm_b200/ links or imports it.
 *
 * PARITY UNPINNED: the reference ships no tests, no golden vectors, no sample payload and no model file
 * (SURVEY.md F3/F4), and its arithmetic lives in un-vendored third-party binaries
 * (org.tensorflow:tensorflow:1.13.1, pom.xml:22-26; jackson-databind 2.10.0.pr1, pom.xml:94-98) that cannot
 * run here (no JVM).  This file therefore restates the *published* semantics of those dependencies at the
 * reference's call sites and is cross-checked against an independent torch-CPU implementation
 * (oracle/torch_ref.py) and hand-derived known-answer vectors (tests/test_oracle_kat.py).
 *
 * What is restated, and where the reference does it:
 *   sibo_execute()        InferenceBolt.execute()            InferenceBolt.java:69-100
 *   sibo_json_decode()    objectMapper.readValue(.., InstObj) InferenceBolt.java:76-77, InstObj.java:7-16
 *                         (Jackson: number text -> double -> (float) narrowing)
 *   sibo_forward_f32()    Tensor.create + Session.run        InferenceBolt.java:80-86
 *                         (TF 1.x Conv2D NHWC/HWIO VALID s1, BiasAdd, Relu, MaxPool 2x2/2 VALID,
 *                          flatten (h,w,c), MatMul [N,K]x[K,M], BiasAdd, Softmax = exp(x-max)/sum in fp32)
 *   sibo_json_encode()    writeValueAsString(predObj)        InferenceBolt.java:89-91, PredObj.java:8-17
 *                         (floats printed per java.lang.Float.toString, JDK>=19 shortest-decimal spec)
 *
 * Plain C99 + optional OpenMP (batch drivers only).  No dependency on the product.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define SIBO_MAX_LAYERS 16
#define SIBO_KIND_CONV 1u  /* conv(kh,kw) VALID s1 + bias + ReLU + maxpool 2x2/2 */
#define SIBO_KIND_DENSE 2u /* x.W + b (+ReLU if flags&1) */

typedef struct {
    uint32_t kind, a, b, c, d, flags, w_off, b_off; /* conv: a=kh b=kw c=cin d=cout ; dense: a=in b=out */
} sibo_layer;

typedef struct {
    uint32_t in_h, in_w, in_c, n_layers, n_classes;
    sibo_layer layer[SIBO_MAX_LAYERS];
    const float *weights; /* points into the blob */
    uint32_t n_weights;
    uint32_t max_act; /* largest activation element count per image */
} sibo_model;

/* ------------------------------------------------------------------ synthetic data */

static uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t sibo_mix64(uint64_t z) { return mix64(z); }

/* counter-based U[0,1): 24 random bits * 2^-24, bit-identical in C / numpy / CUDA */
float sibo_u01(uint64_t seed, uint64_t index) {
    uint64_t r = mix64(mix64(seed) + index);
    return (float)(r >> 40) * (1.0f / 16777216.0f);
}

void sibo_fill_u01(uint64_t seed, uint64_t start, uint64_t n, float *out) {
    for (uint64_t i = 0; i < n; i++) out[i] = sibo_u01(seed, start + i);
}

/* ------------------------------------------------------------------ model blob */
/*
 * Layout (little endian):  "SIBM" | u32 version=1 | u32 in_h,in_w,in_c | u32 n_layers | u32 n_weights | u32 rsvd
 *                          n_layers x { u32 kind,a,b,c,d,flags,w_off,b_off }   (offsets in floats)
 *                          n_weights x f32
 * Stands where the reference keeps model/saved_model.pb (InferenceBolt.java:49).
 */
#define SIBO_HDR_WORDS 8

int sibo_model_parse(const void *blob, size_t n, sibo_model *m) {
    const uint32_t *w = (const uint32_t *)blob;
    if (n < SIBO_HDR_WORDS * 4 || memcmp(blob, "SIBM", 4) != 0 || w[1] != 1) return -1;
    memset(m, 0, sizeof *m);
    m->in_h = w[2]; m->in_w = w[3]; m->in_c = w[4]; m->n_layers = w[5]; m->n_weights = w[6];
    if (m->n_layers == 0 || m->n_layers > SIBO_MAX_LAYERS) return -2;
    size_t need = (size_t)SIBO_HDR_WORDS * 4 + (size_t)m->n_layers * 32 + (size_t)m->n_weights * 4;
    if (n < need) return -3;
    memcpy(m->layer, w + SIBO_HDR_WORDS, (size_t)m->n_layers * 32);
    m->weights = (const float *)(w + SIBO_HDR_WORDS + m->n_layers * 8);
    uint32_t h = m->in_h, wd = m->in_w, c = m->in_c, flat = 0, feat = h * wd * c;
    m->max_act = feat;
    for (uint32_t i = 0; i < m->n_layers; i++) {
        const sibo_layer *L = &m->layer[i];
        if (L->kind == SIBO_KIND_CONV) {
            if (flat || L->c != c || h < L->a || wd < L->b) return -4;
            uint32_t oh = h - L->a + 1, ow = wd - L->b + 1;
            if (oh * ow * L->d > m->max_act) m->max_act = oh * ow * L->d;
            h = oh / 2; wd = ow / 2; c = L->d; feat = h * wd * c;
            if ((size_t)L->w_off + (size_t)L->a * L->b * L->c * L->d > m->n_weights) return -5;
            if ((size_t)L->b_off + L->d > m->n_weights) return -5;
        } else if (L->kind == SIBO_KIND_DENSE) {
            flat = 1;
            if (L->a != feat) return -4;
            feat = L->b;
            if (feat > m->max_act) m->max_act = feat;
            if ((size_t)L->w_off + (size_t)L->a * L->b > m->n_weights) return -5;
            if ((size_t)L->b_off + L->b > m->n_weights) return -5;
        } else return -6;
    }
    if (!flat) return -7;
    m->n_classes = feat;
    return 0;
}

/*
 * Presets: id 0 mnist-lenet5, 1 cifar-lenet5, 2 mnist-wide, 3 cifar-wide.
 * Weights per D2: conv/dense ~ U(-a,a), a = sqrt(6/fan_in); biases ~ U(-0.1,0.1); last layer weights x4.
 * Weight seed 0xC0FFEE00 + preset_id; element index = position in the weight section.
 * Returns blob size (call with out==NULL to size it), or <0.
 */
long sibo_make_preset(int preset, void *out, size_t cap) {
    static const uint32_t cfg[4][8] = {
        /* H  W  C  c1 c2  then dense dims (0 terminated) */
        {28, 28, 1, 6, 16, 120, 84, 10}, {32, 32, 3, 6, 16, 120, 84, 10},
        {28, 28, 1, 16, 32, 64, 10, 0},  {32, 32, 3, 16, 32, 64, 10, 0}};
    if (preset < 0 || preset > 3) return -1;
    const uint32_t *p = cfg[preset];
    sibo_layer L[SIBO_MAX_LAYERS];
    uint32_t nl = 0, off = 0, h = p[0], w = p[1], c = p[2];
    for (int i = 0; i < 2; i++) {
        sibo_layer *l = &L[nl++];
        memset(l, 0, sizeof *l);
        l->kind = SIBO_KIND_CONV; l->a = 5; l->b = 5; l->c = c; l->d = p[3 + i]; l->flags = 1;
        l->w_off = off; off += 25 * c * l->d; l->b_off = off; off += l->d;
        h = (h - 4) / 2; w = (w - 4) / 2; c = l->d;
    }
    uint32_t feat = h * w * c;
    for (int i = 5; i < 8 && p[i]; i++) {
        sibo_layer *l = &L[nl++];
        memset(l, 0, sizeof *l);
        l->kind = SIBO_KIND_DENSE; l->a = feat; l->b = p[i];
        l->flags = (i + 1 < 8 && p[i + 1]) ? 1u : 0u;
        l->w_off = off; off += feat * p[i]; l->b_off = off; off += p[i];
        feat = p[i];
    }
    size_t need = (size_t)SIBO_HDR_WORDS * 4 + (size_t)nl * 32 + (size_t)off * 4;
    if (!out) return (long)need;
    if (cap < need) return -2;
    uint32_t *hd = (uint32_t *)out;
    memcpy(hd, "SIBM", 4);
    hd[1] = 1; hd[2] = p[0]; hd[3] = p[1]; hd[4] = p[2]; hd[5] = nl; hd[6] = off; hd[7] = 0;
    memcpy(hd + SIBO_HDR_WORDS, L, (size_t)nl * 32);
    float *wt = (float *)(hd + SIBO_HDR_WORDS + nl * 8);
    uint64_t seed = 0xC0FFEE00ull + (uint64_t)preset;
    for (uint32_t i = 0; i < nl; i++) {
        uint32_t fan_in = L[i].kind == SIBO_KIND_CONV ? L[i].a * L[i].b * L[i].c : L[i].a;
        uint32_t nw = L[i].kind == SIBO_KIND_CONV ? fan_in * L[i].d : L[i].a * L[i].b;
        uint32_t nb = L[i].kind == SIBO_KIND_CONV ? L[i].d : L[i].b;
        float a = (float)sqrt(6.0 / (double)fan_in);
        if (i + 1 == nl) a *= 4.0f;
        for (uint32_t k = 0; k < nw; k++)
            wt[L[i].w_off + k] = a * (2.0f * sibo_u01(seed, L[i].w_off + k) - 1.0f);
        for (uint32_t k = 0; k < nb; k++)
            wt[L[i].b_off + k] = 0.1f * (2.0f * sibo_u01(seed, L[i].b_off + k) - 1.0f);
    }
    return (long)need;
}

/* ------------------------------------------------------------------ ops, fp32 (TF 1.x semantics) */

/* Conv2D NHWC x HWIO, VALID, stride 1 (cross-correlation, no filter flip). One image.
 * Every output is the sum over (kh, kw, ci) in that order, one fp32 multiply and one fp32 add per term; the output
 * channel is the innermost loop only so that the compiler can vectorise over it (it does not change any sum). */
void sibo_conv2d_valid(const float *in, int h, int w, int cin, const float *filt, int kh, int kw, int cout,
                       float *out) {
    int oh = h - kh + 1, ow = w - kw + 1;
    for (int y = 0; y < oh; y++)
        for (int x = 0; x < ow; x++) {
            float *acc = out + (y * ow + x) * cout;
            for (int co = 0; co < cout; co++) acc[co] = 0.0f;
            for (int i = 0; i < kh; i++)
                for (int j = 0; j < kw; j++)
                    for (int ci = 0; ci < cin; ci++) {
                        const float v = in[((y + i) * w + (x + j)) * cin + ci];
                        const float *wp = filt + ((i * kw + j) * cin + ci) * cout;
                        for (int co = 0; co < cout; co++) acc[co] += v * wp[co];
                    }
        }
}

void sibo_bias_add(float *x, int n_pos, int c, const float *bias) {
    for (int p = 0; p < n_pos; p++)
        for (int k = 0; k < c; k++) x[p * c + k] += bias[k];
}

void sibo_relu(float *x, int n) {
    for (int i = 0; i < n; i++) x[i] = x[i] > 0.0f ? x[i] : 0.0f;
}

/* MaxPool ksize 2, stride 2, VALID: trailing odd row/column is dropped. */
void sibo_maxpool2(const float *in, int h, int w, int c, float *out) {
    int oh = h / 2, ow = w / 2;
    for (int y = 0; y < oh; y++)
        for (int x = 0; x < ow; x++)
            for (int k = 0; k < c; k++) {
                float a = in[((2 * y) * w + 2 * x) * c + k], b = in[((2 * y) * w + 2 * x + 1) * c + k];
                float d = in[((2 * y + 1) * w + 2 * x) * c + k], e = in[((2 * y + 1) * w + 2 * x + 1) * c + k];
                float m = a > b ? a : b, n = d > e ? d : e;
                out[(y * ow + x) * c + k] = m > n ? m : n;
            }
}

/* MatMul [1,K] x [K,M] with W stored [in,out]: out[j] = sum over i in order (inner loop over j vectorises) */
void sibo_dense(const float *x, int k, const float *wt, int m, float *out) {
    for (int j = 0; j < m; j++) out[j] = 0.0f;
    for (int i = 0; i < k; i++) {
        const float v = x[i];
        const float *wp = wt + (size_t)i * m;
        for (int j = 0; j < m; j++) out[j] += v * wp[j];
    }
}

void sibo_softmax(const float *x, int n, float *out) {
    float mx = x[0];
    for (int i = 1; i < n; i++) mx = x[i] > mx ? x[i] : mx;
    float s = 0.0f;
    for (int i = 0; i < n; i++) { out[i] = expf(x[i] - mx); s += out[i]; }
    for (int i = 0; i < n; i++) out[i] = out[i] / s;
}

/* One image through the whole graph. scratch: 3*max_act floats (pre-pool temp + ping-pong). */
static void forward_one_f32(const sibo_model *m, const float *img, float *logits, float *probs, float *scratch) {
    float *tmp = scratch, *nxt = scratch + m->max_act, *oth = scratch + 2 * (size_t)m->max_act;
    int h = (int)m->in_h, w = (int)m->in_w, c = (int)m->in_c;
    const float *cur = img;
    for (uint32_t i = 0; i < m->n_layers; i++) {
        const sibo_layer *L = &m->layer[i];
        if (L->kind == SIBO_KIND_CONV) {
            int oh = h - (int)L->a + 1, ow = w - (int)L->b + 1;
            sibo_conv2d_valid(cur, h, w, c, m->weights + L->w_off, (int)L->a, (int)L->b, (int)L->d, tmp);
            sibo_bias_add(tmp, oh * ow, (int)L->d, m->weights + L->b_off);
            sibo_relu(tmp, oh * ow * (int)L->d);
            sibo_maxpool2(tmp, oh, ow, (int)L->d, nxt);
            h = oh / 2; w = ow / 2; c = (int)L->d;
        } else {
            sibo_dense(cur, (int)L->a, m->weights + L->w_off, (int)L->b, nxt);
            sibo_bias_add(nxt, 1, (int)L->b, m->weights + L->b_off);
            if (L->flags & 1u) sibo_relu(nxt, (int)L->b);
        }
        cur = nxt;
        { float *t = nxt; nxt = oth; oth = t; }
    }
    for (uint32_t k = 0; k < m->n_classes; k++) logits[k] = cur[k];
    if (probs) sibo_softmax(cur, (int)m->n_classes, probs);
}

/* ---- tiny pthread parallel-for (batch drivers only; no OpenMP: libgomp specs are missing from one of the image's gcc installs) */
typedef void (*sibo_range_fn)(void *ctx, long lo, long hi, int tid);
typedef struct { sibo_range_fn fn; void *ctx; long lo, hi; int tid; } sibo_job;
static void *job_main(void *p) { sibo_job *j = (sibo_job *)p; j->fn(j->ctx, j->lo, j->hi, j->tid); return NULL; }

int sibo_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n < 1 ? 1 : (int)n;
}

static void parallel_for(long n, int threads, sibo_range_fn fn, void *ctx) {
    if (threads < 1) threads = sibo_max_threads();
    if (threads > 256) threads = 256;
    if ((long)threads > n) threads = n > 0 ? (int)n : 1;
    if (threads == 1) { fn(ctx, 0, n, 0); return; }
    pthread_t th[256];
    sibo_job jb[256];
    for (int t = 0; t < threads; t++) {
        jb[t].fn = fn; jb[t].ctx = ctx; jb[t].tid = t;
        jb[t].lo = n * t / threads; jb[t].hi = n * (t + 1) / threads;
        if (t) pthread_create(&th[t], NULL, job_main, &jb[t]);
    }
    job_main(&jb[0]);
    for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
}

typedef struct { const sibo_model *m; const float *x; float *logits, *probs; int fail; } fwd_ctx;

static void fwd_range(void *vp, long lo, long hi, int tid) {
    fwd_ctx *c = (fwd_ctx *)vp;
    (void)tid;
    const sibo_model *m = c->m;
    size_t img = (size_t)m->in_h * m->in_w * m->in_c;
    int nc = (int)m->n_classes;
    float *scratch = (float *)malloc(sizeof(float) * 3 * m->max_act);
    float lg[64], pr[64];
    if (!scratch) { c->fail = 1; return; }
    for (long i = lo; i < hi; i++) {
        /* one image per call: the reference's batch-1 Session.run (InferenceBolt.java:81-85) */
        forward_one_f32(m, c->x + (size_t)i * img, lg, pr, scratch);
        if (c->logits) memcpy(c->logits + (size_t)i * nc, lg, sizeof(float) * nc);
        if (c->probs) memcpy(c->probs + (size_t)i * nc, pr, sizeof(float) * nc);
    }
    free(scratch);
}

/* fp32 forward over n images (NHWC contiguous). logits/probs: n x n_classes (either may be NULL).
 * threads: 1 = scalar port; k = k "executors" each owning a contiguous share; <1 = all cores. */
int sibo_forward_f32(const sibo_model *m, const float *nhwc, long n, float *logits, float *probs, int threads) {
    if (m->n_classes > 64) return -2;
    fwd_ctx c = {m, nhwc, logits, probs, 0};
    parallel_for(n, threads, fwd_range, &c);
    return c.fail ? -1 : 0;
}

/* fp64 shadow: same graph, double accumulation and activations. Classifies genuine argmax ties. */
int sibo_forward_f64(const sibo_model *m, const float *nhwc, long n, double *logits, double *probs) {
    size_t img = (size_t)m->in_h * m->in_w * m->in_c;
    int nc = (int)m->n_classes;
    double *a = (double *)malloc(sizeof(double) * 3 * m->max_act);
    if (!a) return -1;
    double *b = a + m->max_act, *p = b + m->max_act;
    for (long n_i = 0; n_i < n; n_i++) {
        int h = (int)m->in_h, w = (int)m->in_w, c = (int)m->in_c;
        for (size_t i = 0; i < img; i++) a[i] = nhwc[(size_t)n_i * img + i];
        double *cur = a, *nxt = b;
        for (uint32_t li = 0; li < m->n_layers; li++) {
            const sibo_layer *L = &m->layer[li];
            const float *wt = m->weights + L->w_off, *bs = m->weights + L->b_off;
            if (L->kind == SIBO_KIND_CONV) {
                int kh = (int)L->a, kw = (int)L->b, co_n = (int)L->d, oh = h - kh + 1, ow = w - kw + 1;
                for (int y = 0; y < oh; y++)
                    for (int x = 0; x < ow; x++)
                        for (int co = 0; co < co_n; co++) {
                            double acc = 0.0;
                            for (int i = 0; i < kh; i++)
                                for (int j = 0; j < kw; j++)
                                    for (int ci = 0; ci < c; ci++)
                                        acc += cur[((y + i) * w + (x + j)) * c + ci] *
                                               (double)wt[((i * kw + j) * c + ci) * co_n + co];
                            acc += (double)bs[co];
                            p[(y * ow + x) * co_n + co] = acc > 0.0 ? acc : 0.0;
                        }
                int ph = oh / 2, pw = ow / 2;
                for (int y = 0; y < ph; y++)
                    for (int x = 0; x < pw; x++)
                        for (int k = 0; k < co_n; k++) {
                            double v0 = p[((2 * y) * ow + 2 * x) * co_n + k], v1 = p[((2 * y) * ow + 2 * x + 1) * co_n + k];
                            double v2 = p[((2 * y + 1) * ow + 2 * x) * co_n + k], v3 = p[((2 * y + 1) * ow + 2 * x + 1) * co_n + k];
                            double m0 = v0 > v1 ? v0 : v1, m1 = v2 > v3 ? v2 : v3;
                            nxt[(y * pw + x) * co_n + k] = m0 > m1 ? m0 : m1;
                        }
                h = ph; w = pw; c = co_n;
            } else {
                int k = (int)L->a, mo = (int)L->b;
                for (int j = 0; j < mo; j++) {
                    double acc = 0.0;
                    for (int i = 0; i < k; i++) acc += cur[i] * (double)wt[i * mo + j];
                    acc += (double)bs[j];
                    nxt[j] = (L->flags & 1u) ? (acc > 0.0 ? acc : 0.0) : acc;
                }
            }
            double *t = cur; cur = nxt; nxt = t;
        }
        double mx = cur[0], s = 0.0;
        for (int k = 1; k < nc; k++) mx = cur[k] > mx ? cur[k] : mx;
        for (int k = 0; k < nc; k++) s += exp(cur[k] - mx);
        for (int k = 0; k < nc; k++) {
            if (logits) logits[(size_t)n_i * nc + k] = cur[k];
            if (probs) probs[(size_t)n_i * nc + k] = exp(cur[k] - mx) / s;
        }
    }
    free(a);
    return 0;
}

/* ------------------------------------------------------------------ JSON codec (Jackson / Float.toString semantics) */

#define SIBO_EJSON (-1)  /* malformed JSON / unknown property  -> reference catches, emits null (InferenceBolt.java:92-98) */
#define SIBO_ESHAPE (-2) /* ragged / wrong rank / null instances -> reference throws uncaught (IllegalArgument/NPE)      */
#define SIBO_ECAP (-3)   /* caller buffer too small */

typedef struct { const char *p, *end; } cur_t;

static void skip_ws(cur_t *c) {
    while (c->p < c->end && (*c->p == ' ' || *c->p == '\t' || *c->p == '\n' || *c->p == '\r')) c->p++;
}

/* JSON number grammar (RFC 8259); value = strtod(text) narrowed to float, as Jackson's
 * ParserBase.getFloatValue() does: (float) getDoubleValue(). */
static int parse_number(cur_t *c, float *out) {
    const char *s = c->p, *q = s;
    if (q < c->end && *q == '-') q++;
    if (q >= c->end) return SIBO_EJSON;
    if (*q == '0') q++;
    else if (*q >= '1' && *q <= '9') { while (q < c->end && *q >= '0' && *q <= '9') q++; }
    else return SIBO_EJSON;
    if (q < c->end && *q == '.') {
        q++;
        if (q >= c->end || *q < '0' || *q > '9') return SIBO_EJSON;
        while (q < c->end && *q >= '0' && *q <= '9') q++;
    }
    if (q < c->end && (*q == 'e' || *q == 'E')) {
        q++;
        if (q < c->end && (*q == '+' || *q == '-')) q++;
        if (q >= c->end || *q < '0' || *q > '9') return SIBO_EJSON;
        while (q < c->end && *q >= '0' && *q <= '9') q++;
    }
    char tmp[512];
    size_t len = (size_t)(q - s);
    char *buf = len < sizeof tmp ? tmp : (char *)malloc(len + 1);
    if (!buf) return SIBO_ECAP;
    memcpy(buf, s, len);
    buf[len] = 0;
    double d = strtod(buf, NULL);
    if (buf != tmp) free(buf);
    *out = (float)d;
    c->p = q;
    return 0;
}

/* parses [[[[f,..]]]] at depth 0..3. Jackson finishes parsing before Tensor.create sees the arrays, so syntax /
 * type errors (SIBO_EJSON) win over shape errors; ragged rows and null rows only set *ragged. */
static int parse_array(cur_t *c, int depth, long dims[4], int first[4], int *ragged, float *out, long cap, long *count) {
    skip_ws(c);
    if (c->p >= c->end) return SIBO_EJSON;
    if (*c->p != '[') {
        if (c->end - c->p >= 4 && memcmp(c->p, "null", 4) == 0) { c->p += 4; *ragged = 1; return 0; }
        return SIBO_EJSON; /* number/string/object where an array is required: MismatchedInputException */
    }
    c->p++;
    long n = 0;
    skip_ws(c);
    if (c->p < c->end && *c->p == ']') { c->p++; }
    else {
        for (;;) {
            if (depth == 3) {
                skip_ws(c);
                float v;
                if (c->end - c->p >= 4 && memcmp(c->p, "null", 4) == 0) { c->p += 4; v = 0.0f; }
                else {
                    int r = parse_number(c, &v); /* '[' here (rank > 4) fails the number grammar -> EJSON */
                    if (r) return r;
                }
                if (*count < cap) out[*count] = v;
                (*count)++;
            } else {
                int r = parse_array(c, depth + 1, dims, first, ragged, out, cap, count);
                if (r) return r;
            }
            n++;
            skip_ws(c);
            if (c->p >= c->end) return SIBO_EJSON;
            if (*c->p == ',') { c->p++; continue; }
            if (*c->p == ']') { c->p++; break; }
            return SIBO_EJSON;
        }
    }
    if (first[depth]) { dims[depth] = n; first[depth] = 0; }
    else if (dims[depth] != n) *ragged = 1;
    return 0;
}

/*
 * {"instances": [[[[..]]]]}  ->  out[N*H*W*C] fp32 NHWC, dims = {N,H,W,C}.
 * Returns number of floats, or SIBO_EJSON / SIBO_ESHAPE / SIBO_ECAP.
 * Unknown property => SIBO_EJSON (Jackson FAIL_ON_UNKNOWN_PROPERTIES default true -> JsonProcessingException).
 * Bytes after the root object are ignored (FAIL_ON_TRAILING_TOKENS default false).
 */
long sibo_json_decode(const char *text, size_t len, float *out, long cap, long dims[4]) {
    cur_t c = {text, text + len};
    int seen = 0, ragged = 0;
    int first[4] = {1, 1, 1, 1};
    long count = 0;
    skip_ws(&c);
    if (c.p >= c.end || *c.p != '{') return SIBO_EJSON;
    c.p++;
    skip_ws(&c);
    if (c.p < c.end && *c.p == '}') { c.p++; }
    else {
        for (;;) {
            skip_ws(&c);
            if (c.p >= c.end || *c.p != '"') return SIBO_EJSON;
            const char *k = ++c.p;
            while (c.p < c.end && *c.p != '"') { if (*c.p == '\\') return SIBO_EJSON; c.p++; }
            if (c.p >= c.end) return SIBO_EJSON;
            size_t kl = (size_t)(c.p - k);
            c.p++;
            if (kl != 9 || memcmp(k, "instances", 9) != 0) return SIBO_EJSON;
            skip_ws(&c);
            if (c.p >= c.end || *c.p != ':') return SIBO_EJSON;
            c.p++;
            first[0] = first[1] = first[2] = first[3] = 1;
            dims[0] = dims[1] = dims[2] = dims[3] = 0;
            count = 0; ragged = 0;
            int r = parse_array(&c, 0, dims, first, &ragged, out, cap, &count);
            if (r) return r;
            seen = 1;
            skip_ws(&c);
            if (c.p >= c.end) return SIBO_EJSON;
            if (*c.p == ',') { c.p++; continue; }
            if (*c.p == '}') { c.p++; break; }
            return SIBO_EJSON;
        }
    }
    if (!seen) return SIBO_ESHAPE; /* instances == null -> NPE in Tensor.create */
    if (ragged || first[3] || count == 0) return SIBO_ESHAPE;
    if (dims[0] * dims[1] * dims[2] * dims[3] != count) return SIBO_ESHAPE;
    if (count > cap) return SIBO_ECAP;
    return count;
}

/*
 * java.lang.Float.toString: shortest decimal that uniquely distinguishes the float (at least 2 digits shown),
 * plain notation for 1e-3 <= |x| < 1e7, otherwise computerised scientific d.dddE[-]n.
 * Jackson writes NaN/Infinity as quoted strings (QUOTE_NON_NUMERIC_NUMBERS default on).
 * Returns length written (no NUL counted); buf must hold >= 24 bytes.
 */
int sibo_float_to_string(float x, char *buf) {
    if (isnan(x)) return sprintf(buf, "\"NaN\"");
    if (isinf(x)) return sprintf(buf, x > 0 ? "\"Infinity\"" : "\"-Infinity\"");
    char *o = buf;
    if (signbit(x)) { *o++ = '-'; x = -x; }
    if (x == 0.0f) { memcpy(o, "0.0", 3); return (int)(o + 3 - buf); }
    /* shortest round-tripping digit string: try 1..9 significant digits; at each length also try the
     * decimal neighbours (asymmetric rounding interval at binade boundaries) and keep the closest. */
    char dig[16];
    int ndig = 0, e10 = 0;
    for (int p = 1; p <= 9 && !ndig; p++) {
        char t[32];
        snprintf(t, sizeof t, "%.*e", p - 1, (double)x);
        /* t = d[.ddd]e[+-]XX */
        long long mant = 0;
        const char *s = t;
        for (; *s && *s != 'e'; s++) if (*s != '.') mant = mant * 10 + (*s - '0');
        int ex = atoi(s + 1);
        double best_err = 1e300;
        long long best_m = -1;
        int best_e = 0;
        for (int dm = -1; dm <= 1; dm++) {
            long long mm = mant + dm;
            int ee = ex;
            if (mm <= 0) continue;
            char u[48];
            snprintf(u, sizeof u, "%llde%d", mm, ee - (p - 1));
            if (strtof(u, NULL) != x) continue;
            double err = fabs(strtod(u, NULL) - (double)x);
            /* Two candidates exactly as close (x = 2^-12 = 2.44140625E-4 lies midway between ..062 and ..063): the even
             * last digit wins (Float.toString's specification since JDK 19; Ryu and std::to_chars agree). The two
             * errors are computed in double, where neither decimal is exact, so "as close" is "equal to 1e-6 of the
             * error": a real difference between two candidates one unit apart is never that small for a 24-bit x. */
            int tie = best_m > 0 && fabs(err - best_err) <= 1e-6 * best_err;
            if ((!tie && err < best_err) || (tie && (mm % 2 == 0))) { best_err = err; best_m = mm; best_e = ee; }
        }
        if (best_m > 0) {
            char u[32];
            int l = snprintf(u, sizeof u, "%lld", best_m);
            /* mant+1 may have gained a digit (99 -> 100) */
            e10 = best_e + (l - p);
            while (l > 1 && u[l - 1] == '0') l--;
            memcpy(dig, u, (size_t)l);
            ndig = l;
        }
    }
    if (ndig == 1) {
        /* Java always shows two digits and picks the 2-digit decimal closest to x (1.4E-45, not 1.0E-45) */
        char t[32];
        snprintf(t, sizeof t, "%.1e", (double)x);
        dig[0] = t[0]; dig[1] = t[2];
        e10 = atoi(strchr(t, 'e') + 1);
        ndig = dig[1] == '0' ? 1 : 2;
    }
    if (x >= 1e-3f && x < 1e7f) {
        if (e10 >= 0) {
            for (int i = 0; i <= e10; i++) *o++ = i < ndig ? dig[i] : '0';
            *o++ = '.';
            if (ndig > e10 + 1) for (int i = e10 + 1; i < ndig; i++) *o++ = dig[i];
            else *o++ = '0';
        } else {
            *o++ = '0'; *o++ = '.';
            for (int i = -1; i > e10; i--) *o++ = '0';
            for (int i = 0; i < ndig; i++) *o++ = dig[i];
        }
    } else {
        *o++ = dig[0]; *o++ = '.';
        if (ndig > 1) for (int i = 1; i < ndig; i++) *o++ = dig[i];
        else *o++ = '0';
        *o++ = 'E';
        o += sprintf(o, "%d", e10);
    }
    return (int)(o - buf);
}

/* {"predictions":[[p,..],..]}  (PredObj.java:8-17). Returns length or SIBO_ECAP. */
long sibo_json_encode(const float *probs, long n, int ncls, char *buf, size_t cap) {
    size_t need = 32 + (size_t)n * ((size_t)ncls * 18 + 4);
    if (cap < need) return SIBO_ECAP;
    char *o = buf;
    o += sprintf(o, "{\"predictions\":[");
    for (long i = 0; i < n; i++) {
        if (i) *o++ = ',';
        *o++ = '[';
        for (int k = 0; k < ncls; k++) {
            if (k) *o++ = ',';
            o += sibo_float_to_string(probs[i * ncls + k], o);
        }
        *o++ = ']';
    }
    *o++ = ']'; *o++ = '}';
    *o = 0;
    return (long)(o - buf);
}

/*
 * Mirror of InferenceBolt.execute() for one tuple (InferenceBolt.java:69-100): JSON in -> JSON out.
 * Returns output length; SIBO_EJSON means "emit null message and ack" (:92-99);
 * SIBO_ESHAPE means the reference would have thrown out of execute() (documented superset: N>1 rows allowed).
 */
long sibo_execute(const sibo_model *m, const char *json_in, size_t len, char *json_out, size_t cap) {
    size_t img = (size_t)m->in_h * m->in_w * m->in_c;
    long maxf = (long)(len / 2 + 8);
    float *x = (float *)malloc(sizeof(float) * (size_t)maxf);
    if (!x) return SIBO_ECAP;
    long dims[4];
    long nf = sibo_json_decode(json_in, len, x, maxf, dims);
    if (nf < 0) { free(x); return nf; }
    if ((uint32_t)dims[1] != m->in_h || (uint32_t)dims[2] != m->in_w || (uint32_t)dims[3] != m->in_c) {
        free(x);
        return SIBO_ESHAPE;
    }
    long n = dims[0];
    float *pr = (float *)malloc(sizeof(float) * (size_t)n * m->n_classes);
    float *lg = (float *)malloc(sizeof(float) * (size_t)n * m->n_classes);
    float *scratch = (float *)malloc(sizeof(float) * 3 * m->max_act);
    long r = SIBO_ECAP;
    if (pr && lg && scratch) {
        for (long i = 0; i < n; i++)
            forward_one_f32(m, x + (size_t)i * img, lg + i * m->n_classes, pr + i * m->n_classes, scratch);
        r = sibo_json_encode(pr, n, (int)m->n_classes, json_out, cap);
    }
    free(x); free(pr); free(lg); free(scratch);
    return r;
}

typedef struct { const sibo_model *m; const char *tuples; const long *offs; long ok[256], bytes[256]; } tup_ctx;

static void tup_range(void *vp, long lo, long hi, int tid) {
    tup_ctx *c = (tup_ctx *)vp;
    long ok = 0, bytes = 0;
    for (long i = lo; i < hi; i++) {
        char out[4096];
        long r = sibo_execute(c->m, c->tuples + c->offs[i], (size_t)(c->offs[i + 1] - c->offs[i]), out, sizeof out);
        if (r > 0) { ok++; bytes += r; }
    }
    c->ok[tid] = ok; c->bytes[tid] = bytes;
}

/*
 * Timed per-tuple driver (the reference's shape: batch 1 per call, one thread per "executor").
 * tuples: n_tuples JSON strings packed back to back, offs[n_tuples+1]. Each thread owns a contiguous share.
 * Returns number of tuples that produced a message; out_bytes accumulates output sizes (anti-DCE).
 */
long sibo_run_tuples(const sibo_model *m, const char *tuples, const long *offs, long n_tuples, int threads,
                     long *out_bytes) {
    tup_ctx *c = (tup_ctx *)calloc(1, sizeof *c);
    if (!c) return -1;
    c->m = m; c->tuples = tuples; c->offs = offs;
    parallel_for(n_tuples, threads, tup_range, c);
    long ok = 0, bytes = 0;
    for (int t = 0; t < 256; t++) { ok += c->ok[t]; bytes += c->bytes[t]; }
    free(c);
    if (out_bytes) *out_bytes = bytes;
    return ok;
}

size_t sibo_sizeof_model(void) { return sizeof(sibo_model); }

/* ------------------------------------------------------------------ open-loop latency of the CPU path */
/*
 * The reference's shape under load: `threads` executors (INFERENCE_BOLT_PARAL, MainTopology.java:27) each run
 * execute() on one tuple at a time; tuples arrive on a fixed schedule (tuple i at i/rate seconds) and are dealt to
 * whichever executor is free next (what shuffle grouping converges to). Latency = completion - scheduled arrival,
 * so queueing counts; when rate > capacity the latencies grow without bound (reported as is, the run is bounded by
 * n_tuples). mode 0: JSON in -> JSON out (sibo_execute); mode 1: fp32 image in -> probabilities (model only).
 */
#include <time.h>
static long long now_ns(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (long long)ts.tv_sec * 1000000000ll + ts.tv_nsec;
}

typedef struct {
    const sibo_model *m;
    const void *const *payloads;
    const size_t *lens;
    int n_payloads, mode;
    double gap_ns;
    long n_tuples;
    long long t0;
    long long *lat;
    volatile long next;
} lat_ctx;

static void *lat_worker(void *vp) {
    lat_ctx *c = (lat_ctx *)vp;
    float *scratch = (float *)malloc(sizeof(float) * 3 * c->m->max_act);
    char out[4096];
    float lg[64], pr[64];
    for (;;) {
        long i = __sync_fetch_and_add(&c->next, 1);
        if (i >= c->n_tuples) break;
        long long due = c->t0 + (long long)((double)i * c->gap_ns);
        while (now_ns() < due) { /* spin: arrival not yet due */ }
        int k = (int)(i % c->n_payloads);
        if (c->mode == 0) sibo_execute(c->m, (const char *)c->payloads[k], c->lens[k], out, sizeof out);
        else forward_one_f32(c->m, (const float *)c->payloads[k], lg, pr, scratch);
        c->lat[i] = now_ns() - due;
    }
    free(scratch);
    return NULL;
}

int sibo_latency_run(const sibo_model *m, const void *const *payloads, const size_t *lens, int n_payloads, int mode,
                     double rate_per_s, long n_tuples, int threads, long long *lat_out) {
    if (threads < 1) threads = sibo_max_threads();
    if (threads > 256) threads = 256;
    lat_ctx c;
    c.m = m; c.payloads = payloads; c.lens = lens; c.n_payloads = n_payloads; c.mode = mode;
    c.gap_ns = 1e9 / rate_per_s; c.n_tuples = n_tuples; c.lat = lat_out; c.next = 0;
    c.t0 = now_ns() + 2000000;
    pthread_t th[256];
    for (int t = 0; t < threads; t++) pthread_create(&th[t], NULL, lat_worker, &c);
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    return 0;
}
