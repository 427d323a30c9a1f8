/*
 * sib.h -- C ABI of libsib.so: the B200-native replacement for the InferenceBolt hot path of
 * HyoJong-Moon/Distributed-Inference-System-based-Storm.
 *
 * Every entry point is what a JNI / Panama binding of the reference's bolt would bind (INTEGRATION.md shows
 * the Java side).  Plain pointers and sizes only; no C++ or torch types; no exception crosses this boundary.
 * All functions return 0 / a non-negative count on success and a negative SIB_E* code on failure;
 * sib_last_error() gives text for the calling thread's last failure.
 *
 * Reference citations are relative to /root/reference/src/main/java/dke/model/.
 */
#ifndef SIB_H
#define SIB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes ------------------------------------------------------------------------------------- */
#define SIB_OK 0
#define SIB_EJSON (-1)   /* malformed JSON / unknown property: the reference catches JsonProcessingException,
                            emits a null message and acks (InferenceBolt.java:92-99) */
#define SIB_ESHAPE (-2)  /* ragged / wrong rank / wrong HxWxC / missing instances: the reference throws out of
                            execute() (Tensor.create / copyTo IllegalArgumentException, InferenceBolt.java:80,86) */
#define SIB_ECAP (-3)    /* caller buffer too small / tuple larger than max_batch */
#define SIB_EAGAIN (-4)  /* every batch slot is busy: call sib_poll() and retry */
#define SIB_EINVAL (-5)  /* bad argument / bad model blob */
#define SIB_ECUDA (-6)   /* CUDA runtime failure, or no usable GPU (there is NO CPU fallback) */
#define SIB_ECLOSED (-7) /* engine is shutting down */

/* ---- precision of the forward pass -------------------------------------------------------------------- */
#define SIB_PREC_FP16 0 /* tensor-core path, half operands (11-bit significand), fp32 accumulate, fp32 margin guard.
                           Default: same tensor rate as BF16 with 8x smaller logit error; inputs are in [0,1) and
                           activations of these nets stay far below 65504 */
#define SIB_PREC_BF16 1 /* tensor-core path, bfloat16 operands (8-bit significand, fp32 range), fp32 accumulate, guard */
#define SIB_PREC_FP32 2 /* CUDA-core fp32 path (any model the blob can describe; also the guard's recompute path) */

typedef struct sib_engine sib_engine;

/* Engine configuration. Zero / negative fields take the defaults noted. Read from the Storm conf Map that
 * prepare() receives (InferenceBolt.java:44) so the topology CLI keeps its 3 arguments (MainTopology.java:36-38). */
typedef struct sib_config {
    int32_t device;      /* CUDA ordinal; <0 = task_index % device_count  (MainTopology.java:62 shuffle grouping
                            x parallelism hint is the multi-GPU sharding; no collective) */
    int32_t task_index;  /* TopologyContext.getThisTaskIndex() of this executor (InferenceBolt.java:44) */
    int32_t max_batch;   /* images per micro-batch, flush-by-size              (default 4096) */
    int32_t deadline_us; /* age of the oldest queued tuple that forces a flush (default 200)  */
    int32_t precision;   /* SIB_PREC_*                                          (default FP16) */
    int32_t n_slots;     /* pinned batch slots in the ring, >= 2               (default 4)    */
    int32_t guard;       /* 1 = recompute low-margin images in fp32 so argmax equals the fp32 path (default 1) */
    float guard_tau;     /* top1-top2 logit margin below which the guard recomputes (default 0.05 FP16, 0.30 BF16) */
    int32_t json_on_gpu; /* 1 = sib_submit_json ships the raw UTF-8 to the GPU and a CUDA kernel decodes it (one
                            instance per tuple, canonical nesting); tuples it cannot take are re-read by the host
                            decoder when their batch completes, so results are the same either way. In this mode a
                            malformed or wrong-shaped tuple is reported in sib_result.status (SIB_EJSON / SIB_ESHAPE,
                            n = 0) instead of by sib_submit_json's return value. 2 = auto: an idle engine decodes on
                            the submitting thread (lowest single-tuple latency), a busy one ships the text (highest
                            rate); a bad tuple is then reported one way or the other. (default 0) */
    int32_t strict_batching; /* 0 (default): a tuple that finds no batch queued or in flight is launched as soon as it
                            is written, so at low arrival rates nobody sits out deadline_us, and under load batches
                            fill while the previous one runs (p50 39 us instead of 215 us at 10 k tuples/s, 115 us
                            instead of 265 us at 1 M tuples/s: profiles/latency_r01.jsonl). 1: batches are launched
                            by max_batch, deadline_us and sib_flush only. */
} sib_config;

/* One completed tuple. probs/argmax are engine-owned and stay valid until the next sib_poll() on the engine. */
typedef struct sib_result {
    uint64_t tag;          /* caller's tuple handle, returned untouched */
    int32_t n;             /* images in the tuple (rows of predictions) */
    int32_t n_classes;     /* columns of predictions (10 for the reference: InferenceBolt.java:86) */
    int32_t status;        /* SIB_OK, SIB_ECUDA, or (json_on_gpu) SIB_EJSON / SIB_ESHAPE with n = 0 and probs = NULL */
    int32_t reserved;
    const float *probs;    /* n x n_classes softmax probabilities, "output/Softmax:0" (InferenceBolt.java:83) */
    const int32_t *argmax; /* n predicted classes (first maximum) */
} sib_result;

typedef struct sib_stats {
    uint64_t tuples, images, batches;
    uint64_t flush_by_size, flush_by_deadline,
        flush_explicit; /* sib_flush, engine close, and launches made because the GPU was idle (strict_batching = 0) */
    uint64_t guard_recomputed; /* images the margin guard re-ran in fp32 */
    uint64_t batch_hist[16];   /* batch sizes, bucket b counts batches with 2^b <= images < 2^(b+1) */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t kernel_launches;
    uint64_t json_gpu_decoded; /* json_on_gpu: tuples the decode kernel took */
    uint64_t json_host_redone; /* json_on_gpu: tuples it declined, re-read by the host decoder at completion */
} sib_stats;

typedef struct sib_model_info {
    int32_t in_h, in_w, in_c, n_layers, n_classes, n_weights;
    int64_t flops_per_image; /* 2 * sum(M*N*K) over conv/dense layers */
    int64_t bytes_per_image; /* H*W*C*4 in + n_classes*4 out */
} sib_model_info;

/* ---- life cycle: replaces InferenceBolt.prepare()'s SavedModelBundle.load / session (InferenceBolt.java:43-62) */
int sib_device_count(void);
int sib_model_describe(const void *model_blob, size_t n, sib_model_info *out);
/* Everywhere a model blob is taken, the reference's own artefact -- the bytes of model/saved_model.pb, a TensorFlow
 * SavedModel (or frozen GraphDef) with constant weights, InferenceBolt.java:49-58 -- is accepted as well and
 * converted on the fly. This call exposes the conversion: writes the blob into out (capacity cap) and returns its
 * length; out == NULL returns the length needed; SIB_ECAP if cap is too small; SIB_EINVAL (text in
 * sib_last_error) if the graph is not Placeholder -> [Conv2D VALID s1 + BiasAdd + Relu + MaxPool 2x2/2]* ->
 * Reshape -> [MatMul + BiasAdd (+ Relu)]* -> Softmax with frozen weights. Replaces SavedModelBundle.load (:57). */
long sib_model_import_pb(const void *saved_model_pb, size_t n, void *out, size_t cap);
int sib_engine_open(const void *model_blob, size_t n, const sib_config *cfg, sib_engine **out);
int sib_engine_close(sib_engine *e); /* flushes, waits for in-flight batches, frees everything */
int sib_engine_info(sib_engine *e, sib_model_info *out, sib_config *effective_cfg);

/* ---- one engine for all the bolts of a worker process ------------------------------------------------------------
 * Storm puts several executors of the bolt into one worker JVM (MainTopology.java:27, :65-66); the reference gives
 * each its own model copy (InferenceBolt.java:57-58). sib_client_open returns a view of a process-wide,
 * reference-counted engine keyed by (resolved device, model bytes, configuration): co-located bolts then fill the
 * same batches through one set of streams and one weight upload. A client sees only its own tuples, in its own
 * submission order; its results live in client-owned memory until its next sib_client_poll, so one bolt's polling
 * pace never holds the others back. Same return codes as the sib_engine calls of the same name. The last
 * sib_client_close closes the engine. No counterpart in the reference. */
typedef struct sib_client sib_client;
int sib_client_open(const void *model, size_t n, const sib_config *cfg, sib_client **out);
int sib_client_close(sib_client *c);
int sib_client_submit_json(sib_client *c, uint64_t tag, const char *utf8, size_t len);
int sib_client_submit_f32(sib_client *c, uint64_t tag, const float *nhwc, int n);
int sib_client_poll(sib_client *c, sib_result *out, int max, int timeout_us);
int sib_client_flush(sib_client *c);
int sib_client_stats(sib_client *c, sib_stats *out); /* of the shared engine */

/* ---- per-tuple path: replaces the body of InferenceBolt.execute() (InferenceBolt.java:69-100) --------- */

/* readValue(inputJson, InstObj.class) + Tensor.create(data) (InferenceBolt.java:76-80): decodes
 * {"instances":[N][H][W][C]} straight into the pinned batch slot. Returns N (>0) or a negative code:
 * SIB_EJSON => caller emits a null message and acks; SIB_ESHAPE => caller throws IllegalArgumentException. */
int sib_submit_json(sib_engine *e, uint64_t tag, const char *utf8, size_t len);

/* Same, for a caller that already holds fp32 NHWC (benchmarks; a TensorFlow-Java style caller). Input is copied
 * before return. */
int sib_submit_f32(sib_engine *e, uint64_t tag, const float *nhwc, int n);

/* Zero-copy form of sib_submit_f32 for a caller whose fp32 NHWC already sits in page-locked host memory (a direct
 * ByteBuffer registered with cudaHostRegister, cudaHostAlloc memory): the DMA reads the caller's buffer, nothing is
 * staged. The buffer must stay untouched until the tuple's result has been returned by sib_poll. */
int sib_submit_pinned(sib_engine *e, uint64_t tag, const float *pinned_nhwc, int n);

/* sess.runner().feed().fetch().run() + copyTo (InferenceBolt.java:81-86), decoupled: returns up to max completed
 * tuples in submission order. timeout_us: 0 = don't wait, <0 = wait until at least one is ready or nothing is
 * pending. Returns the count. Invalidates pointers handed out by the previous poll. */
int sib_poll(sib_engine *e, sib_result *out, int max, int timeout_us);

/* Batched egress towards KafkaBolt (KafkaBolt.java:116-166 takes one "message" String per tuple):
 * sib_poll + sib_format_json for up to max tuples in ONE native pass. The messages are written back to back into
 * `arena` (no NUL between them); message i is arena[offsets[i] .. offsets[i+1]) and offsets must hold max+1 entries.
 * status[i] != SIB_OK (SIB_EJSON, SIB_ESHAPE, ...) => empty range => the bolt emits a null message for that tuple
 * (InferenceBolt.java:92-98). A tuple that no longer fits the arena stays queued for the next call; cap must hold at
 * least one message (32 + rows * (18 * classes + 4) bytes). A binding turns the arena into Strings and emits them
 * under one collector monitor per batch instead of one JNI crossing, one allocation and one lock per tuple.
 * Returns the tuple count like sib_poll. */
int sib_poll_json_batch(sib_engine *e, char *arena, size_t cap, uint64_t *tags, uint32_t *offsets, int32_t *status,
                        int max, int timeout_us);

/* Launch the partially filled batch now (bolt cleanup(), tests). */
int sib_flush(sib_engine *e);

/* new ObjectMapper().writeValueAsString(predObj) (InferenceBolt.java:89-91, PredObj.java:8-17):
 * {"predictions":[[p0,..,p9],..]} with java.lang.Float.toString number text. Returns length (no NUL) or SIB_ECAP. */
long sib_format_json(const sib_result *r, char *buf, size_t cap);

int sib_stats_get(sib_engine *e, sib_stats *out);

/* ---- whole-batch synchronous forms (tests, benchmarks, oracle diff) ----------------------------------- */

/* Host buffers in, host buffers out: H2D + forward + D2H on the engine's stream, then waits.
 * probs: n x n_classes; logits (may be NULL): n x n_classes; argmax (may be NULL): n. */
int sib_forward_sync(sib_engine *e, const float *nhwc_host, int n, float *probs, float *logits, int32_t *argmax);

/* Device-resident form (kernel-level timing): pointers are CUDA device pointers on the engine's device, stream is
 * a cudaStream_t (NULL = the CUDA default stream). Asynchronous: the work is only ordered on that stream. d_logits / d_argmax may be NULL. Needs n <= max_batch.
 * With the guard on, all calls on one engine share the guard's scratch (margins, low-margin list, counter): calls whose
 * GPU work overlaps (different streams, no ordering between them) race on it -- order them, or give each stream its own
 * engine. */
int sib_forward_device(sib_engine *e, const float *d_nhwc, int n, float *d_probs, float *d_logits,
                       int32_t *d_argmax, void *stream);

/* No counterpart in the reference (its TF session allocates as it pleases). Same work as sib_forward_device on the
 * guarded tensor-core path, but the guard's scratch (margins, low-margin list, counter) is a fixed array inside
 * libsib.so's device image instead of engine-owned cudaMalloc memory, and the call itself allocates nothing.
 * For callers that must not ask the driver for device memory: a task attaching to a GPU whose memory is already
 * fully handed out (cudaMalloc would fail), a latency-critical loop that cannot afford cudaMalloc/cudaFree's
 * device-wide synchronisation stalling the other executors on the GPU, or a phase run under a no-allocation rule.
 * Limits: n <= SIB_STATIC_MAX_IMAGES per call (larger n => SIB_ECAP); engine must be SIB_PREC_FP16/BF16 with guard on;
 * d_argmax is required. The first call on a device resolves the array's address -- make it before any stream
 * capture on that device. All calls on a device share the array: calls whose GPU work overlaps race on it. */
#define SIB_STATIC_MAX_IMAGES 16384
int sib_forward_device_static(sib_engine *e, const float *d_nhwc, int n, float *d_probs, float *d_logits,
                              int32_t *d_argmax, void *stream);

/* Test hook: intermediate activations of the tensor-core kernel for the first 16-image tile, fp32 (what fetching
 * inner tensors of the TF graph would give, InferenceBolt.java:81-85). Returns floats written or a code. */
long sib_debug_stage_dump(sib_engine *e, const float *nhwc_host, int n, float *out, long cap);

/* ---- stand-alone codec (same code the engine uses) ---------------------------------------------------- */

/* Jackson semantics: number text -> double -> (float). dims receives {N,H,W,C}. Returns float count or code. */
long sib_json_decode(const char *utf8, size_t len, float *out, long cap, int64_t dims[4]);
long sib_json_encode(const float *probs, long n, int n_classes, char *buf, size_t cap);
int sib_float_to_string(float x, char *buf24); /* java.lang.Float.toString; returns length */

const char *sib_last_error(void);
const char *sib_strerror(int code);
const char *sib_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SIB_H */
