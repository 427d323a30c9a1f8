/*
 * sib_loadgen.h -- open-loop arrival generator for the latency configs of BASELINE.json ("p50/p99 end-to-end tuple
 * latency at fixed arrival rates"). It plays the KafkaSpout side of MainTopology.java:61-62:
 * tuples arrive on a schedule that does not care how fast the bolt is, a consumer thread plays the downstream
 * KafkaBolt (takes every result, formats the PredObj JSON in JSON mode). Latency of a tuple = time its result is
 * available - its SCHEDULED arrival time, so queueing delay counts.
 */
#ifndef SIB_LOADGEN_H
#define SIB_LOADGEN_H
#include <stddef.h>
#include <stdint.h>

#include "sib.h"
#ifdef __cplusplus
extern "C" {
#endif

#define SIB_LG_JSON 0 /* payloads are {"instances":...} texts -> sib_submit_json, result formatted with sib_format_json (T2) */
#define SIB_LG_F32 1  /* payloads are fp32 NHWC arrays (payload_len = bytes) -> sib_submit_f32 (T1) */

typedef struct sib_loadgen_cfg {
    double rate_per_s;  /* tuples per second, all producers together */
    int64_t n_tuples;
    int32_t mode;       /* SIB_LG_* */
    int32_t producers;  /* submitting threads (executors feeding the engine), >= 1 */
    int32_t poisson;    /* 0 = fixed interval, 1 = exponential gaps */
    int32_t reserved;
    uint64_t seed;
} sib_loadgen_cfg;

typedef struct sib_loadgen_report {
    int64_t completed, rejected;   /* rejected: tuples the engine refused (SIB_EJSON / SIB_ESHAPE) */
    int64_t images;
    double wall_s, achieved_tuples_per_s;
    int64_t p50_ns, p90_ns, p99_ns, p999_ns, max_ns, mean_ns;
    int64_t eagain_spins;          /* back-pressure events seen by the producers */
    int64_t late_submits;          /* submits that started > 100 us after their scheduled arrival (producer-bound) */
} sib_loadgen_report;

/* payload i of the run is payloads[i % n_payloads]. latencies_ns (may be NULL) receives n_tuples values, -1 for rejected. */
int sib_loadgen_run(sib_engine *e, const sib_loadgen_cfg *cfg, const void *const *payloads, const size_t *payload_lens,
                    int n_payloads, int64_t *latencies_ns, sib_loadgen_report *out);

#ifdef __cplusplus
}
#endif
#endif
