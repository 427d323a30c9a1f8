/*
 * sib_bolt.h -- C entry points of the bolt-contract harness.
 *
 * libsib.so carries a C++ mirror of the reference's bolt (class dke::model::InferenceBolt in csrc/sib_bolt.hpp:
 * prepare / execute / declareOutputFields / cleanup, same argument meaning and error behaviour as
 * InferenceBolt.java:43-105) because no JVM exists on the build or GPU boxes. These functions play the Storm
 * executor around it: one thread calls execute() per tuple, a recording OutputCollector logs every emit and ack.
 * tests/test_bolt_contract.py drives them; the Java bolt in java/dke/model/ follows the same logic over JNI.
 */
#ifndef SIB_BOLT_H
#define SIB_BOLT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct sib_harness sib_harness;

#define SIB_EV_EMIT 1 /* OutputCollector.emit(new Values(message))  InferenceBolt.java:98 */
#define SIB_EV_ACK 2  /* OutputCollector.ack(tuple)                 InferenceBolt.java:99 */

/* conf: "key=value\n" lines standing in for the Storm conf Map of prepare() (InferenceBolt.java:44). Keys:
 * sib.max.batch, sib.deadline.us, sib.precision (fp16|bf16|fp32), sib.device, sib.slots, sib.guard, sib.guard.tau.
 * task_index: TopologyContext.getThisTaskIndex(). Runs new InferenceBolt() + prepare(). */
int sib_harness_open(const void *model_blob, size_t n, const char *conf, int task_index, sib_harness **out);
/* execute(tuple) with tuple.getString(0) = utf8. Returns 0, or SIB_ESHAPE when execute() threw the
 * IllegalArgumentException the reference would have thrown (the tuple is then neither emitted nor acked). */
int sib_harness_execute(sib_harness *h, uint64_t tuple_id, const char *utf8, size_t len);
/* cleanup(): flush, wait for every pending tuple, stop the completion thread. */
int sib_harness_cleanup(sib_harness *h);
int sib_harness_close(sib_harness *h);
/* recorded collector events, in the order the collector saw them */
long sib_harness_event_count(sib_harness *h);
int sib_harness_event(sib_harness *h, long i, int *kind, uint64_t *tuple_id, const char **msg, size_t *msg_len,
                      int64_t *t_ns);
/* declareOutputFields(): writes the declared field names, comma separated ("message"). */
int sib_harness_output_fields(sib_harness *h, char *buf, size_t cap);
/* wait until at least n events were recorded or timeout_us passed; returns the event count */
long sib_harness_wait_events(sib_harness *h, long n, long timeout_us);

#ifdef __cplusplus
}
#endif
#endif
