// sib_jni.cpp -- Java_dke_model_NativeInfer_* shims over the C ABI (java/dke/model/NativeInfer.java).
// Compiled only where a JDK's jni.h is on the include path; neither the build box nor the GPU box has one, so on
// those this translation unit is empty and the contract is exercised through csrc/sib_bolt.cpp instead.
#if defined(__has_include)
#if __has_include(<jni.h>)
#include <jni.h>

#include <string>
#include <vector>

#include "../../include/sib.h"

// One Java long = one JHandle: the executor's own engine, or (cfg[9] = 1, Storm conf sib.shared) its view of the
// worker process's shared engine (sib_client_*).
struct JHandle {
    sib_engine* e = nullptr;
    sib_client* c = nullptr;
};

extern "C" {

JNIEXPORT jlong JNICALL Java_dke_model_NativeInfer_open(JNIEnv* env, jclass, jbyteArray blob, jintArray cfg, jfloat tau) {
    jsize n = env->GetArrayLength(blob);
    std::vector<jbyte> b((size_t)n);
    env->GetByteArrayRegion(blob, 0, n, b.data());
    jint c[10] = {-1, 0, 0, 0, 0, 0, 1, 0, 0, 0};  // device, task, max_batch, deadline_us, precision, slots, guard, json_on_gpu, strict_batching, shared
    const jsize nc = env->GetArrayLength(cfg);
    env->GetIntArrayRegion(cfg, 0, nc < 10 ? nc : 10, c);
    sib_config sc{c[0], c[1], c[2], c[3], c[4], c[5], c[6], tau, c[7], c[8]};
    JHandle* h = new JHandle;
    const int rc = c[9] ? sib_client_open(b.data(), (size_t)n, &sc, &h->c) : sib_engine_open(b.data(), (size_t)n, &sc, &h->e);
    if (rc != SIB_OK) { delete h; return 0; }
    return (jlong)(intptr_t)h;
}

JNIEXPORT void JNICALL Java_dke_model_NativeInfer_close(JNIEnv*, jclass, jlong hh) {
    JHandle* h = (JHandle*)(intptr_t)hh;
    if (!h) return;
    if (h->c) sib_client_close(h->c); else sib_engine_close(h->e);
    delete h;
}

JNIEXPORT jint JNICALL Java_dke_model_NativeInfer_submitJson(JNIEnv* env, jclass, jlong hh, jlong tag, jbyteArray utf8) {
    // The submit takes the engine's mutex and may answer SIB_EAGAIN after waiting on it: none of that is allowed inside
    // a Get/ReleasePrimitiveArrayCritical region (no blocking, no other JNI calls), so the text is copied out first
    // (10-40 KB, one memcpy; the decoder then writes the floats straight into the pinned slot).
    JHandle* h = (JHandle*)(intptr_t)hh;
    const jsize n = env->GetArrayLength(utf8);
    static thread_local std::vector<jbyte> text;
    if ((jsize)text.size() < n) text.resize((size_t)n);
    env->GetByteArrayRegion(utf8, 0, n, text.data());
    return h->c ? sib_client_submit_json(h->c, (uint64_t)tag, (const char*)text.data(), (size_t)n)
                : sib_submit_json(h->e, (uint64_t)tag, (const char*)text.data(), (size_t)n);
}

// Batched egress: one crossing per poll. `arena` is a direct ByteBuffer the bolt keeps; the messages land in it back to
// back, tags/offsets/status describe them (sib_poll_json_batch). Own engines only (h->e).
JNIEXPORT jint JNICALL Java_dke_model_NativeInfer_pollJsonBatch(JNIEnv* env, jclass, jlong hh, jobject arena, jlongArray tags,
                                                                 jintArray offsets, jintArray status, jint timeout_us) {
    JHandle* h = (JHandle*)(intptr_t)hh;
    if (!h || !h->e) return SIB_EINVAL;
    char* base = (char*)env->GetDirectBufferAddress(arena);
    const jlong cap = env->GetDirectBufferCapacity(arena);
    const int max = env->GetArrayLength(tags);
    if (!base || cap <= 0 || env->GetArrayLength(offsets) < max + 1 || env->GetArrayLength(status) < max) return SIB_EINVAL;
    static thread_local std::vector<uint64_t> t;
    static thread_local std::vector<uint32_t> o;
    static thread_local std::vector<int32_t> st;
    t.resize((size_t)max); o.resize((size_t)max + 1); st.resize((size_t)max);
    const int n = sib_poll_json_batch(h->e, base, (size_t)cap, t.data(), o.data(), st.data(), max, timeout_us);
    if (n > 0) {
        env->SetLongArrayRegion(tags, 0, n, (const jlong*)t.data());
        env->SetIntArrayRegion(offsets, 0, n + 1, (const jint*)o.data());
        env->SetIntArrayRegion(status, 0, n, (const jint*)st.data());
    }
    return n;
}

JNIEXPORT void JNICALL Java_dke_model_NativeInfer_flush(JNIEnv*, jclass, jlong hh) {
    JHandle* h = (JHandle*)(intptr_t)hh;
    if (h->c) sib_client_flush(h->c); else sib_flush(h->e);
}

JNIEXPORT jint JNICALL Java_dke_model_NativeInfer_pollJson(JNIEnv* env, jclass, jlong hh, jlongArray tags, jobjectArray msgs, jint timeout_us) {
    JHandle* h = (JHandle*)(intptr_t)hh;
    const int cap = env->GetArrayLength(tags);
    std::vector<sib_result> res((size_t)cap);
    int n = h->c ? sib_client_poll(h->c, res.data(), cap, timeout_us) : sib_poll(h->e, res.data(), cap, timeout_us);
    std::vector<char> buf(1 << 14);
    for (int i = 0; i < n; i++) {
        jlong t = (jlong)res[i].tag;
        env->SetLongArrayRegion(tags, i, 1, &t);
        jstring s = nullptr;
        if (res[i].status == SIB_OK) {
            size_t need = 64 + (size_t)res[i].n * ((size_t)res[i].n_classes * 18 + 4);
            if (buf.size() < need) buf.resize(need);
            if (sib_format_json(&res[i], buf.data(), buf.size()) > 0) s = env->NewStringUTF(buf.data());
        }
        env->SetObjectArrayElement(msgs, i, s);
    }
    return n < 0 ? 0 : n;
}

JNIEXPORT jstring JNICALL Java_dke_model_NativeInfer_lastError(JNIEnv* env, jclass) { return env->NewStringUTF(sib_last_error()); }

}  // extern "C"
#endif
#endif
