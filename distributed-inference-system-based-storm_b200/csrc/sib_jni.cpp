// sib_jni.cpp -- Java_dke_model_NativeInfer_* shims over the C ABI (java/dke/model/NativeInfer.java).
// Compiled only where a JDK's jni.h is on the include path; neither the build box nor the GPU box has one, so on
// those this translation unit is empty and the contract is exercised through csrc/sib_bolt.cpp instead.
#if defined(__has_include)
#if __has_include(<jni.h>)
#include <jni.h>

#include <string>
#include <vector>

#include "../../include/sib.h"

extern "C" {

JNIEXPORT jlong JNICALL Java_dke_model_NativeInfer_open(JNIEnv* env, jclass, jbyteArray blob, jintArray cfg, jfloat tau) {
    jsize n = env->GetArrayLength(blob);
    std::vector<jbyte> b((size_t)n);
    env->GetByteArrayRegion(blob, 0, n, b.data());
    jint c[9] = {-1, 0, 0, 0, 0, 0, 1, 0, 0};  // device, task, max_batch, deadline_us, precision, slots, guard, json_on_gpu, strict_batching
    const jsize nc = env->GetArrayLength(cfg);
    env->GetIntArrayRegion(cfg, 0, nc < 9 ? nc : 9, c);
    sib_config sc{c[0], c[1], c[2], c[3], c[4], c[5], c[6], tau, c[7], c[8]};
    sib_engine* e = nullptr;
    if (sib_engine_open(b.data(), (size_t)n, &sc, &e) != SIB_OK) return 0;
    return (jlong)(intptr_t)e;
}

JNIEXPORT void JNICALL Java_dke_model_NativeInfer_close(JNIEnv*, jclass, jlong h) { sib_engine_close((sib_engine*)(intptr_t)h); }

JNIEXPORT jint JNICALL Java_dke_model_NativeInfer_submitJson(JNIEnv* env, jclass, jlong h, jlong tag, jbyteArray utf8) {
    jsize n = env->GetArrayLength(utf8);
    jbyte* p = (jbyte*)env->GetPrimitiveArrayCritical(utf8, nullptr);   // no copy: decoded straight into the pinned slot
    jint rc = sib_submit_json((sib_engine*)(intptr_t)h, (uint64_t)tag, (const char*)p, (size_t)n);
    env->ReleasePrimitiveArrayCritical(utf8, p, JNI_ABORT);
    return rc;
}

JNIEXPORT void JNICALL Java_dke_model_NativeInfer_flush(JNIEnv*, jclass, jlong h) { sib_flush((sib_engine*)(intptr_t)h); }

JNIEXPORT jint JNICALL Java_dke_model_NativeInfer_pollJson(JNIEnv* env, jclass, jlong h, jlongArray tags, jobjectArray msgs, jint timeout_us) {
    const int cap = env->GetArrayLength(tags);
    std::vector<sib_result> res((size_t)cap);
    int n = sib_poll((sib_engine*)(intptr_t)h, res.data(), cap, timeout_us);
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
