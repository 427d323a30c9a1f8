package dke.model;

/**
 * JNI face of libsib.so (include/sib.h). One long handle = the executor's sib_engine, or (cfg[9] = 1) its sib_client
 * view of the engine shared by the bolts of this worker JVM.
 * NOT COMPILED in this repository's CI: there is no JDK on the build or GPU boxes. Every method maps 1:1 to a
 * C entry point; csrc/sib_jni.cpp holds the Java_dke_model_NativeInfer_* shims (built when jni.h is present).
 */
public final class NativeInfer {
    static { System.loadLibrary("sib"); }

    private NativeInfer() {}

    /** sib_engine_open / sib_client_open. cfg = {device, taskIndex, maxBatch, deadlineUs, precision, nSlots, guard,
     *  jsonOnGpu, strictBatching, shared} (shorter arrays take defaults for the rest); returns handle, 0 on failure. */
    public static native long open(byte[] modelBlob, int[] cfg, float guardTau);
    public static native void close(long engine);
    /** sib_submit_json on the tuple's UTF-8 bytes: returns N (>0), or SIB_EJSON(-1) / SIB_ESHAPE(-2) / SIB_EAGAIN(-4). */
    public static native int submitJson(long engine, long tag, byte[] utf8);
    public static native void flush(long engine);
    /** sib_poll + sib_format_json: fills tags[] and messages[] (null message = failed batch); returns count. */
    public static native int pollJson(long engine, long[] tags, String[] messages, int timeoutUs);
    /** sib_poll_json_batch: one crossing per poll. The messages of up to tags.length tuples land back to back in the
     *  direct buffer `arena`; message i is arena[offsets[i], offsets[i+1]) (offsets.length >= tags.length + 1), an empty
     *  range with status[i] != 0 is a null message. Own engines only (not sib.shared). Returns the tuple count. */
    public static native int pollJsonBatch(long engine, java.nio.ByteBuffer arena, long[] tags, int[] offsets, int[] status, int timeoutUs);
    public static native String lastError();
}
