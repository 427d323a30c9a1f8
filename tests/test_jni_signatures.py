"""No JDK exists on either box, so java/dke/model/NativeInfer.java and csrc/sib_jni.cpp are delivered uncompiled
(INTEGRATION.md). What can be checked without one: every `native` method the Java class declares has exactly one
`Java_dke_model_NativeInfer_<name>` definition in the shim, with the JNI type each Java parameter and the return type map
to, in the same order — the mismatch a JVM would only report as UnsatisfiedLinkError at the first call — and the Java
bolt calls no native method that is not declared."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-inference-system-based-storm_b200")

JNI_TYPE = {
    "long": "jlong", "int": "jint", "float": "jfloat", "void": "void", "String": "jstring",
    "byte[]": "jbyteArray", "int[]": "jintArray", "long[]": "jlongArray", "String[]": "jobjectArray",
    "java.nio.ByteBuffer": "jobject", "ByteBuffer": "jobject",
}


def java_natives():
    src = open(os.path.join(PKG, "java", "dke", "model", "NativeInfer.java")).read()
    out = {}
    for ret, name, params in re.findall(r"public\s+static\s+native\s+([\w.\[\]]+)\s+(\w+)\s*\(([^)]*)\)\s*;", src):
        ptypes = [p.strip().rsplit(" ", 1)[0].strip() for p in params.split(",") if p.strip()]
        out[name] = (ret, ptypes)
    return out


def shim_definitions():
    src = open(os.path.join(PKG, "csrc", "sib_jni.cpp")).read()
    out = {}
    for ret, name, params in re.findall(r"JNIEXPORT\s+(\w+)\s+JNICALL\s+Java_dke_model_NativeInfer_(\w+)\s*\(([^)]*)\)", src, re.S):
        ptypes = [p.strip().split()[0].rstrip("*") for p in params.replace("\n", " ").split(",") if p.strip()]
        assert name not in out, f"{name} is defined twice"
        out[name] = (ret, ptypes)
    return out


def test_every_native_method_has_a_matching_shim():
    j, c = java_natives(), shim_definitions()
    assert len(j) >= 7 and set(j) == set(c), (sorted(j), sorted(c))
    for name, (ret, ptypes) in j.items():
        cret, cparams = c[name]
        assert cparams[:2] == ["JNIEnv", "jclass"], name        # static native: (JNIEnv*, jclass, ...)
        assert cret == JNI_TYPE[ret], (name, ret, cret)
        assert cparams[2:] == [JNI_TYPE[t] for t in ptypes], (name, ptypes, cparams[2:])


def test_the_bolt_calls_only_declared_natives():
    bolt = open(os.path.join(PKG, "java", "dke", "model", "InferenceBolt.java")).read()
    called = set(re.findall(r"NativeInfer\.(\w+)\s*\(", bolt))
    assert called and called <= set(java_natives()), called - set(java_natives())
    assert {"open", "submitJson", "pollJsonBatch", "close"} <= called
