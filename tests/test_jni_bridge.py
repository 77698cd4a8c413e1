"""JNI bridge (native_bridge/VectorDBJNI_b200.cpp).

No JDK exists on this image, so the bridge is compiled against a declarations-only stand-in header (tests/jni_stub/jni.h)
and has never been loaded by a JVM — INTEGRATION.md says so. What CAN be held here:
  CPU: it compiles; every `Java_com_vectordb_jni_*` name the reference's javah headers declare
       (native/jni_headers/com_vectordb_jni_*.h:15-48, committed literally below because /root/reference does not travel) is
       exported with C linkage, plus the new NativeHnswPqIndex / NativeFlatIndex natives; the Java sources declare exactly the
       natives the headers list; the stand-in JNIEnv function table has the JNI specification's slot numbers.
  GPU: the bridge is driven through a mock JNIEnv (tests/jni_stub/mock_env.cpp) with Java-style arrays and direct buffers and
       must return what the C ABI returns, release every array it pinned, and turn error codes into Java exceptions."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BRIDGE = os.path.join(ROOT, "native_bridge")

# the reference's javah output, name -> JVM descriptor (native/jni_headers/com_vectordb_jni_NativeIndex.h:15-48,
# ..NativeHnswIndex.h:15-16, ..NativePqIndex.h:15-24, ..NativeIvfIndex.h, ..NativeLshIndex.h, ..NativeAnnoyIndex.h)
REFERENCE_NATIVES = {
    "Java_com_vectordb_jni_NativeIndex_nativeAdd": "(JI[F)V",
    "Java_com_vectordb_jni_NativeIndex_nativeSearch": "(J[FI[I[F)I",
    "Java_com_vectordb_jni_NativeIndex_nativeDestroy": "(J)V",
    "Java_com_vectordb_jni_NativeIndex_nativeAddBatch": "(JLjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;II)V",
    "Java_com_vectordb_jni_NativeIndex_nativeSearchBatch": "(JLjava/nio/ByteBuffer;IIILjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;)I",
    "Java_com_vectordb_jni_NativeHnswIndex_nativeCreateHNSW": "(IIIII)J",
    "Java_com_vectordb_jni_NativePqIndex_nativeCreatePQ": "(IIII)J",
    "Java_com_vectordb_jni_NativePqIndex_nativeTrain": "(JI[F)V",
    "Java_com_vectordb_jni_NativeIvfIndex_nativeCreateIVF": "(IIII)J",
    "Java_com_vectordb_jni_NativeIvfIndex_nativeTrain": "(JI[F)V",
    "Java_com_vectordb_jni_NativeLshIndex_nativeCreateLSH": "(IIII)J",
    "Java_com_vectordb_jni_NativeAnnoyIndex_nativeCreateAnnoy": "(III)J",
    "Java_com_vectordb_jni_NativeAnnoyIndex_nativeBuild": "(J)V",
}
NEW_NATIVES = {
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeCreateHNSWPQ": "(IIIIIIII)J",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeTrain": "(JI[F)V",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeBuild": "(J)V",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeSize": "(J)I",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeRemove": "(JI)Z",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeSave": "(JLjava/nio/ByteBuffer;)V",
    "Java_com_vectordb_jni_NativeHnswPqIndex_nativeLoad": "(JLjava/nio/ByteBuffer;)V",
    "Java_com_vectordb_jni_NativeFlatIndex_nativeCreateFlat": "(III)J",
}
JTYPE = {"int": "I", "long": "J", "float": "F", "boolean": "Z", "void": "V", "ByteBuffer": "Ljava/nio/ByteBuffer;"}


def _compile_object(tmp_path):
    obj = str(tmp_path / "bridge.o")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-fPIC", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "tests", "jni_stub"),
                           "-I" + BRIDGE, "-c", os.path.join(BRIDGE, "VectorDBJNI_b200.cpp"), "-o", obj])
    return obj


def test_bridge_compiles_and_exports_every_reference_native_with_c_linkage(tmp_path):
    obj = _compile_object(tmp_path)
    out = subprocess.check_output(["nm", "--defined-only", obj], text=True)
    defined = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for name in list(REFERENCE_NATIVES) + list(NEW_NATIVES):
        assert name in defined, name + " is not exported with C linkage"
    # nothing JNI-facing may be C++-mangled (the reference's own bridge would be: it includes none of its javah headers)
    assert not [s for s in defined if s.startswith("_Z") and "Java_com_vectordb" in s]
    # and nothing but the JNI entry points is exported by the bridge
    assert {s for s in defined if not s.startswith("_Z")} == set(REFERENCE_NATIVES) | set(NEW_NATIVES)


def test_reference_javah_headers_agree_with_the_committed_table():
    """runs only where /root/reference exists (this container): the literal table above IS the reference's javah output"""
    hdr_dir = "/root/reference/native/jni_headers"
    if not os.path.isdir(hdr_dir):
        pytest.skip("reference not present on this machine")
    found = {}
    for f in sorted(os.listdir(hdr_dir)):
        text = open(os.path.join(hdr_dir, f)).read()
        for sig, name in re.findall(r"Signature:\s*(\S+)\s*\*/\s*JNIEXPORT\s+\w+\s+JNICALL\s+(Java_\w+)", text):
            found[name] = sig
    assert found == REFERENCE_NATIVES


def _java_natives(path, cls):
    src = open(path).read()
    out = {}
    for ret, name, args in re.findall(r"native\s+(\w+)\s+(\w+)\s*\(([^)]*)\)\s*;", src):
        desc = "("
        for a in [a.strip() for a in args.replace("\n", " ").split(",") if a.strip()]:
            t = a.split()[0]
            desc += "[" + JTYPE[t[:-2]] if t.endswith("[]") else JTYPE[t]
        out["Java_com_vectordb_jni_%s_%s" % (cls, name)] = desc + ")" + JTYPE[ret]
    return out


def test_java_classes_declare_exactly_the_natives_of_their_headers():
    jdir = os.path.join(BRIDGE, "java", "com", "vectordb", "jni")
    got = {}
    got.update(_java_natives(os.path.join(jdir, "NativeHnswPqIndex.java"), "NativeHnswPqIndex"))
    got.update(_java_natives(os.path.join(jdir, "NativeFlatIndex.java"), "NativeFlatIndex"))
    assert got == NEW_NATIVES
    for cls in ("NativeHnswPqIndex", "NativeFlatIndex"):
        hdr = open(os.path.join(BRIDGE, "jni_headers", "com_vectordb_jni_%s.h" % cls)).read()
        listed = dict((n, s) for s, n in re.findall(r"Signature:\s*(\S+)\s*\*/\s*JNIEXPORT\s+\w+\s+JNICALL\s+(Java_\w+)", hdr))
        assert listed == {k: v for k, v in NEW_NATIVES.items() if "_%s_" % cls in k}
    src = open(os.path.join(jdir, "NativeHnswPqIndex.java")).read()
    assert "extends NativeIndex" in src and "package com.vectordb.jni;" in src


def test_stand_in_jni_header_has_the_specified_function_table_slots(tmp_path):
    so = str(tmp_path / "mock.so")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "tests", "jni_stub"),
                           os.path.join(ROOT, "tests", "jni_stub", "mock_env.cpp"), "-o", so])
    L = C.CDLL(so)
    L.mock_slot_of.argtypes = [C.c_char_p]
    spec = {"FindClass": 6, "ThrowNew": 14, "GetArrayLength": 171, "GetIntArrayElements": 187, "GetFloatArrayElements": 189,
            "ReleaseIntArrayElements": 195, "ReleaseFloatArrayElements": 197, "ExceptionCheck": 228,
            "GetDirectBufferAddress": 230, "GetDirectBufferCapacity": 231}
    for name, slot in spec.items():
        assert L.mock_slot_of(name.encode()) == slot, name
