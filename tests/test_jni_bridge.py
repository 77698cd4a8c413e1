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


# ---------------------------------------------------------------------------------------------------------------------
# GPU: the bridge driven through the mock JNIEnv (Java-style arrays / direct buffers), against the C ABI's own answers
# ---------------------------------------------------------------------------------------------------------------------
class Bridge:
    def __init__(self):
        import __graft_entry__ as ge
        ge.build_product()
        subprocess.check_call(["make", "-C", BRIDGE, "stub"], stdout=subprocess.DEVNULL)
        self.L = L = C.CDLL(os.path.join(BRIDGE, "libvectordb_jni_stub.so"))
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.mock_env.restype = vp
        L.mock_new_float_array.argtypes = [vp, C.c_int]; L.mock_new_float_array.restype = vp
        L.mock_new_int_array.argtypes = [C.c_int]; L.mock_new_int_array.restype = vp
        L.mock_array_read.argtypes = [vp, vp]
        L.mock_free_array.argtypes = [vp]
        L.mock_new_direct_buffer.argtypes = [vp, i64]; L.mock_new_direct_buffer.restype = vp
        L.mock_free_buffer.argtypes = [vp]
        L.mock_exception_class.restype = C.c_char_p
        L.mock_exception_message.restype = C.c_char_p
        P = "Java_com_vectordb_jni_"
        sig = {"NativeHnswPqIndex_nativeCreateHNSWPQ": ([vp, vp] + [i32] * 8, i64), "NativePqIndex_nativeCreatePQ": ([vp, vp] + [i32] * 4, i64),
               "NativeHnswIndex_nativeCreateHNSW": ([vp, vp] + [i32] * 5, i64), "NativeFlatIndex_nativeCreateFlat": ([vp, vp] + [i32] * 3, i64),
               "NativeIvfIndex_nativeCreateIVF": ([vp, vp] + [i32] * 4, i64),
               "NativePqIndex_nativeTrain": ([vp, vp, i64, i32, vp], None), "NativeHnswPqIndex_nativeBuild": ([vp, vp, i64], None),
               "NativeHnswPqIndex_nativeSize": ([vp, vp, i64], i32), "NativeHnswPqIndex_nativeRemove": ([vp, vp, i64, i32], C.c_uint8),
               "NativeIndex_nativeAdd": ([vp, vp, i64, i32, vp], None), "NativeIndex_nativeSearch": ([vp, vp, i64, vp, i32, vp, vp], i32),
               "NativeIndex_nativeDestroy": ([vp, vp, i64], None), "NativeIndex_nativeAddBatch": ([vp, vp, i64, vp, vp, i32, i32], None),
               "NativeIndex_nativeSearchBatch": ([vp, vp, i64, vp, i32, i32, i32, vp, vp], i32)}
        for name, (args, res) in sig.items():
            f = getattr(L, P + name); f.argtypes = args; f.restype = res
            setattr(self, name.split("_")[1], f) if name.split("_")[1] not in ("nativeTrain",) else None
        self.nativeTrainPQ = getattr(L, P + "NativePqIndex_nativeTrain")
        self.env = L.mock_env()

    def farray(self, a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        return self.L.mock_new_float_array(a.ctypes.data, a.size)

    def buf(self, a):
        return self.L.mock_new_direct_buffer(a.ctypes.data, a.nbytes)

    def exception(self):
        if not self.L.mock_exception_pending():
            return None
        e = (self.L.mock_exception_class().decode(), self.L.mock_exception_message().decode())
        self.L.mock_exception_clear()
        return e


@pytest.mark.gpu
def test_bridge_through_mock_jnienv_matches_c_abi(po):
    import vector_db_b200 as v
    b = Bridge()
    n, dim, M, k, nq = 2500, 32, 4, 10, 40
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    # reference answer through the C ABI (ctypes mirror)
    g = v.NativeHnswPqIndex(dim, 5000, pqM=M, level_seed=1234)
    g.addBatch(np.arange(n, dtype=np.int32), X)
    want_i, want_d = g.searchBatch(Q, k)
    # same index through the bridge: NativeIndex.addBatch / searchBatch with direct ByteBuffers (NativeIndex.java:153-230)
    h = b.nativeCreateHNSWPQ(b.env, None, dim, 5000, 32, 64, 64, M, 8, 25)
    assert h == 1                                                   # handles count up from 1 (VectorDBJNI.cpp:15)
    ids = np.arange(n, dtype=np.int32)
    ib, vb = b.buf(ids), b.buf(X)
    b.nativeAddBatch(b.env, None, h, ib, vb, n, dim)
    assert b.exception() is None and b.nativeSize(b.env, None, h) == n
    out_i = np.zeros((nq, k), dtype=np.int32); out_d = np.zeros((nq, k), dtype=np.float32)
    qb, oib, odb = b.buf(Q), b.buf(out_i), b.buf(out_d)
    assert b.nativeSearchBatch(b.env, None, h, qb, nq, k, dim, oib, odb) == nq and b.exception() is None
    assert np.array_equal(out_i, want_i) and np.array_equal(out_d.view(np.uint32), want_d.view(np.uint32))
    # NativeIndex.search(float[], k): Java arrays, copy-back on release, count = hits
    qa, ia, da = b.farray(Q[0]), b.L.mock_new_int_array(k), b.L.mock_new_float_array(None, k)
    assert b.nativeSearch(b.env, None, h, qa, k, ia, da) == k
    ri = np.zeros(k, dtype=np.int32); rd = np.zeros(k, dtype=np.float32)
    b.L.mock_array_read(ia, ri.ctypes.data); b.L.mock_array_read(da, rd.ctypes.data)
    assert np.array_equal(ri, want_i[0]) and np.array_equal(rd.view(np.uint32), want_d[0].view(np.uint32))
    assert b.L.mock_outstanding_elements() == 0                     # every Get...ArrayElements was released
    # nativeAdd(id, float[]) + removeVector + size
    va = b.farray(Q[1])
    b.nativeAdd(b.env, None, h, 9000, va)
    assert b.exception() is None and b.nativeSize(b.env, None, h) == n + 1
    assert b.nativeRemove(b.env, None, h, 9000) == 1 and b.nativeRemove(b.env, None, h, 9000) == 0
    assert b.nativeSize(b.env, None, h) == n
    # errors become Java exceptions, never C++ exceptions across the frame
    short = b.farray(Q[0][:5])
    b.nativeAdd(b.env, None, h, 9001, short)
    assert b.exception()[0] == "java/lang/IllegalArgumentException"
    assert b.nativeSearchBatch(b.env, None, h, qb, nq, k, dim + 1, oib, odb) == 0
    assert b.exception()[0] == "java/lang/IllegalArgumentException"
    small = np.zeros(4, dtype=np.int32); sb = b.buf(small)
    assert b.nativeSearchBatch(b.env, None, h, qb, nq, k, dim, sb, odb) == 0          # result buffer too small: refused, not overrun
    assert b.exception()[0] == "java/lang/IllegalArgumentException"
    # PQ index: add before train -> IllegalStateException (the reference would let std::runtime_error escape, PQIndex.cpp:124-126)
    p = b.nativeCreatePQ(b.env, None, dim, 5000, M, 8)
    assert p == 2
    b.nativeAddBatch(b.env, None, p, ib, vb, n, dim)
    assert b.exception()[0] == "java/lang/IllegalStateException"
    ta = b.farray(X[:600])
    b.nativeTrainPQ(b.env, None, p, 600, ta)
    assert b.exception() is None
    b.nativeAddBatch(b.env, None, p, ib, vb, n, dim)
    assert b.exception() is None
    assert b.nativeSearchBatch(b.env, None, p, qb, nq, k, dim, oib, odb) == nq
    pq = v.NativePqIndex(dim, 5000, M=M)
    pq.train(X[:600]); pq.addBatch(ids, X)
    pi, pd = pq.searchBatch(Q, k)
    assert np.array_equal(out_i, pi) and np.array_equal(out_d.view(np.uint32), pd.view(np.uint32))
    # out-of-scope families report creation failure the reference's way; unknown handles are silent no-ops
    assert b.nativeCreateIVF(b.env, None, dim, 100, 4, 2) == 0
    assert b.nativeSearchBatch(b.env, None, 777, qb, nq, k, dim, oib, odb) == 0 and b.exception() is None
    b.nativeDestroy(b.env, None, h); b.nativeDestroy(b.env, None, p); b.nativeDestroy(b.env, None, h)
    assert b.nativeSize(b.env, None, h) == -1
    assert b.L.mock_outstanding_elements() == 0
    g.close(); pq.close()


@pytest.mark.gpu
def test_bridge_spreads_one_index_over_devices_from_the_environment(po, monkeypatch):
    """VECTORDB_DEVICES = vdb_config.n_devices: the Java constructors keep the reference's signatures, the bridge reads the device
    count at creation; ONE nativeSearchBatch then fans out over the shards and returns the single-index answer bit for bit
    (on a one-GPU box the shards share the device: VDB_MULTI_SAME_DEVICE)."""
    import vector_db_b200 as v
    b = Bridge()
    n, dim, k, nq = 6000, 48, 12, 33
    X, Q = po.synth(42, n, dim), po.synth(43, nq, dim)
    ref = v.NativeFlatIndex(dim, n)
    ids = np.arange(n, dtype=np.int32)
    ref.addBatch(ids, X)
    want_i, want_d = ref.searchBatch(Q, k)
    ref.close()
    monkeypatch.setenv("VDB_MULTI_SAME_DEVICE", "1")
    monkeypatch.setenv("VECTORDB_DEVICES", "3")
    h = b.nativeCreateFlat(b.env, None, dim, n, 0)
    monkeypatch.delenv("VECTORDB_DEVICES")
    assert h > 0
    ib, vb = b.buf(ids), b.buf(X)
    b.nativeAddBatch(b.env, None, h, ib, vb, n, dim)
    assert b.exception() is None and b.nativeSize(b.env, None, h) == n
    out_i = np.zeros((nq, k), dtype=np.int32); out_d = np.zeros((nq, k), dtype=np.float32)
    qb, oib, odb = b.buf(Q), b.buf(out_i), b.buf(out_d)
    assert b.nativeSearchBatch(b.env, None, h, qb, nq, k, dim, oib, odb) == nq and b.exception() is None
    assert np.array_equal(out_i, want_i) and np.array_equal(out_d.view(np.uint32), want_d.view(np.uint32))
    # a graph index refuses more than one device (replicas): creation fails the reference's way (handle 0)
    monkeypatch.setenv("VECTORDB_DEVICES", "2")
    assert b.nativeCreateHNSWPQ(b.env, None, dim, 5000, 32, 64, 64, 6, 8, 25) == 0
    b.nativeDestroy(b.env, None, h)
