// native_bridge/VectorDBJNI_b200.cpp — the JNI bridge of the B200 library: every `Java_com_vectordb_jni_*` entry point the
// reference's bridge defines (/root/reference/native/bridge/VectorDBJNI.cpp:39-200, declared in
// /root/reference/native/jni_headers/com_vectordb_jni_{NativeIndex,NativeHnswIndex,NativePqIndex}.h:15-48) re-expressed over
// the C ABI of include/vdb.h, plus the class the reference lacks (SURVEY.md §0 finding 4 / §8b "missing piece"):
// com.vectordb.jni.NativeHnswPqIndex (native_bridge/java/..., header native_bridge/jni_headers/).
//
// Differences from the reference bridge, on purpose:
//   * the declarations ARE included, so the symbols get C linkage (the reference's bridge includes none of its javah headers:
//     its functions would be C++-mangled and the JVM could not resolve them, VectorDBJNI.cpp:1-12);
//   * nothing throws across the JNI frame: the C ABI returns status codes, which become Java exceptions through ThrowNew
//     (the reference lets std::runtime_error escape from nativeAdd/nativeTrain, VectorDBJNI.cpp:67-75,132-139);
//   * result slots without a hit are id = -1 / distance = +inf, which NativeIndex.searchBatch already skips
//     (jni/NativeIndex.java:222);
//   * IVF / LSH / Annoy are not part of this library (out of scope): their nativeCreate* return 0, the
//     reference's "creation failed" value (VectorDBJNI.cpp:48-50).
// Handle table: jlong -> vdb_index*, monotonically increasing from 1, one global mutex (VectorDBJNI.cpp:13-36).
//
// Build: `make -C native_bridge JAVA_HOME=...` links it with libvectordb.so's objects into the one library the Java side
// loads (NativeLoader.java:16). On this image there is no JDK: the file is compiled against tests/jni_stub/jni.h and driven
// through a mock JNIEnv by the tests — it has never been loaded by a JVM.
#include <jni.h>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>

#include "../include/vdb.h"
#include "vectordb_jni_exports.h"                              // the entry points the reference's Java classes declare
#include "jni_headers/com_vectordb_jni_NativeHnswPqIndex.h"    // new classes (javah layout)
#include "jni_headers/com_vectordb_jni_NativeFlatIndex.h"

namespace {
struct Entry { vdb_index* idx; int dim; };
std::unordered_map<jlong, Entry> g_indices;
std::mutex g_mutex;
jlong g_next = 1;

jlong register_index(vdb_index* idx, int dim) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const jlong h = g_next++;
    g_indices[h] = Entry{idx, dim};
    return h;
}
bool lookup(jlong h, Entry* out) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_indices.find(h);
    if (it == g_indices.end()) return false;
    *out = it->second;
    return true;
}
// status code -> Java exception (what the pure-Java indexes throw: IllegalArgumentException for bad arguments,
// IllegalStateException before training / when full, RuntimeException otherwise)
void throw_for(JNIEnv* env, int rc) {
    if (rc == VDB_OK || env->ExceptionCheck()) return;
    const char* cls = "java/lang/RuntimeException";
    if (rc == VDB_ERR_INVALID) cls = "java/lang/IllegalArgumentException";
    else if (rc == VDB_ERR_UNTRAINED || rc == VDB_ERR_FULL) cls = "java/lang/IllegalStateException";
    jclass c = env->FindClass(cls);
    if (c) env->ThrowNew(c, vdb_last_error());
}
// devices for a new index: VECTORDB_DEVICE (ordinal, default 0) and VECTORDB_DEVICES (how many GPUs of the box one index may
// spread its rows over, default 1) — the Java constructors keep the reference's signatures
void apply_env(vdb_config* c) {
    if (const char* d = getenv("VECTORDB_DEVICE")) c->device = atoi(d);
    if (const char* n = getenv("VECTORDB_DEVICES")) c->n_devices = atoi(n);
}
jlong create(vdb_config* c) {
    apply_env(c);
    vdb_index* idx = nullptr;
    if (vdb_create(c, &idx) != VDB_OK || !idx) return 0;       // 0 = creation failed (VectorDBJNI.cpp:48-50)
    return register_index(idx, c->dim);
}
void train_common(JNIEnv* env, jlong handle, jint nSamples, jfloatArray samples) {
    Entry e;
    if (!lookup(handle, &e)) return;                 // unknown handle: silent no-op (VectorDBJNI.cpp:69-70)
    if (!samples || nSamples <= 0 || (jlong)env->GetArrayLength(samples) < (jlong)nSamples * e.dim) {
        jclass c = env->FindClass("java/lang/IllegalArgumentException");
        if (c) env->ThrowNew(c, "Invalid training samples");     // PQIndex.cpp:30-32
        return;
    }
    jfloat* data = env->GetFloatArrayElements(samples, nullptr);
    if (!data) return;
    const int rc = vdb_train(e.idx, data, nSamples);
    env->ReleaseFloatArrayElements(samples, data, JNI_ABORT);
    throw_for(env, rc);
}
}  // namespace

// ---- constructors ----------------------------------------------------------------------------------------------------
// NativeHnswIndex.nativeCreateHNSW (IIIII)J — VectorDBJNI.cpp:39-51. The class VectorDatabase runs for IndexType.HNSW without
// compression is Java HnswIndex (cosine on normalised vectors, index/HnswIndex.java): that is what VDB_INDEX_HNSW reproduces.
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeHnswIndex_nativeCreateHNSW(JNIEnv*, jclass, jint dimension, jint maxElements,
                                                                                jint M, jint efConstruction, jint ef) {
    vdb_config c;
    if (vdb_default_config(VDB_INDEX_HNSW, dimension, maxElements, &c) != VDB_OK) return 0;
    if (M > 0) c.hnsw_m = M;
    if (efConstruction > 0) c.ef_construction = efConstruction;
    if (ef > 0) c.ef_search = ef;
    return create(&c);
}
// NativePqIndex.nativeCreatePQ (IIII)J — VectorDBJNI.cpp:53-65 (native PQIndex: LUT + ADC scan + (dist,id) top-k)
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativePqIndex_nativeCreatePQ(JNIEnv*, jclass, jint dimension, jint maxElements, jint M,
                                                                            jint nBits) {
    vdb_config c;
    if (vdb_default_config(VDB_INDEX_PQ, dimension, maxElements, &c) != VDB_OK) return 0;
    c.pq_m = M; c.pq_bits = nBits;
    return create(&c);
}
// NativeHnswPqIndex.nativeCreateHNSWPQ (IIIIIIII)J — new: Java HnswPqIndex semantics (index/HnswPqIndex.java:67-108)
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeCreateHNSWPQ(JNIEnv*, jclass, jint dimension, jint maxElements,
                                                                                    jint M, jint efConstruction, jint ef, jint pqM,
                                                                                    jint pqBits, jint pqIters) {
    vdb_config c;
    if (vdb_default_config(VDB_INDEX_HNSW_PQ, dimension, maxElements, &c) != VDB_OK) return 0;
    if (M > 0) c.hnsw_m = M;
    if (efConstruction > 0) c.ef_construction = efConstruction;
    if (ef > 0) c.ef_search = ef;
    if (pqM > 0) c.pq_m = pqM;                       // 0 = CompressionConfig.recommendedConfig(dim)
    if (pqBits > 0) c.pq_bits = pqBits;
    if (pqIters > 0) c.pq_iters = pqIters;
    return create(&c);
}
// NativeFlatIndex.nativeCreateFlat (III)J — new: exact brute force on the tensor cores (ground truth of
// RecallOptimizationTest.java:151-164); metric 0 = squared L2, 1 = cosine
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeFlatIndex_nativeCreateFlat(JNIEnv*, jclass, jint dimension, jint maxElements,
                                                                                jint metric) {
    vdb_config c;
    if (vdb_default_config(VDB_INDEX_FLAT, dimension, maxElements, &c) != VDB_OK) return 0;
    c.metric = metric;
    return create(&c);
}

// ---- training / build / CRUD of the new class ---------------------------------------------------------------------------
// NativePqIndex.nativeTrain (JI[F)V — VectorDBJNI.cpp:67-75
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativePqIndex_nativeTrain(JNIEnv* env, jobject, jlong handle, jint nSamples,
                                                                        jfloatArray samples) {
    train_common(env, handle, nSamples, samples);
}
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeTrain(JNIEnv* env, jobject, jlong handle, jint nSamples,
                                                                            jfloatArray samples) {
    train_common(env, handle, nSamples, samples);
}
// NativeHnswPqIndex.nativeBuild (J)V — VectorIndex.buildIndex / VectorDatabase.rebuildIndex (HnswPqIndex.java:676-718); pattern of
// NativeAnnoyIndex.nativeBuild (VectorDBJNI.cpp:123-129)
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeBuild(JNIEnv* env, jobject, jlong handle) {
    Entry e;
    if (!lookup(handle, &e)) return;
    throw_for(env, vdb_build(e.idx));
}
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeSize(JNIEnv*, jobject, jlong handle) {
    Entry e;
    if (!lookup(handle, &e)) return -1;              // NativeIndex.size() answers -1 when unknown (NativeIndex.java:79-82)
    return (jint)vdb_size(e.idx);
}
// VectorIndex.removeVector (HnswPqIndex.java:542-596): true when the id was present
JNIEXPORT jboolean JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeRemove(JNIEnv* env, jobject, jlong handle, jint id) {
    Entry e;
    if (!lookup(handle, &e)) return 0;
    int32_t removed = 0;
    throw_for(env, vdb_remove(e.idx, id, &removed));
    return removed ? 1 : 0;
}
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeSave(JNIEnv* env, jobject, jlong handle, jobject pathBuf) {
    Entry e;
    if (!lookup(handle, &e)) return;
    const char* p = pathBuf ? static_cast<const char*>(env->GetDirectBufferAddress(pathBuf)) : nullptr;   // NUL-terminated UTF-8
    if (!p) { throw_for(env, VDB_ERR_INVALID); return; }
    throw_for(env, vdb_save(e.idx, p));
}
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeLoad(JNIEnv* env, jobject, jlong handle, jobject pathBuf) {
    Entry e;
    if (!lookup(handle, &e)) return;
    const char* p = pathBuf ? static_cast<const char*>(env->GetDirectBufferAddress(pathBuf)) : nullptr;
    if (!p) { throw_for(env, VDB_ERR_INVALID); return; }
    throw_for(env, vdb_load(e.idx, p));
}

// ---- NativeIndex: the five natives every index shares (jni/NativeIndex.java:137-146) ----------------------------------------
// nativeAdd (JI[F)V — VectorDBJNI.cpp:132-139
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeAdd(JNIEnv* env, jobject, jlong handle, jint id, jfloatArray vector) {
    Entry e;
    if (!lookup(handle, &e)) return;
    if (!vector || env->GetArrayLength(vector) != e.dim) { throw_for(env, VDB_ERR_INVALID); return; }
    jfloat* data = env->GetFloatArrayElements(vector, nullptr);
    if (!data) return;
    const int32_t vid = id;
    const int rc = vdb_add_batch(e.idx, &vid, data, 1);
    env->ReleaseFloatArrayElements(vector, data, JNI_ABORT);
    throw_for(env, rc);
}
// nativeSearch (J[FI[I[F)I — VectorDBJNI.cpp:141-159: returns the number of hits, ids/distances filled in ascending distance
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeIndex_nativeSearch(JNIEnv* env, jobject, jlong handle, jfloatArray query, jint k,
                                                                       jintArray resultIds, jfloatArray resultDistances) {
    Entry e;
    if (!lookup(handle, &e)) return 0;
    if (!query || !resultIds || !resultDistances || k <= 0 || env->GetArrayLength(query) != e.dim ||
        env->GetArrayLength(resultIds) < k || env->GetArrayLength(resultDistances) < k) {
        throw_for(env, VDB_ERR_INVALID);
        return 0;
    }
    jfloat* q = env->GetFloatArrayElements(query, nullptr);
    jint* ids = env->GetIntArrayElements(resultIds, nullptr);
    jfloat* d = env->GetFloatArrayElements(resultDistances, nullptr);
    int rc = VDB_ERR_INVALID;
    jint count = 0;
    if (q && ids && d) {
        rc = vdb_search_batch(e.idx, q, 1, k, reinterpret_cast<int32_t*>(ids), d);
        if (rc == VDB_OK) while (count < k && ids[count] >= 0) count++;
    }
    if (q) env->ReleaseFloatArrayElements(query, q, JNI_ABORT);
    if (ids) env->ReleaseIntArrayElements(resultIds, ids, 0);
    if (d) env->ReleaseFloatArrayElements(resultDistances, d, 0);
    throw_for(env, rc);
    return count;
}
// nativeDestroy (J)V — VectorDBJNI.cpp:161-164
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeDestroy(JNIEnv*, jobject, jlong handle) {
    Entry e;
    {
        std::lock_guard<std::mutex> lk(g_mutex);
        auto it = g_indices.find(handle);
        if (it == g_indices.end()) return;
        e = it->second;
        g_indices.erase(it);
    }
    vdb_destroy(e.idx);
}
// nativeAddBatch (JLjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;II)V — VectorDBJNI.cpp:167-182: direct buffers in native byte
// order, ids[count] i32 and vectors[count][dimension] f32, borrowed for the call (no copy on the Java side)
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeAddBatch(JNIEnv* env, jobject, jlong handle, jobject idsBuffer,
                                                                         jobject vectorsBuffer, jint count, jint dimension) {
    Entry e;
    if (!lookup(handle, &e)) return;
    const int32_t* ids = idsBuffer ? static_cast<const int32_t*>(env->GetDirectBufferAddress(idsBuffer)) : nullptr;
    const float* vecs = vectorsBuffer ? static_cast<const float*>(env->GetDirectBufferAddress(vectorsBuffer)) : nullptr;
    if (!ids || !vecs || count < 0 || dimension != e.dim || env->GetDirectBufferCapacity(idsBuffer) < (jlong)count * 4 ||
        env->GetDirectBufferCapacity(vectorsBuffer) < (jlong)count * dimension * 4) {
        throw_for(env, VDB_ERR_INVALID);
        return;
    }
    throw_for(env, vdb_add_batch(e.idx, ids, vecs, count));
}
// nativeSearchBatch (JLjava/nio/ByteBuffer;IIILjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;)I — VectorDBJNI.cpp:184-200: queries
// [nQueries][dimension] f32 -> ids [nQueries][k] i32, distances [nQueries][k] f32; returns nQueries. ONE call = one batched
// GPU search (all GPUs of the index when it was created with VECTORDB_DEVICES > 1).
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeIndex_nativeSearchBatch(JNIEnv* env, jobject, jlong handle, jobject queriesBuffer,
                                                                            jint nQueries, jint k, jint dimension,
                                                                            jobject resultIdsBuffer, jobject resultDistancesBuffer) {
    Entry e;
    if (!lookup(handle, &e)) return 0;
    const float* q = queriesBuffer ? static_cast<const float*>(env->GetDirectBufferAddress(queriesBuffer)) : nullptr;
    int32_t* ids = resultIdsBuffer ? static_cast<int32_t*>(env->GetDirectBufferAddress(resultIdsBuffer)) : nullptr;
    float* d = resultDistancesBuffer ? static_cast<float*>(env->GetDirectBufferAddress(resultDistancesBuffer)) : nullptr;
    if (!q || !ids || !d || nQueries < 0 || k <= 0 || dimension != e.dim ||
        env->GetDirectBufferCapacity(queriesBuffer) < (jlong)nQueries * dimension * 4 ||
        env->GetDirectBufferCapacity(resultIdsBuffer) < (jlong)nQueries * k * 4 ||
        env->GetDirectBufferCapacity(resultDistancesBuffer) < (jlong)nQueries * k * 4) {
        throw_for(env, VDB_ERR_INVALID);
        return 0;
    }
    const int rc = vdb_search_batch(e.idx, q, nQueries, k, ids, d);
    throw_for(env, rc);
    return rc == VDB_OK ? nQueries : 0;
}

// ---- families outside this library: creation fails the reference's way (handle 0) ----------------
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeIvfIndex_nativeCreateIVF(JNIEnv*, jclass, jint, jint, jint, jint) { return 0; }
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIvfIndex_nativeTrain(JNIEnv*, jobject, jlong, jint, jfloatArray) {}
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeLshIndex_nativeCreateLSH(JNIEnv*, jclass, jint, jint, jint, jint) { return 0; }
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeAnnoyIndex_nativeCreateAnnoy(JNIEnv*, jclass, jint, jint, jint) { return 0; }
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeAnnoyIndex_nativeBuild(JNIEnv*, jobject, jlong) {}
