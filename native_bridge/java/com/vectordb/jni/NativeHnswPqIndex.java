package com.vectordb.jni;

import java.nio.ByteBuffer;
import java.nio.charset.StandardCharsets;

/**
 * HNSW+PQ index on the GPU (libvectordb.so built from vector_db_b200): the class the reference lacks — its JNI bridge has no
 * entry for HNSW+PQ (native/bridge/VectorDBJNI.cpp:2-6). Shape mirrors NativePqIndex / NativeAnnoyIndex
 * (jni/NativePqIndex.java:6-25, jni/NativeAnnoyIndex.java:19-25); semantics are those of the Java HnswPqIndex that
 * VectorDatabase instantiates for IndexType.HNSW with compression on (index/HnswPqIndex.java): squared L2, implicit PQ
 * training after min(10000, maxElements/10) inserts, PQ-distance greedy descent, exact layer-0 beam, exact rerank.
 *
 * Drop-in: VectorDatabase.Builder.withIndex(new NativeHnswPqIndex(dim, maxElements, ...)) (core/VectorDatabase.java:402-405),
 * or the one-line switch in VectorDatabase.createIndex (:71-103) shown in INTEGRATION.md.
 *
 * GPU selection: environment VECTORDB_DEVICE (ordinal) and VECTORDB_DEVICES (GPUs of the box the index may use).
 */
public class NativeHnswPqIndex extends NativeIndex {

    public NativeHnswPqIndex(int dimension, int maxElements, int M, int efConstruction, int ef,
                             int pqSubspaces, int pqBits, int pqIterations) {
        super(dimension, nativeCreateHNSWPQ(dimension, maxElements, M, efConstruction, ef,
                                            pqSubspaces, pqBits, pqIterations));
        if (nativeHandle == 0) {
            throw new RuntimeException("Failed to create GPU HNSW+PQ index (no B200 device or bad configuration)");
        }
    }

    /** HnswPqIndex defaults (index/HnswPqIndex.java:27-32): M=32, efConstruction=64, ef=64; pqSubspaces=0 picks
     *  CompressionConfig.recommendedConfig(dimension). */
    public NativeHnswPqIndex(int dimension, int maxElements) {
        this(dimension, maxElements, 32, 64, 64, 0, 8, 25);
    }

    /** Explicit codebook training (optional: addVector trains implicitly like the Java class). */
    public void train(int nSamples, float[] samples) {
        nativeTrain(nativeHandle, nSamples, samples);
    }

    /** VectorIndex.buildIndex / VectorDatabase.rebuildIndex: train if needed, re-encode, rebuild the graph by ascending id. */
    @Override
    public boolean buildIndex() {
        nativeBuild(nativeHandle);
        return true;
    }

    @Override
    public int size() {
        return nativeSize(nativeHandle);
    }

    @Override
    public boolean removeVector(int id) {
        return nativeRemove(nativeHandle, id);
    }

    public void save(String path) {
        nativeSave(nativeHandle, cString(path));
    }

    public void load(String path) {
        nativeLoad(nativeHandle, cString(path));
    }

    private static ByteBuffer cString(String s) {
        byte[] b = s.getBytes(StandardCharsets.UTF_8);
        ByteBuffer buf = ByteBuffer.allocateDirect(b.length + 1);
        buf.put(b).put((byte) 0);
        return buf;
    }

    private static native long nativeCreateHNSWPQ(int dimension, int maxElements, int M, int efConstruction, int ef,
                                                  int pqSubspaces, int pqBits, int pqIterations);
    private native void nativeTrain(long handle, int nSamples, float[] samples);
    private native void nativeBuild(long handle);
    private native int nativeSize(long handle);
    private native boolean nativeRemove(long handle, int id);
    private native void nativeSave(long handle, ByteBuffer path);
    private native void nativeLoad(long handle, ByteBuffer path);
}
