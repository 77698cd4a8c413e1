package com.vectordb.jni;

/**
 * Exact brute-force search on the GPU tensor cores (bf16 tcgen05 GEMM shortlist + certified fp32 re-score): the ground
 * truth loop of benchmark/RecallOptimizationTest.java:151-164 as an index. metric 0 = squared L2 (HnswPqIndex convention),
 * 1 = cosine distance on normalised vectors (HnswIndex convention).
 */
public class NativeFlatIndex extends NativeIndex {

    public NativeFlatIndex(int dimension, int maxElements, int metric) {
        super(dimension, nativeCreateFlat(dimension, maxElements, metric));
        if (nativeHandle == 0) {
            throw new RuntimeException("Failed to create GPU flat index");
        }
    }

    public NativeFlatIndex(int dimension, int maxElements) {
        this(dimension, maxElements, 0);
    }

    private static native long nativeCreateFlat(int dimension, int maxElements, int metric);
}
