/* native_bridge/vectordb_jni_exports.h — C-linkage prototypes of the JNI entry points the REFERENCE's Java classes already
 * declare (jni/NativeIndex.java:137-146, NativeHnswIndex.java:24, NativePqIndex.java:23-24, NativeIvfIndex / NativeLshIndex /
 * NativeAnnoyIndex) and its javah headers list (/root/reference/native/jni_headers/com_vectordb_jni_*.h:15-48). A real build
 * may include those generated headers instead; this file states the same names and JNI signatures in one place so that the
 * bridge gets `extern "C"` (the reference's VectorDBJNI.cpp includes none of them and would export mangled names).
 * Signature strings are the JVM's method descriptors. */
#ifndef VECTORDB_JNI_EXPORTS_H_
#define VECTORDB_JNI_EXPORTS_H_
#include <jni.h>
#ifdef __cplusplus
extern "C" {
#endif
/* com.vectordb.jni.NativeIndex */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeAdd(JNIEnv*, jobject, jlong, jint, jfloatArray);                       /* (JI[F)V */
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeIndex_nativeSearch(JNIEnv*, jobject, jlong, jfloatArray, jint, jintArray, jfloatArray); /* (J[FI[I[F)I */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeDestroy(JNIEnv*, jobject, jlong);                                        /* (J)V */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIndex_nativeAddBatch(JNIEnv*, jobject, jlong, jobject, jobject, jint, jint);        /* (JLjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;II)V */
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeIndex_nativeSearchBatch(JNIEnv*, jobject, jlong, jobject, jint, jint, jint, jobject, jobject); /* (JLjava/nio/ByteBuffer;IIILjava/nio/ByteBuffer;Ljava/nio/ByteBuffer;)I */
/* com.vectordb.jni.NativeHnswIndex */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeHnswIndex_nativeCreateHNSW(JNIEnv*, jclass, jint, jint, jint, jint, jint);         /* (IIIII)J */
/* com.vectordb.jni.NativePqIndex */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativePqIndex_nativeCreatePQ(JNIEnv*, jclass, jint, jint, jint, jint);                   /* (IIII)J */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativePqIndex_nativeTrain(JNIEnv*, jobject, jlong, jint, jfloatArray);                    /* (JI[F)V */
/* families outside this library: exported so the classes link, creation reports failure (handle 0) */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeIvfIndex_nativeCreateIVF(JNIEnv*, jclass, jint, jint, jint, jint);                 /* (IIII)J */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeIvfIndex_nativeTrain(JNIEnv*, jobject, jlong, jint, jfloatArray);                   /* (JI[F)V */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeLshIndex_nativeCreateLSH(JNIEnv*, jclass, jint, jint, jint, jint);                 /* (IIII)J */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeAnnoyIndex_nativeCreateAnnoy(JNIEnv*, jclass, jint, jint, jint);                   /* (III)J */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeAnnoyIndex_nativeBuild(JNIEnv*, jobject, jlong);                                     /* (J)V */
#ifdef __cplusplus
}
#endif
#endif
