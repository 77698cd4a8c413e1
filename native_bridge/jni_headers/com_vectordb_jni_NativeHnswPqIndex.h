/* javah-style header for the NEW class com.vectordb.jni.NativeHnswPqIndex (native_bridge/java/com/vectordb/jni/
 * NativeHnswPqIndex.java). Written by hand in javah's layout: no JDK exists on the build image. */
#include <jni.h>
/* Header for class com_vectordb_jni_NativeHnswPqIndex */

#ifndef _Included_com_vectordb_jni_NativeHnswPqIndex
#define _Included_com_vectordb_jni_NativeHnswPqIndex
#ifdef __cplusplus
extern "C" {
#endif
/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeCreateHNSWPQ
 * Signature: (IIIIIIII)J
 */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeCreateHNSWPQ
  (JNIEnv *, jclass, jint, jint, jint, jint, jint, jint, jint, jint);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeTrain
 * Signature: (JI[F)V
 */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeTrain
  (JNIEnv *, jobject, jlong, jint, jfloatArray);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeBuild
 * Signature: (J)V
 */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeBuild
  (JNIEnv *, jobject, jlong);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeSize
 * Signature: (J)I
 */
JNIEXPORT jint JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeSize
  (JNIEnv *, jobject, jlong);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeRemove
 * Signature: (JI)Z
 */
JNIEXPORT jboolean JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeRemove
  (JNIEnv *, jobject, jlong, jint);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeSave
 * Signature: (JLjava/nio/ByteBuffer;)V
 */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeSave
  (JNIEnv *, jobject, jlong, jobject);

/*
 * Class:     com_vectordb_jni_NativeHnswPqIndex
 * Method:    nativeLoad
 * Signature: (JLjava/nio/ByteBuffer;)V
 */
JNIEXPORT void JNICALL Java_com_vectordb_jni_NativeHnswPqIndex_nativeLoad
  (JNIEnv *, jobject, jlong, jobject);

#ifdef __cplusplus
}
#endif
#endif
