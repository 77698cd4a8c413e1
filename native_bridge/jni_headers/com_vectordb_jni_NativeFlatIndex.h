/* javah-style header for the NEW class com.vectordb.jni.NativeFlatIndex (exact brute force on the tensor cores). */
#include <jni.h>
/* Header for class com_vectordb_jni_NativeFlatIndex */

#ifndef _Included_com_vectordb_jni_NativeFlatIndex
#define _Included_com_vectordb_jni_NativeFlatIndex
#ifdef __cplusplus
extern "C" {
#endif
/*
 * Class:     com_vectordb_jni_NativeFlatIndex
 * Method:    nativeCreateFlat
 * Signature: (III)J
 */
JNIEXPORT jlong JNICALL Java_com_vectordb_jni_NativeFlatIndex_nativeCreateFlat
  (JNIEnv *, jclass, jint, jint, jint);

#ifdef __cplusplus
}
#endif
#endif
