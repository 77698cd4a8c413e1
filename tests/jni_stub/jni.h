/* tests/jni_stub/jni.h — declarations-only STAND-IN for the JDK's <jni.h> (this image has no JDK).
 *
 * It exists so that native_bridge/VectorDBJNI_b200.cpp can be COMPILED and its exported symbol names checked here, and so
 * that tests can drive the shim through a mock JNIEnv (tests/jni_stub/mock_env.cpp). It is NOT used for a real build: the
 * Makefile takes $JAVA_HOME/include/jni.h when a JDK exists. The function-table slots below follow the JNI specification's
 * numbering (slot 6 FindClass, 14 ThrowNew, 171 GetArrayLength, 187/189 Get{Int,Float}ArrayElements,
 * 195/197 Release{Int,Float}ArrayElements, 228 ExceptionCheck, 230/231 GetDirectBuffer{Address,Capacity}) so the layout is
 * the JVM's, but nothing here has ever been loaded by a JVM — INTEGRATION.md says so.
 */
#ifndef VDB_TEST_JNI_STUB_H_
#define VDB_TEST_JNI_STUB_H_
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2
#define JNI_COMMIT 1
#define JNI_OK 0
#define JNI_VERSION_1_8 0x00010008

typedef int32_t jint;
typedef int64_t jlong;
typedef float jfloat;
typedef uint8_t jboolean;
typedef jint jsize;

#ifdef __cplusplus
class _jobject {};
class _jclass : public _jobject {};
class _jarray : public _jobject {};
class _jintArray : public _jarray {};
class _jfloatArray : public _jarray {};
typedef _jobject* jobject;
typedef _jclass* jclass;
typedef _jarray* jarray;
typedef _jintArray* jintArray;
typedef _jfloatArray* jfloatArray;
#else
typedef void* jobject;
typedef jobject jclass;
typedef jobject jarray;
typedef jarray jintArray;
typedef jarray jfloatArray;
#endif

struct JNIEnv_;
struct JavaVM_;
#ifdef __cplusplus
typedef JNIEnv_ JNIEnv;
typedef JavaVM_ JavaVM;
#else
typedef const struct JNINativeInterface_* JNIEnv;
#endif

struct JNINativeInterface_ {
    void* slot0_5[6];                                                         /* 0..5 */
    jclass (*FindClass)(JNIEnv*, const char*);                                /* 6 */
    void* slot7_13[7];                                                        /* 7..13 */
    jint (*ThrowNew)(JNIEnv*, jclass, const char*);                           /* 14 */
    void* slot15_170[156];                                                    /* 15..170 */
    jsize (*GetArrayLength)(JNIEnv*, jarray);                                 /* 171 */
    void* slot172_186[15];                                                    /* 172..186 */
    jint* (*GetIntArrayElements)(JNIEnv*, jintArray, jboolean*);              /* 187 */
    void* slot188;                                                            /* 188 */
    jfloat* (*GetFloatArrayElements)(JNIEnv*, jfloatArray, jboolean*);        /* 189 */
    void* slot190_194[5];                                                     /* 190..194 */
    void (*ReleaseIntArrayElements)(JNIEnv*, jintArray, jint*, jint);         /* 195 */
    void* slot196;                                                            /* 196 */
    void (*ReleaseFloatArrayElements)(JNIEnv*, jfloatArray, jfloat*, jint);   /* 197 */
    void* slot198_227[30];                                                    /* 198..227 */
    jboolean (*ExceptionCheck)(JNIEnv*);                                      /* 228 */
    void* slot229;                                                            /* 229 */
    void* (*GetDirectBufferAddress)(JNIEnv*, jobject);                        /* 230 */
    jlong (*GetDirectBufferCapacity)(JNIEnv*, jobject);                       /* 231 */
};

#ifdef __cplusplus
struct JNIEnv_ {
    const struct JNINativeInterface_* functions;
    jclass FindClass(const char* n) { return functions->FindClass(this, n); }
    jint ThrowNew(jclass c, const char* m) { return functions->ThrowNew(this, c, m); }
    jsize GetArrayLength(jarray a) { return functions->GetArrayLength(this, a); }
    jint* GetIntArrayElements(jintArray a, jboolean* c) { return functions->GetIntArrayElements(this, a, c); }
    jfloat* GetFloatArrayElements(jfloatArray a, jboolean* c) { return functions->GetFloatArrayElements(this, a, c); }
    void ReleaseIntArrayElements(jintArray a, jint* e, jint m) { functions->ReleaseIntArrayElements(this, a, e, m); }
    void ReleaseFloatArrayElements(jfloatArray a, jfloat* e, jint m) { functions->ReleaseFloatArrayElements(this, a, e, m); }
    jboolean ExceptionCheck() { return functions->ExceptionCheck(this); }
    void* GetDirectBufferAddress(jobject b) { return functions->GetDirectBufferAddress(this, b); }
    jlong GetDirectBufferCapacity(jobject b) { return functions->GetDirectBufferCapacity(this, b); }
};
#endif
#endif /* VDB_TEST_JNI_STUB_H_ */
