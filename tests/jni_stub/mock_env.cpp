// tests/jni_stub/mock_env.cpp — TEST INFRASTRUCTURE: a mock JNIEnv that lets the tests drive native_bridge/VectorDBJNI_b200.cpp
// without a JVM. Arrays are heap objects; Get<T>ArrayElements hands out a COPY (as a JVM may) and Release copies back unless the
// mode is JNI_ABORT, so the bridge's release modes are exercised for real. Direct buffers wrap caller memory (numpy arrays).
#include <jni.h>

#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {
struct MockArray : _jarray { int kind; std::vector<unsigned char> data; int len; };      // kind 0 = int, 1 = float
struct MockBuffer : _jobject { void* addr; jlong cap; };
struct MockClass : _jclass { std::string name; };
thread_local std::string g_exc_class, g_exc_msg;
thread_local bool g_exc = false;
int g_outstanding = 0;            // Get...Elements without a matching Release

jclass FindClass(JNIEnv*, const char* n) { auto* c = new MockClass(); c->name = n; return c; }      // leaked: test process only
jint ThrowNew(JNIEnv*, jclass c, const char* m) {
    g_exc = true; g_exc_class = static_cast<MockClass*>(c)->name; g_exc_msg = m ? m : "";
    return 0;
}
jsize GetArrayLength(JNIEnv*, jarray a) { return static_cast<MockArray*>(a)->len; }
template <typename T, typename A>
T* get_elems(A a, jboolean* isCopy) {
    MockArray* m = static_cast<MockArray*>(static_cast<jarray>(a));
    T* copy = static_cast<T*>(malloc(m->data.size() ? m->data.size() : 1));
    memcpy(copy, m->data.data(), m->data.size());
    if (isCopy) *isCopy = 1;
    g_outstanding++;
    return copy;
}
template <typename T, typename A>
void release_elems(A a, T* e, jint mode) {
    MockArray* m = static_cast<MockArray*>(static_cast<jarray>(a));
    if (mode != JNI_ABORT) memcpy(m->data.data(), e, m->data.size());
    if (mode != JNI_COMMIT) { free(e); g_outstanding--; }
}
jint* GetIntArrayElements(JNIEnv*, jintArray a, jboolean* c) { return get_elems<jint>(a, c); }
jfloat* GetFloatArrayElements(JNIEnv*, jfloatArray a, jboolean* c) { return get_elems<jfloat>(a, c); }
void ReleaseIntArrayElements(JNIEnv*, jintArray a, jint* e, jint mode) { release_elems(a, e, mode); }
void ReleaseFloatArrayElements(JNIEnv*, jfloatArray a, jfloat* e, jint mode) { release_elems(a, e, mode); }
jboolean ExceptionCheck(JNIEnv*) { return g_exc ? 1 : 0; }
void* GetDirectBufferAddress(JNIEnv*, jobject b) { return static_cast<MockBuffer*>(b)->addr; }
jlong GetDirectBufferCapacity(JNIEnv*, jobject b) { return static_cast<MockBuffer*>(b)->cap; }

JNINativeInterface_ make_table() {
    JNINativeInterface_ t;
    memset(&t, 0, sizeof t);
    t.FindClass = FindClass; t.ThrowNew = ThrowNew; t.GetArrayLength = GetArrayLength;
    t.GetIntArrayElements = GetIntArrayElements; t.GetFloatArrayElements = GetFloatArrayElements;
    t.ReleaseIntArrayElements = ReleaseIntArrayElements; t.ReleaseFloatArrayElements = ReleaseFloatArrayElements;
    t.ExceptionCheck = ExceptionCheck; t.GetDirectBufferAddress = GetDirectBufferAddress; t.GetDirectBufferCapacity = GetDirectBufferCapacity;
    return t;
}
const JNINativeInterface_ g_table = make_table();
JNIEnv g_env = {&g_table};
}  // namespace

#define MOCK_API extern "C" __attribute__((visibility("default")))
MOCK_API void* mock_env() { return &g_env; }
MOCK_API void* mock_new_float_array(const float* src, int n) {
    auto* a = new MockArray(); a->kind = 1; a->len = n; a->data.resize((size_t)n * 4);
    if (src) memcpy(a->data.data(), src, (size_t)n * 4);
    return static_cast<jarray>(a);
}
MOCK_API void* mock_new_int_array(int n) { auto* a = new MockArray(); a->kind = 0; a->len = n; a->data.assign((size_t)n * 4, 0); return static_cast<jarray>(a); }
MOCK_API void mock_array_read(void* arr, void* out) { auto* a = static_cast<MockArray*>(static_cast<jarray>(arr)); memcpy(out, a->data.data(), a->data.size()); }
MOCK_API void mock_free_array(void* arr) { delete static_cast<MockArray*>(static_cast<jarray>(arr)); }
MOCK_API void* mock_new_direct_buffer(void* addr, int64_t cap) { auto* b = new MockBuffer(); b->addr = addr; b->cap = cap; return static_cast<jobject>(b); }
MOCK_API void mock_free_buffer(void* b) { delete static_cast<MockBuffer*>(static_cast<jobject>(b)); }
MOCK_API int mock_exception_pending() { return g_exc ? 1 : 0; }
MOCK_API const char* mock_exception_class() { return g_exc_class.c_str(); }
MOCK_API const char* mock_exception_message() { return g_exc_msg.c_str(); }
MOCK_API void mock_exception_clear() { g_exc = false; g_exc_class.clear(); g_exc_msg.clear(); }
MOCK_API int mock_outstanding_elements() { return g_outstanding; }
/* layout check for the stand-in header: slot indices of the JNI function table (JNI specification numbering) */
MOCK_API int mock_slot_of(const char* name) {
    const size_t P = sizeof(void*);
    std::string n(name);
    if (n == "FindClass") return (int)(offsetof(JNINativeInterface_, FindClass) / P);
    if (n == "ThrowNew") return (int)(offsetof(JNINativeInterface_, ThrowNew) / P);
    if (n == "GetArrayLength") return (int)(offsetof(JNINativeInterface_, GetArrayLength) / P);
    if (n == "GetIntArrayElements") return (int)(offsetof(JNINativeInterface_, GetIntArrayElements) / P);
    if (n == "GetFloatArrayElements") return (int)(offsetof(JNINativeInterface_, GetFloatArrayElements) / P);
    if (n == "ReleaseIntArrayElements") return (int)(offsetof(JNINativeInterface_, ReleaseIntArrayElements) / P);
    if (n == "ReleaseFloatArrayElements") return (int)(offsetof(JNINativeInterface_, ReleaseFloatArrayElements) / P);
    if (n == "ExceptionCheck") return (int)(offsetof(JNINativeInterface_, ExceptionCheck) / P);
    if (n == "GetDirectBufferAddress") return (int)(offsetof(JNINativeInterface_, GetDirectBufferAddress) / P);
    if (n == "GetDirectBufferCapacity") return (int)(offsetof(JNINativeInterface_, GetDirectBufferCapacity) / P);
    return -1;
}
