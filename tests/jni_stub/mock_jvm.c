/* A mock JNIEnv for tests/test_jni_shim.py: just enough of a JVM's object model (strings, byte arrays, object arrays,
 * one pending exception) to drive textsummarization-on-flink_b200/jni/pgen_jni.c from Python through ctypes.  Goes with
 * the stand-in jni.h in this directory; nothing here is a JVM, and nothing here ships. */
#include <stdlib.h>
#include <string.h>

#include "jni.h"

enum { K_CLASS = 1, K_STRING, K_BYTES, K_OBJECTS };
struct _jobject {
  int kind;
  jsize len;
  char* data;             /* K_STRING: NUL-terminated, K_BYTES: len bytes, K_CLASS: the class name */
  struct _jobject** elems; /* K_OBJECTS */
  int borrowed;           /* GetByteArrayElements / GetStringUTFChars not yet released */
};

static char* g_exc_class = 0;
static char* g_exc_msg = 0;
static int g_local_refs_deleted = 0;

static jobject new_obj(int kind) {
  jobject o = (jobject)calloc(1, sizeof(struct _jobject));
  o->kind = kind;
  return o;
}
static jclass m_FindClass(JNIEnv* e, const char* name) {
  jobject o = new_obj(K_CLASS);
  o->data = strdup(name);
  return o;
}
static jint m_ThrowNew(JNIEnv* e, jclass c, const char* msg) {
  free(g_exc_class); free(g_exc_msg);
  g_exc_class = strdup(c && c->data ? c->data : "?");
  g_exc_msg = strdup(msg ? msg : "");
  return 0;
}
static const char* m_GetStringUTFChars(JNIEnv* e, jstring s, jboolean* copy) { s->borrowed++; return s->data; }
static void m_ReleaseStringUTFChars(JNIEnv* e, jstring s, const char* p) { s->borrowed--; }
static jsize m_GetArrayLength(JNIEnv* e, jarray a) { return a->len; }
static jbyte* m_GetByteArrayElements(JNIEnv* e, jbyteArray a, jboolean* copy) { a->borrowed++; return (jbyte*)a->data; }
static void m_ReleaseByteArrayElements(JNIEnv* e, jbyteArray a, jbyte* p, jint mode) { a->borrowed--; }
static jbyteArray m_NewByteArray(JNIEnv* e, jsize n) {
  jobject o = new_obj(K_BYTES);
  o->len = n;
  o->data = (char*)calloc((size_t)n + 1, 1);
  return o;
}
static void m_SetByteArrayRegion(JNIEnv* e, jbyteArray a, jsize start, jsize n, const jbyte* src) {
  if (start >= 0 && n >= 0 && start + n <= a->len) memcpy(a->data + start, src, (size_t)n);
}
static jobjectArray m_NewObjectArray(JNIEnv* e, jsize n, jclass c, jobject init) {
  jobject o = new_obj(K_OBJECTS);
  o->len = n;
  o->elems = (struct _jobject**)calloc((size_t)n + 1, sizeof(struct _jobject*));
  return o;
}
static void m_SetObjectArrayElement(JNIEnv* e, jobjectArray a, jsize i, jobject v) {
  if (i >= 0 && i < a->len) a->elems[i] = v;
}
static void m_DeleteLocalRef(JNIEnv* e, jobject o) { g_local_refs_deleted++; }   /* objects stay alive for the test to read */

static const struct JNINativeInterface_ g_table = {
    m_FindClass, m_ThrowNew, m_GetStringUTFChars, m_ReleaseStringUTFChars, m_GetArrayLength, m_GetByteArrayElements,
    m_ReleaseByteArrayElements, m_NewByteArray, m_SetByteArrayRegion, m_NewObjectArray, m_SetObjectArrayElement, m_DeleteLocalRef};
static JNIEnv g_env = &g_table;

/* ---- what the Python test calls ---- */
JNIEXPORT JNIEnv* mock_env(void) { return &g_env; }
JNIEXPORT jobject mock_new_string(const char* s) {
  jobject o = new_obj(K_STRING);
  o->data = strdup(s);
  o->len = (jsize)strlen(s);
  return o;
}
JNIEXPORT jobject mock_new_bytes(const char* p, int n) {
  jobject o = m_NewByteArray(&g_env, n);
  if (n) memcpy(o->data, p, (size_t)n);
  return o;
}
JNIEXPORT const char* mock_exception_class(void) { return g_exc_class; }
JNIEXPORT const char* mock_exception_message(void) { return g_exc_msg; }
JNIEXPORT void mock_clear_exception(void) { free(g_exc_class); free(g_exc_msg); g_exc_class = g_exc_msg = 0; }
JNIEXPORT int mock_length(jobject o) { return o ? o->len : -1; }
JNIEXPORT jobject mock_element(jobject o, int i) { return (o && o->kind == K_OBJECTS && i >= 0 && i < o->len) ? o->elems[i] : 0; }
JNIEXPORT const char* mock_data(jobject o) { return o ? o->data : 0; }
JNIEXPORT int mock_borrowed(jobject o) { return o ? o->borrowed : -1; }
JNIEXPORT int mock_local_refs_deleted(void) { return g_local_refs_deleted; }
