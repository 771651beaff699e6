/* JNI shim between PgenNative and libpgen_cuda.so: what a maintainer of the reference adds
 * next to TFModel.java so that the flatMap operator calls the C ABI of include/pgen.h instead of handing rows to the
 * Python / TensorFlow worker (src/main/java/org/apache/flink/table/ml/lib/tensorflow/TFModel.java:76).
 * Build: gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude pgen_jni.c -lpgen_cuda
 * NOT compiled against a JDK here (the image has none).  tests/test_jni_shim.py compiles it against a stand-in
 * jni.h (tests/jni_stub/, declarations of the members used below only) and drives it with a mock JNIEnv. */
#include <jni.h>
#include "pgen.h"

static void throw_rt(JNIEnv* e, const char* msg) {
  (*e)->ThrowNew(e, (*e)->FindClass(e, "java/lang/RuntimeException"), msg);
}
JNIEXPORT jlong JNICALL Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_create
    (JNIEnv* e, jclass c, jstring hyper, jint device) {
  const char* s = (*e)->GetStringUTFChars(e, hyper, 0);
  pgen_t* h = 0;
  int rc = pgen_create(s, device, &h);
  (*e)->ReleaseStringUTFChars(e, hyper, s);
  if (rc != PGEN_OK) { throw_rt(e, pgen_last_error(0)); return 0; }
  return (jlong)(intptr_t)h;
}
JNIEXPORT void JNICALL Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_submit
    (JNIEnv* e, jclass c, jlong hh, jbyteArray uuid, jbyteArray art, jbyteArray ref) {
  pgen_t* h = (pgen_t*)(intptr_t)hh;
  jsize nu = (*e)->GetArrayLength(e, uuid), na = (*e)->GetArrayLength(e, art), nr = (*e)->GetArrayLength(e, ref);
  jbyte *pu = (*e)->GetByteArrayElements(e, uuid, 0), *pa = (*e)->GetByteArrayElements(e, art, 0),
        *pr = (*e)->GetByteArrayElements(e, ref, 0);
  int rc = pgen_submit(h, (char*)pu, nu, (char*)pa, na, (char*)pr, nr);   /* inputs borrowed for the call */
  (*e)->ReleaseByteArrayElements(e, uuid, pu, JNI_ABORT);
  (*e)->ReleaseByteArrayElements(e, art, pa, JNI_ABORT);
  (*e)->ReleaseByteArrayElements(e, ref, pr, JNI_ABORT);
  if (rc != PGEN_OK) throw_rt(e, pgen_last_error(h));
}
static jbyteArray bytes_of(JNIEnv* e, const char* p, size_t n) {
  jbyteArray a = (*e)->NewByteArray(e, (jsize)n);
  if (a && n) (*e)->SetByteArrayRegion(e, a, 0, (jsize)n, (const jbyte*)p);
  return a;
}
/* rows x {uuid, article, summary, reference}; an empty array when nothing has finished yet */
JNIEXPORT jobjectArray JNICALL Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_poll
    (JNIEnv* e, jclass c, jlong hh, jboolean flush) {
  pgen_t* h = (pgen_t*)(intptr_t)hh;
  pgen_rows_t rows;
  int rc = pgen_poll(h, flush ? 1 : 0, &rows);
  /* on an error `rows` is empty; rows that had finished stay queued inside the handle and come with the next poll
   * (a row that could not be mapped back to words among them, with an empty summary) */
  jclass row_cls = (*e)->FindClass(e, "[[B"), col_cls = (*e)->FindClass(e, "[B");
  jobjectArray out = (*e)->NewObjectArray(e, (jsize)rows.n_rows, row_cls, 0);
  for (size_t i = 0; out && i < rows.n_rows; ++i) {
    const pgen_row_t* r = &rows.rows[i];
    jobjectArray cols = (*e)->NewObjectArray(e, 4, col_cls, 0);
    (*e)->SetObjectArrayElement(e, cols, 0, bytes_of(e, r->uuid, r->uuid_len));
    (*e)->SetObjectArrayElement(e, cols, 1, bytes_of(e, r->article, r->article_len));
    (*e)->SetObjectArrayElement(e, cols, 2, bytes_of(e, r->summary, r->summary_len));
    (*e)->SetObjectArrayElement(e, cols, 3, bytes_of(e, r->reference, r->reference_len));
    (*e)->SetObjectArrayElement(e, out, (jsize)i, cols);
    (*e)->DeleteLocalRef(e, cols);                     /* long polls: do not pile up local references */
  }
  pgen_rows_free(&rows);                               /* the rows were copied into Java arrays above */
  if (rc != PGEN_OK) throw_rt(e, pgen_last_error(h));
  return out;
}
JNIEXPORT void JNICALL Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_destroy
    (JNIEnv* e, jclass c, jlong hh) {
  pgen_destroy((pgen_t*)(intptr_t)hh);                 /* frees device memory, streams, graphs; NULL is fine */
}
JNIEXPORT jint JNICALL Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_deviceCount
    (JNIEnv* e, jclass c) {
  return (jint)pgen_device_count();                    /* 0: no CUDA device, create() would fail */
}
