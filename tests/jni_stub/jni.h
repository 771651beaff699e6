/* STAND-IN for the JDK's jni.h, for compile and logic checks only (the image has no JDK).  It declares the types and
 * the JNIEnv members that textsummarization-on-flink_b200/jni/pgen_jni.c uses, with the JNI specification's signatures;
 * the function table has ONLY those members, so its layout is NOT the JVM's: objects built with this header work with
 * tests/jni_stub/mock_jvm.c and with nothing else. */
#ifndef PGEN_TEST_JNI_STUB_H_
#define PGEN_TEST_JNI_STUB_H_
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef jint jsize;
struct _jobject;
typedef struct _jobject* jobject;
typedef jobject jclass;
typedef jobject jstring;
typedef jobject jarray;
typedef jarray jbyteArray;
typedef jarray jobjectArray;
typedef jobject jthrowable;

struct JNINativeInterface_;
typedef const struct JNINativeInterface_* JNIEnv;

struct JNINativeInterface_ {
  jclass (*FindClass)(JNIEnv*, const char*);
  jint (*ThrowNew)(JNIEnv*, jclass, const char*);
  const char* (*GetStringUTFChars)(JNIEnv*, jstring, jboolean*);
  void (*ReleaseStringUTFChars)(JNIEnv*, jstring, const char*);
  jsize (*GetArrayLength)(JNIEnv*, jarray);
  jbyte* (*GetByteArrayElements)(JNIEnv*, jbyteArray, jboolean*);
  void (*ReleaseByteArrayElements)(JNIEnv*, jbyteArray, jbyte*, jint);
  jbyteArray (*NewByteArray)(JNIEnv*, jsize);
  void (*SetByteArrayRegion)(JNIEnv*, jbyteArray, jsize, jsize, const jbyte*);
  jobjectArray (*NewObjectArray)(JNIEnv*, jsize, jclass, jobject);
  void (*SetObjectArrayElement)(JNIEnv*, jobjectArray, jsize, jobject);
  void (*DeleteLocalRef)(JNIEnv*, jobject);
};
#endif
