// minimal stand-in for <jni.h>: just enough declarations to type-check emrf_jni.cpp without a JDK (tests/test_tools.py)
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef double jdouble; typedef int8_t jbyte; typedef jint jsize;
struct _jobject {}; typedef _jobject* jobject; typedef jobject jclass; typedef jobject jarray;
typedef jobject jintArray; typedef jobject jlongArray; typedef jobject jdoubleArray; typedef jobject jbyteArray;
#define JNIEXPORT
#define JNICALL
#define JNI_ABORT 2
struct JNIEnv {
  jclass FindClass(const char*); jint ThrowNew(jclass, const char*);
  jsize GetArrayLength(jarray);
  void GetIntArrayRegion(jintArray, jsize, jsize, jint*); void GetLongArrayRegion(jlongArray, jsize, jsize, jlong*);
  void GetDoubleArrayRegion(jdoubleArray, jsize, jsize, jdouble*); void GetByteArrayRegion(jbyteArray, jsize, jsize, jbyte*);
  void SetLongArrayRegion(jlongArray, jsize, jsize, const jlong*); void SetDoubleArrayRegion(jdoubleArray, jsize, jsize, const jdouble*);
  void SetByteArrayRegion(jbyteArray, jsize, jsize, const jbyte*);
  jlongArray NewLongArray(jsize); jdoubleArray NewDoubleArray(jsize); jbyteArray NewByteArray(jsize);
  void* GetDirectBufferAddress(jobject);
  jint* GetIntArrayElements(jintArray, bool*); jlong* GetLongArrayElements(jlongArray, bool*); jdouble* GetDoubleArrayElements(jdoubleArray, bool*);
  void ReleaseIntArrayElements(jintArray, jint*, jint); void ReleaseLongArrayElements(jlongArray, jlong*, jint);
  void ReleaseDoubleArrayElements(jdoubleArray, jdouble*, jint);
};
