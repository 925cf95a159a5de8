// emrf_jni.cpp — thin JNI shim: com.euromillioner.gpu.Native -> include/emrf.h, 1:1.
// SOURCE ONLY: there is no jni.h / JDK in the build image, so this file is not part of the build
// (compile with: g++ -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../../include
//  emrf_jni.cpp -L.. -lemrf -o libemrf_jni.so).
#include <jni.h>

#include <vector>

#include "emrf.h"

static void throw_arg(JNIEnv* env, const char* msg) { env->ThrowNew(env->FindClass("java/lang/IllegalArgumentException"), msg); }
static void throw_for(JNIEnv* env, int rc) {
  const char* cls = rc == EMRF_ERR_ARG ? "java/lang/IllegalArgumentException" : "java/lang/RuntimeException";
  env->ThrowNew(env->FindClass(cls), emrf_last_error());
}
#define CHECK(rc, ret) do { int _r = (rc); if (_r != EMRF_OK) { throw_for(env, _r); return ret; } } while (0)

extern "C" {
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_ctxCreate(JNIEnv* env, jclass, jint device) {
  emrf_ctx_t* c = nullptr;
  CHECK(emrf_ctx_create(device, &c), 0);
  return (jlong)c;
}
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_ctxDestroy(JNIEnv*, jclass, jlong ctx) { emrf_ctx_destroy((emrf_ctx_t*)ctx); }
JNIEXPORT jbyteArray JNICALL Java_com_euromillioner_gpu_Native_ncclUniqueId(JNIEnv* env, jclass) {
  uint8_t uid[128];
  CHECK(emrf_nccl_unique_id(uid), nullptr);
  jbyteArray out = env->NewByteArray(128);
  env->SetByteArrayRegion(out, 0, 128, (const jbyte*)uid);
  return out;
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_ctxCreateDist(JNIEnv* env, jclass, jint device, jint rank, jint world,
                                                                        jbyteArray juid) {
  if (!juid || env->GetArrayLength(juid) != 128) { throw_arg(env, "the NCCL unique id has 128 bytes"); return 0; }
  uint8_t uid[128];
  env->GetByteArrayRegion(juid, 0, 128, (jbyte*)uid);
  emrf_ctx_t* c = nullptr;
  CHECK(emrf_ctx_create_dist(device, rank, world, uid, &c), 0);
  return (jlong)c;
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_ctxReleaseBuffers(JNIEnv* env, jclass, jlong ctx) {
  int64_t freed = 0;
  CHECK(emrf_ctx_release_buffers((emrf_ctx_t*)ctx, &freed), 0);
  return (jlong)freed;
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_hostAllocPinned(JNIEnv* env, jclass, jlong bytes) {
  void* p = nullptr;
  CHECK(emrf_host_alloc_pinned(bytes, &p), 0);
  return (jlong)p;
}
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_hostFreePinned(JNIEnv*, jclass, jlong addr) { emrf_host_free_pinned((void*)addr); }
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_putDoubles(JNIEnv* env, jclass, jlong dst, jlong off, jdoubleArray src, jint len) {
  env->GetDoubleArrayRegion(src, 0, len, (jdouble*)dst + off);   // (throws ArrayIndexOutOfBoundsException itself when len is too large)
}

static emrf_params_t decode_params(JNIEnv* env, jintArray jp, jdoubleArray jd, jlong seed) {
  jint p[9]; jdouble d[3];
  env->GetIntArrayRegion(jp, 0, 9, p);
  env->GetDoubleArrayRegion(jd, 0, 3, d);
  emrf_params_t P = {};
  P.num_trees = p[0]; P.max_depth = p[1]; P.max_bins = p[2]; P.min_instances_per_node = p[3]; P.num_classes = p[4];
  P.impurity = p[5]; P.feature_subset_kind = p[6]; P.max_memory_mb = p[7]; P.find_splits_policy = p[8];
  P.feature_subset_value = d[0]; P.min_info_gain = d[1]; P.subsampling_rate = d[2]; P.seed = seed;
  return P;
}
static std::vector<jlong> long_array(JNIEnv* env, jlongArray a) {
  std::vector<jlong> v(env->GetArrayLength(a));
  env->GetLongArrayRegion(a, 0, (jsize)v.size(), v.data());
  return v;
}
static jlong fit_common(JNIEnv* env, jlong ctx, jintArray jp, jdoubleArray jd, jlong seed, const double* X, const double* y, jlong n,
                        jint F, jlongArray jpo) {
  const emrf_params_t P = decode_params(env, jp, jd, seed);
  const std::vector<jlong> po = long_array(env, jpo);
  emrf_forest_t* f = nullptr;
  CHECK(emrf_fit((emrf_ctx_t*)ctx, &P, X, y, n, F, (const int64_t*)po.data(), (int32_t)po.size() - 1, &f), 0);
  return (jlong)f;
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_fit(JNIEnv* env, jclass, jlong ctx, jintArray jp, jdoubleArray jd, jlong seed,
                                                              jobject features, jobject labels, jlong n, jint F, jlongArray jpo) {
  return fit_common(env, ctx, jp, jd, seed, (const double*)env->GetDirectBufferAddress(features),
                    (const double*)env->GetDirectBufferAddress(labels), n, F, jpo);
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_fitAt(JNIEnv* env, jclass, jlong ctx, jintArray jp, jdoubleArray jd, jlong seed,
                                                                jlong features, jlong labels, jlong n, jint F, jlongArray jpo) {
  return fit_common(env, ctx, jp, jd, seed, (const double*)features, (const double*)labels, n, F, jpo);
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_fitShardedAt(JNIEnv* env, jclass, jlong ctx, jintArray jp, jdoubleArray jd,
                                                                       jlong seed, jlong features, jlong labels, jlong n_local, jint F,
                                                                       jlongArray jgo, jint first_partition, jint n_partitions_local) {
  const emrf_params_t P = decode_params(env, jp, jd, seed);
  const std::vector<jlong> go = long_array(env, jgo);
  emrf_forest_t* f = nullptr;
  CHECK(emrf_fit_sharded((emrf_ctx_t*)ctx, &P, (const double*)features, (const double*)labels, n_local, F, (const int64_t*)go.data(),
                         (int32_t)go.size() - 1, first_partition, n_partitions_local, &f), 0);
  return (jlong)f;
}
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_forestFree(JNIEnv*, jclass, jlong f) { emrf_forest_free((emrf_forest_t*)f); }
JNIEXPORT jlongArray JNICALL Java_com_euromillioner_gpu_Native_forestInfo(JNIEnv* env, jclass, jlong f) {
  int32_t T, C, F; int64_t n;
  CHECK(emrf_forest_info((emrf_forest_t*)f, &T, &C, &F, &n), nullptr);
  jlong v[4] = {T, C, F, n};
  jlongArray out = env->NewLongArray(4);
  env->SetLongArrayRegion(out, 0, 4, v);
  return out;
}
JNIEXPORT jdoubleArray JNICALL Java_com_euromillioner_gpu_Native_featureImportances(JNIEnv* env, jclass, jlong f) {
  int32_t T, C, F; int64_t n;
  CHECK(emrf_forest_info((emrf_forest_t*)f, &T, &C, &F, &n), nullptr);
  std::vector<double> imp(F);
  CHECK(emrf_forest_feature_importances((emrf_forest_t*)f, imp.data()), nullptr);
  jdoubleArray out = env->NewDoubleArray(F);
  env->SetDoubleArrayRegion(out, 0, F, imp.data());
  return out;
}
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_predict(JNIEnv* env, jclass, jlong ctx, jlong f, jobject features, jlong n,
                                                                 jobject raw, jobject prob, jobject pred) {
  CHECK(emrf_predict((emrf_ctx_t*)ctx, (emrf_forest_t*)f, (const double*)env->GetDirectBufferAddress(features), n,
                     raw ? (double*)env->GetDirectBufferAddress(raw) : nullptr,
                     prob ? (double*)env->GetDirectBufferAddress(prob) : nullptr,
                     pred ? (double*)env->GetDirectBufferAddress(pred) : nullptr), );
}
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_predictThresholds(JNIEnv* env, jclass, jlong ctx, jlong f, jobject features,
                                                                           jlong n, jdoubleArray thresholds, jobject raw, jobject prob,
                                                                           jobject pred) {
  jdouble* t = env->GetDoubleArrayElements(thresholds, nullptr);
  const int rc = emrf_predict_thresholds((emrf_ctx_t*)ctx, (emrf_forest_t*)f, (const double*)env->GetDirectBufferAddress(features), n, t,
                                         raw ? (double*)env->GetDirectBufferAddress(raw) : nullptr,
                                         prob ? (double*)env->GetDirectBufferAddress(prob) : nullptr,
                                         pred ? (double*)env->GetDirectBufferAddress(pred) : nullptr);
  env->ReleaseDoubleArrayElements(thresholds, t, JNI_ABORT);
  CHECK(rc, );
}
// forestExport: ten GetPrimitiveArrayCritical / emrf_forest_export / Release pairs; omitted arrays are an error.
JNIEXPORT void JNICALL Java_com_euromillioner_gpu_Native_forestExport(JNIEnv* env, jclass, jlong f, jlongArray treeOff, jintArray nodeId,
                                                                      jintArray left, jintArray right, jintArray feature, jdoubleArray thr,
                                                                      jdoubleArray gain, jdoubleArray imp, jdoubleArray pred, jdoubleArray stats) {
  jlong* a0 = env->GetLongArrayElements(treeOff, nullptr);
  jint *a1 = env->GetIntArrayElements(nodeId, nullptr), *a2 = env->GetIntArrayElements(left, nullptr),
       *a3 = env->GetIntArrayElements(right, nullptr), *a4 = env->GetIntArrayElements(feature, nullptr);
  jdouble *a5 = env->GetDoubleArrayElements(thr, nullptr), *a6 = env->GetDoubleArrayElements(gain, nullptr),
          *a7 = env->GetDoubleArrayElements(imp, nullptr), *a8 = env->GetDoubleArrayElements(pred, nullptr),
          *a9 = env->GetDoubleArrayElements(stats, nullptr);
  int rc = emrf_forest_export((emrf_forest_t*)f, (int64_t*)a0, a1, a2, a3, a4, a5, a6, a7, a8, a9);
  env->ReleaseLongArrayElements(treeOff, a0, 0);
  env->ReleaseIntArrayElements(nodeId, a1, 0); env->ReleaseIntArrayElements(left, a2, 0);
  env->ReleaseIntArrayElements(right, a3, 0); env->ReleaseIntArrayElements(feature, a4, 0);
  env->ReleaseDoubleArrayElements(thr, a5, 0); env->ReleaseDoubleArrayElements(gain, a6, 0);
  env->ReleaseDoubleArrayElements(imp, a7, 0); env->ReleaseDoubleArrayElements(pred, a8, 0);
  env->ReleaseDoubleArrayElements(stats, a9, 0);
  if (rc != EMRF_OK) throw_for(env, rc);
}
JNIEXPORT jlong JNICALL Java_com_euromillioner_gpu_Native_forestImport(JNIEnv* env, jclass, jlong ctx, jint T, jint C, jint F,
                                                                       jlongArray treeOff, jintArray nodeId, jintArray left, jintArray right,
                                                                       jintArray feature, jdoubleArray thr, jdoubleArray gain,
                                                                       jdoubleArray imp, jdoubleArray pred, jdoubleArray stats) {
  if (env->GetArrayLength(treeOff) != T + 1) { throw_arg(env, "treeOff must hold numTrees + 1 offsets"); return 0; }
  jlong* a0 = env->GetLongArrayElements(treeOff, nullptr);
  const jsize n = (jsize)a0[T];
  const bool sized = n >= 0 && env->GetArrayLength(nodeId) == n && env->GetArrayLength(left) == n && env->GetArrayLength(right) == n &&
                     env->GetArrayLength(feature) == n && env->GetArrayLength(thr) == n && env->GetArrayLength(gain) == n &&
                     env->GetArrayLength(imp) == n && env->GetArrayLength(pred) == n && (jlong)env->GetArrayLength(stats) == (jlong)n * C;
  if (!sized) {
    env->ReleaseLongArrayElements(treeOff, a0, JNI_ABORT);
    throw_arg(env, "node arrays must hold treeOff[numTrees] entries (stats: times numClasses)");
    return 0;
  }
  jint *a1 = env->GetIntArrayElements(nodeId, nullptr), *a2 = env->GetIntArrayElements(left, nullptr),
       *a3 = env->GetIntArrayElements(right, nullptr), *a4 = env->GetIntArrayElements(feature, nullptr);
  jdouble *a5 = env->GetDoubleArrayElements(thr, nullptr), *a6 = env->GetDoubleArrayElements(gain, nullptr),
          *a7 = env->GetDoubleArrayElements(imp, nullptr), *a8 = env->GetDoubleArrayElements(pred, nullptr),
          *a9 = env->GetDoubleArrayElements(stats, nullptr);
  emrf_forest_t* f = nullptr;
  const int rc = emrf_forest_import((emrf_ctx_t*)ctx, T, C, F, (const int64_t*)a0, a1, a2, a3, a4, a5, a6, a7, a8, a9, &f);
  env->ReleaseLongArrayElements(treeOff, a0, JNI_ABORT);
  env->ReleaseIntArrayElements(nodeId, a1, JNI_ABORT); env->ReleaseIntArrayElements(left, a2, JNI_ABORT);
  env->ReleaseIntArrayElements(right, a3, JNI_ABORT); env->ReleaseIntArrayElements(feature, a4, JNI_ABORT);
  env->ReleaseDoubleArrayElements(thr, a5, JNI_ABORT); env->ReleaseDoubleArrayElements(gain, a6, JNI_ABORT);
  env->ReleaseDoubleArrayElements(imp, a7, JNI_ABORT); env->ReleaseDoubleArrayElements(pred, a8, JNI_ABORT);
  env->ReleaseDoubleArrayElements(stats, a9, JNI_ABORT);
  CHECK(rc, 0);
  return (jlong)f;
}
}
