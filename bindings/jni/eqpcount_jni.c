/* eqpcount_jni.c — JNI glue between EqpCountNative.java and libeqpcount.so (include/eqp_count.h).
 *
 * The reference's host language for the compute-counts path is Java (exon-pipeline-scripts/java/source/bed-source/ComputeCounts.java
 * and the other mains compute-counts.sh:733-1037 spawns); this file lets the Java side keep I/O and the CLI and call the CUDA path
 * through the C ABI.  No JDK exists in the build image, so the translation unit is guarded: without -DEQP_HAVE_JNI it compiles to
 * nothing (tests/test_bindings.py checks that it at least parses as C with a stub jni.h when the guard is on).
 *
 *   gcc -shared -fPIC -DEQP_HAVE_JNI -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude bindings/jni/eqpcount_jni.c \
 *       -Leqp-cluster_b200/lib -leqpcount -o libeqpcount_jni.so
 */
#ifdef EQP_HAVE_JNI
#include <jni.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "eqp_count.h"

/* ComputeCounts.java:996-999: any failure becomes a message containing ERROR; the Java main prints it and exits 1 */
static void throw_io(JNIEnv* e, eqp_ctx* c) {
  char msg[700]; snprintf(msg, sizeof msg, "ERROR: %s", eqp_last_error(c));
  (*e)->ThrowNew(e, (*e)->FindClass(e, "java/io/IOException"), msg);
}
static void params_from(JNIEnv* e, jintArray ip, jdoubleArray dp, eqp_params* p) {
  jint iv[10]; jdouble dv[2];
  eqp_default_params(p);
  (*e)->GetIntArrayRegion(e, ip, 0, 10, iv); (*e)->GetDoubleArrayRegion(e, dp, 0, 2, dv);
  p->paired_end = iv[0]; p->stranded = iv[1]; p->gene_overlap = iv[2]; p->exon_overlap = iv[3]; p->junction_overlap = iv[4];
  p->unweighted = iv[5]; p->unambiguous = iv[6]; p->containment = iv[7]; p->nonsplice_conforming = iv[8]; p->modes = iv[9];
  p->read_weight_threshold = dv[0]; p->distance_slack = dv[1];
}
static void results_to(JNIEnv* e, const eqp_counts* c, jdoubleArray gene, jdoubleArray exon, jdoubleArray junc, jbyteArray act, jlongArray sc, jdoubleArray sums) {
  jlong s[12];
  (*e)->SetDoubleArrayRegion(e, gene, 0, (jsize)c->n_gene, c->gene);
  (*e)->SetDoubleArrayRegion(e, exon, 0, (jsize)c->n_exon, c->exon);
  (*e)->SetDoubleArrayRegion(e, junc, 0, (jsize)c->n_junction, c->junction);
  (*e)->SetByteArrayRegion(e, act, 0, (jsize)c->n_junction_act, (const jbyte*)c->junction_active);
  s[0] = c->number_mapped_reads; s[1] = c->number_expressed_reads;
  for (int m = 0; m < 3; m++) { s[2 + m] = c->num_weight_objects[m]; s[5 + m] = c->num_weight_objects_included[m]; s[8 + m] = c->num_weight_objects_excluded[m]; }
  s[11] = c->kernel_launches;
  (*e)->SetLongArrayRegion(e, sc, 0, 12, s);
  (*e)->SetDoubleArrayRegion(e, sums, 0, 3, c->sum_weights);
}

JNIEXPORT jlong JNICALL Java_EqpCountNative_create(JNIEnv* e, jclass k, jint device, jobjectArray files) {
  eqp_ctx* ctx = NULL; const char* p[9]; jstring js[9]; eqp_annotation_files af; int rc;
  (void)k;
  if (eqp_create(device, &ctx)) { throw_io(e, NULL); return 0; }
  for (int i = 0; i < 9; i++) { js[i] = (jstring)(*e)->GetObjectArrayElement(e, files, i); p[i] = (*e)->GetStringUTFChars(e, js[i], NULL); }
  af.combined_exons_bed = p[0]; af.transcript_exons_bed = p[1]; af.exon_gene_map = p[2]; af.exon_exon_map = p[3]; af.exon_junction_map = p[4];
  af.exon_junction_transcript_map = p[5]; af.junction_junction_coordinate_map = p[6]; af.equivalent_junctions_map = p[7]; af.genome_fai = p[8];
  rc = eqp_load_annotation(ctx, &af);
  for (int i = 0; i < 9; i++) (*e)->ReleaseStringUTFChars(e, js[i], p[i]);
  if (rc) { throw_io(e, ctx); eqp_destroy(ctx); return 0; }
  return (jlong)(intptr_t)ctx;
}

JNIEXPORT jlongArray JNICALL Java_EqpCountNative_sizes(JNIEnv* e, jclass k, jlong h) {     /* n_gene, n_exon, n_junction, n_junction_act (array sizes) */
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; jlong v[4]; jlongArray out = (*e)->NewLongArray(e, 4); int64_t acts = 0, n;
  (void)k;
  for (int m = 0; m < 3; m++) v[m] = eqp_feature_count(ctx, m);
  n = eqp_output_row_count(ctx, 2);
  for (int64_t i = 0; i < n; i++) { eqp_output_row r; if (!eqp_get_output_row(ctx, 2, i, &r) && r.act + 1 > acts) acts = r.act + 1; }
  v[3] = acts;
  (*e)->SetLongArrayRegion(e, out, 0, 4, v);
  return out;
}

/* direct ByteBuffers hold the inflated SAM text: what `zcat` feeds the mains today (compute-intersection-bed-file.sh:644-662) */
JNIEXPORT void JNICALL Java_EqpCountNative_countChunk(JNIEnv* e, jclass k, jlong h, jintArray ip, jdoubleArray dp,
    jobject t, jlong tl, jobject j, jlong jl, jobject g, jlong gl, jdoubleArray gene, jdoubleArray exon, jdoubleArray junc, jbyteArray act, jlongArray sc, jdoubleArray sums) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; eqp_params p; eqp_counts c;
  (void)k;
  params_from(e, ip, dp, &p);
  if (eqp_count_chunk(ctx, &p, (const uint8_t*)(*e)->GetDirectBufferAddress(e, t), (size_t)tl, (const uint8_t*)(*e)->GetDirectBufferAddress(e, j), (size_t)jl,
                      (const uint8_t*)(*e)->GetDirectBufferAddress(e, g), (size_t)gl, /*on_device=*/0, &c)) { throw_io(e, ctx); return; }
  results_to(e, &c, gene, exon, junc, act, sc, sums);
}
JNIEXPORT void JNICALL Java_EqpCountNative_prefetchChunk(JNIEnv* e, jclass k, jlong h, jobject t, jlong tl, jobject j, jlong jl, jobject g, jlong gl) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h;
  (void)k;
  if (eqp_prefetch_chunk(ctx, (const uint8_t*)(*e)->GetDirectBufferAddress(e, t), (size_t)tl, (const uint8_t*)(*e)->GetDirectBufferAddress(e, j), (size_t)jl,
                         (const uint8_t*)(*e)->GetDirectBufferAddress(e, g), (size_t)gl)) throw_io(e, ctx);
}

/* streaming ingest: GZIPInputStream.read(byte[]) -> chunkFeed, one inflate thread per input, calls serialised by the Java side */
JNIEXPORT void JNICALL Java_EqpCountNative_chunkBegin(JNIEnv* e, jclass k, jlong h, jlong th, jlong jh, jlong gh) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; (void)k;
  if (eqp_chunk_begin(ctx, (size_t)th, (size_t)jh, (size_t)gh)) throw_io(e, ctx);
}
JNIEXPORT void JNICALL Java_EqpCountNative_chunkFeed(JNIEnv* e, jclass k, jlong h, jint input, jbyteArray buf, jint len) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; jbyte* p = (*e)->GetPrimitiveArrayCritical(e, buf, NULL); int rc;
  (void)k;
  rc = eqp_chunk_feed(ctx, input, (const uint8_t*)p, (size_t)len);           /* copies into the page-locked staging ring and returns */
  (*e)->ReleasePrimitiveArrayCritical(e, buf, p, JNI_ABORT);
  if (rc) throw_io(e, ctx);
}
/* one complete .sam.gz member (a direct ByteBuffer holding the file's bytes): compressed over PCIe, inflated on the device.
 * JNI_FALSE: a valid file the device decoder does not take (several members / BGZF): inflate it with GZIPInputStream and chunkFeed. */
JNIEXPORT jboolean JNICALL Java_EqpCountNative_chunkFeedGz(JNIEnv* e, jclass k, jlong h, jint input, jobject gz, jlong len) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; const uint8_t* p = (const uint8_t*)(*e)->GetDirectBufferAddress(e, gz); int rc;
  (void)k;
  if (!p) { jclass io = (*e)->FindClass(e, "java/io/IOException"); if (io) (*e)->ThrowNew(e, io, "chunkFeedGz: not a direct ByteBuffer"); return JNI_FALSE; }
  rc = eqp_chunk_feed_gz(ctx, input, p, (size_t)len);
  if (rc == EQP_ERR_UNSUPPORTED) return JNI_FALSE;
  if (rc) throw_io(e, ctx);
  return JNI_TRUE;
}
JNIEXPORT void JNICALL Java_EqpCountNative_chunkEnd(JNIEnv* e, jclass k, jlong h, jintArray ip, jdoubleArray dp,
    jdoubleArray gene, jdoubleArray exon, jdoubleArray junc, jbyteArray act, jlongArray sc, jdoubleArray sums) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; eqp_params p; eqp_counts c;
  (void)k;
  params_from(e, ip, dp, &p);
  if (eqp_chunk_end(ctx, &p, &c)) { throw_io(e, ctx); return; }
  results_to(e, &c, gene, exon, junc, act, sc, sums);
}

/* count-file rows in print order (ComputeCounts.java:875-950): ids / feature index / activation index / always flag */
JNIEXPORT jobjectArray JNICALL Java_EqpCountNative_rowIds(JNIEnv* e, jclass k, jlong h, jint mode) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; int64_t n = eqp_output_row_count(ctx, mode);
  jobjectArray out = (*e)->NewObjectArray(e, (jsize)n, (*e)->FindClass(e, "java/lang/String"), NULL);
  (void)k;
  for (int64_t i = 0; i < n; i++) { eqp_output_row r; eqp_get_output_row(ctx, mode, i, &r); jstring s = (*e)->NewStringUTF(e, r.id); (*e)->SetObjectArrayElement(e, out, (jsize)i, s); (*e)->DeleteLocalRef(e, s); }
  return out;
}
JNIEXPORT jintArray JNICALL Java_EqpCountNative_rowTable(JNIEnv* e, jclass k, jlong h, jint mode) {   /* 3 ints per row: feature, act, always; [3n] = map columns */
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; int64_t n = eqp_output_row_count(ctx, mode); jintArray out = (*e)->NewIntArray(e, (jsize)(3 * n + 1)); jint cols = eqp_map_columns(ctx, mode);
  (void)k;
  for (int64_t i = 0; i < n; i++) { eqp_output_row r; jint v[3]; eqp_get_output_row(ctx, mode, i, &r); v[0] = (jint)r.feature; v[1] = (jint)r.act; v[2] = r.always; (*e)->SetIntArrayRegion(e, out, (jsize)(3 * i), 3, v); }
  (*e)->SetIntArrayRegion(e, out, (jsize)(3 * n), 1, &cols);
  return out;
}

/* sample level: wide accumulator on the device (handle = device pointer), NCCL inside the library */
JNIEXPORT jbyteArray JNICALL Java_EqpCountNative_commId(JNIEnv* e, jclass k) {
  char id[EQP_COMM_ID_BYTES]; jbyteArray out; (void)k;
  if (eqp_comm_id(id)) { throw_io(e, NULL); return NULL; }
  out = (*e)->NewByteArray(e, EQP_COMM_ID_BYTES); (*e)->SetByteArrayRegion(e, out, 0, EQP_COMM_ID_BYTES, (const jbyte*)id); return out;
}
JNIEXPORT void JNICALL Java_EqpCountNative_commInit(JNIEnv* e, jclass k, jlong h, jbyteArray id, jint nRanks, jint rank) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; char buf[EQP_COMM_ID_BYTES]; (void)k;
  (*e)->GetByteArrayRegion(e, id, 0, EQP_COMM_ID_BYTES, (jbyte*)buf);
  if (eqp_comm_init(ctx, buf, nRanks, rank)) throw_io(e, ctx);
}
JNIEXPORT void JNICALL Java_EqpCountNative_accumulate(JNIEnv* e, jclass k, jlong h, jlong devAcc) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; (void)k;
  if (eqp_accumulate(ctx, (uint64_t*)(intptr_t)devAcc)) throw_io(e, ctx);
}
JNIEXPORT void JNICALL Java_EqpCountNative_reduceSample(JNIEnv* e, jclass k, jlong h, jlong devAcc, jint root) {
  eqp_ctx* ctx = (eqp_ctx*)(intptr_t)h; (void)k;
  if (eqp_reduce_sample(ctx, (uint64_t*)(intptr_t)devAcc, root)) throw_io(e, ctx);
}
JNIEXPORT void JNICALL Java_EqpCountNative_destroy(JNIEnv* e, jclass k, jlong h) { (void)e; (void)k; eqp_destroy((eqp_ctx*)(intptr_t)h); }
#endif /* EQP_HAVE_JNI */
