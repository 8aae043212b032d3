// EqpCountNative — JNI face of libeqpcount.so (include/eqp_count.h); glue: bindings/jni/eqpcount_jni.c.
// Goes next to the mains it serves: exon-pipeline-scripts/java/source/bed-source/.  Every failure surfaces as an IOException whose
// message starts with "ERROR:", which ComputeCountsGpu prints before System.exit(1) (ComputeCounts.java:996-999).
import java.io.IOException;
import java.nio.ByteBuffer;

public final class EqpCountNative {
  static { System.loadLibrary("eqpcount_jni"); }
  private EqpCountNative() {}

  /** intParams: paired_end, stranded (0 none, 1 forward, 2 backward), gene_overlap, exon_overlap, junction_overlap, unweighted,
   *  unambiguous, containment, nonsplice_conforming, modes (bit0 gene, bit1 exon, bit2 junction).  doubleParams: read_weight_threshold, distance_slack. */
  public static final int N_INT_PARAMS = 10, N_DOUBLE_PARAMS = 2;
  /** scalars: number_mapped_reads, number_expressed_reads, num_weight_objects[3], included[3], excluded[3], kernel_launches */
  public static final int N_SCALARS = 12;

  /** eqp_create + eqp_load_annotation; annotationFiles = the 9 paths in the order of eqp_annotation_files */
  public static native long create(int device, String[] annotationFiles) throws IOException;
  /** n_gene, n_exon, n_junction, n_junction_act: sizes of the result arrays */
  public static native long[] sizes(long ctx);
  public static native void countChunk(long ctx, int[] intParams, double[] doubleParams, ByteBuffer t, long tLen, ByteBuffer j, long jLen, ByteBuffer g, long gLen,
                                       double[] gene, double[] exon, double[] junction, byte[] junctionActive, long[] scalars, double[] sumWeights) throws IOException;
  public static native void prefetchChunk(long ctx, ByteBuffer t, long tLen, ByteBuffer j, long jLen, ByteBuffer g, long gLen) throws IOException;
  public static native void chunkBegin(long ctx, long tHint, long jHint, long gHint) throws IOException;
  public static native void chunkFeed(long ctx, int input, byte[] buf, int len) throws IOException;
  /** One complete .sam.gz member in a direct buffer, inflated on the device; false: inflate it on the host (GZIPInputStream) and use chunkFeed. */
  public static native boolean chunkFeedGz(long ctx, int input, ByteBuffer gz, long len) throws IOException;
  public static native void chunkEnd(long ctx, int[] intParams, double[] doubleParams, double[] gene, double[] exon, double[] junction, byte[] junctionActive,
                                     long[] scalars, double[] sumWeights) throws IOException;
  public static native String[] rowIds(long ctx, int mode);
  /** 3 ints per row (feature index, activation index, always flag), then the number of map columns */
  public static native int[] rowTable(long ctx, int mode);
  public static native byte[] commId() throws IOException;
  public static native void commInit(long ctx, byte[] id, int nRanks, int rank) throws IOException;
  public static native void accumulate(long ctx, long deviceAccumulatorWide) throws IOException;
  public static native void reduceSample(long ctx, long deviceAccumulatorWide, int root) throws IOException;
  public static native void destroy(long ctx);
}
