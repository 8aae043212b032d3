// EqpCountPanama — the same binding without C glue, on java.lang.foreign (JDK >= 22): downcall handles straight onto libeqpcount.so.
// Struct layouts follow include/eqp_count.h (tests/test_oracle_and_host.py pins sizeof(eqp_params) == 64 and the ctypes mirror of eqp_counts).
import java.lang.foreign.*;
import java.lang.invoke.MethodHandle;
import static java.lang.foreign.ValueLayout.*;

public final class EqpCountPanama {
  static final Linker LINKER = Linker.nativeLinker();
  static final SymbolLookup LIB = SymbolLookup.libraryLookup("libeqpcount.so", Arena.global());
  static MethodHandle fn(String name, FunctionDescriptor d) { return LINKER.downcallHandle(LIB.find(name).orElseThrow(), d); }

  static final MethodHandle CREATE = fn("eqp_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS));
  static final MethodHandle LOAD = fn("eqp_load_annotation", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
  static final MethodHandle DEFAULTS = fn("eqp_default_params", FunctionDescriptor.ofVoid(ADDRESS));
  static final MethodHandle COUNT = fn("eqp_count_chunk", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, JAVA_INT, ADDRESS));
  static final MethodHandle BEGIN = fn("eqp_chunk_begin", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_LONG, JAVA_LONG));
  static final MethodHandle FEED = fn("eqp_chunk_feed", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_LONG));
  static final MethodHandle END = fn("eqp_chunk_end", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
  // one complete .sam.gz member (a mapped or direct segment), inflated on the device; returns EQP_ERR_UNSUPPORTED (-5) for files to inflate on the host
  static final MethodHandle FEED_GZ = fn("eqp_chunk_feed_gz", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_LONG));
  static final MethodHandle COUNT_GZ = fn("eqp_count_chunk_gz", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS));
  static final MethodHandle ACCUMULATE = fn("eqp_accumulate", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
  static final MethodHandle COMM_ID = fn("eqp_comm_id", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  static final MethodHandle COMM_INIT = fn("eqp_comm_init", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT));
  static final MethodHandle REDUCE = fn("eqp_reduce_sample", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT));
  static final MethodHandle LAST_ERROR = fn("eqp_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
  static final MethodHandle DESTROY = fn("eqp_destroy", FunctionDescriptor.ofVoid(ADDRESS));

  // eqp_params: 5 x int32, 4 bytes padding, double, 4 x int32, double, 2 x int32 = 64 bytes
  static final StructLayout PARAMS = MemoryLayout.structLayout(JAVA_INT.withName("paired_end"), JAVA_INT.withName("stranded"), JAVA_INT.withName("gene_overlap"),
      JAVA_INT.withName("exon_overlap"), JAVA_INT.withName("junction_overlap"), MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("read_weight_threshold"),
      JAVA_INT.withName("unweighted"), JAVA_INT.withName("unambiguous"), JAVA_INT.withName("containment"), JAVA_INT.withName("nonsplice_conforming"),
      JAVA_DOUBLE.withName("distance_slack"), JAVA_INT.withName("modes"), JAVA_INT.withName("reserved"));
  // eqp_counts: 3 x int64 sizes, 3 x double*, 3 x uint64*, uint8* junction_active, int64 n_junction_act, 2 x int64, 4 x int64[3], double[3], 2 x double, int64
  static final StructLayout COUNTS = MemoryLayout.structLayout(JAVA_LONG.withName("n_gene"), JAVA_LONG.withName("n_exon"), JAVA_LONG.withName("n_junction"),
      ADDRESS.withName("gene"), ADDRESS.withName("exon"), ADDRESS.withName("junction"), ADDRESS.withName("gene_fixed"), ADDRESS.withName("exon_fixed"),
      ADDRESS.withName("junction_fixed"), ADDRESS.withName("junction_active"), JAVA_LONG.withName("n_junction_act"), JAVA_LONG.withName("number_mapped_reads"),
      JAVA_LONG.withName("number_expressed_reads"), MemoryLayout.sequenceLayout(3, JAVA_LONG).withName("num_weight_objects"),
      MemoryLayout.sequenceLayout(3, JAVA_LONG).withName("num_weight_objects_included"), MemoryLayout.sequenceLayout(3, JAVA_LONG).withName("num_weight_objects_excluded"),
      MemoryLayout.sequenceLayout(3, JAVA_LONG).withName("n_records"), MemoryLayout.sequenceLayout(3, JAVA_DOUBLE).withName("sum_weights"),
      JAVA_DOUBLE.withName("device_ms"), JAVA_DOUBLE.withName("dominant_kernel_ms"), JAVA_LONG.withName("kernel_launches"));

  static void check(int rc, MemorySegment ctx) throws Throwable {
    if (rc != 0) throw new java.io.IOException("ERROR: " + ((MemorySegment) LAST_ERROR.invokeExact(ctx)).reinterpret(1024).getString(0));
  }

  /** count one chunk whose inflated texts sit in native segments; returns the gene vector as an example of reading eqp_counts */
  public static double[] countChunk(MemorySegment ctx, MemorySegment t, MemorySegment j, MemorySegment g, Arena arena) throws Throwable {
    MemorySegment p = arena.allocate(PARAMS), c = arena.allocate(COUNTS);
    DEFAULTS.invokeExact(p);
    check((int) COUNT.invokeExact(ctx, p, t, t.byteSize(), j, j.byteSize(), g, g.byteSize(), 0, c), ctx);
    long n = c.get(JAVA_LONG, COUNTS.byteOffset(MemoryLayout.PathElement.groupElement("n_gene")));
    MemorySegment gene = c.get(ADDRESS, COUNTS.byteOffset(MemoryLayout.PathElement.groupElement("gene"))).reinterpret(n * 8);
    return gene.toArray(JAVA_DOUBLE);
  }
}
