// ComputeCountsGpu — the one Java main that replaces lines 733-1037 of compute-counts.sh: it parses the options the replaced mains took
// (ComputeCounts.java:546, CombineSamFiles.java:575, ConvertSamBed.java:161), inflates the three .sam.gz streams the way `zcat | java ...`
// did (compute-intersection-bed-file.sh:644-662), feeds them to the GPU through EqpCountNative while they inflate, writes the three
// .cnt files exactly as ComputeCounts.java:875-950 does and prints the log lines getJobStatus.py / extractPipelineMetaData.py scrape.
// No counting happens in Java.
//
// java -Djava.library.path=$JAVA_DIR/native -cp classes ComputeCountsGpu [-g] [-e] [-j] [-n] [-u] [-U] [-C] [-N] [-s forward|backward]
//      [-W w] [-E n] [-J n] [-D slack] [-d device] -P <project dir> -B <file base> -R <radius> -t <T.sam.gz> -i <J.sam.gz> -G <G.sam.gz> -o <out base>
import java.io.*;
import java.util.zip.GZIPInputStream;

public class ComputeCountsGpu {
  static final String[] MODES = {"gene", "exon", "junction"};

  static InputStream open(String path) throws IOException {               // UtilLib.getBufferedReader: ".gz" => GZIPInputStream (UtilLib.java:322-383)
    InputStream in = new FileInputStream(path);
    return path.endsWith(".gz") ? new GZIPInputStream(in, 1 << 20) : in;
  }

  public static void main(String[] args) {
    int[] ip = {1, 0, 1, 5, 8, 0, 0, 0, 0, 0}; double[] dp = {0.01, 0.25};
    boolean nonZeroJunctions = false; int device = 0; String project = null, base = "ensembl_rna_hs", radius = "100", out = null;
    final String[] sam = new String[3];
    try {
      for (int i = 0; i < args.length; i++) {
        String a = args[i];
        if (a.equals("-g")) ip[9] |= 1; else if (a.equals("-e")) ip[9] |= 2; else if (a.equals("-j")) ip[9] |= 4;
        else if (a.equals("-n")) nonZeroJunctions = true; else if (a.equals("-u")) ip[5] = 1; else if (a.equals("-U")) ip[6] = 1;
        else if (a.equals("-C")) ip[7] = 1; else if (a.equals("-N")) ip[8] = 1;
        else if (a.equals("-s")) ip[1] = args[++i].equals("backward") ? 2 : 1;
        else if (a.equals("-W")) dp[0] = Double.parseDouble(args[++i]); else if (a.equals("-E")) ip[3] = Integer.parseInt(args[++i]);
        else if (a.equals("-J")) ip[4] = Integer.parseInt(args[++i]); else if (a.equals("-D")) dp[1] = Double.parseDouble(args[++i]);
        else if (a.equals("-d")) device = Integer.parseInt(args[++i]); else if (a.equals("-P")) project = args[++i];
        else if (a.equals("-B")) base = args[++i]; else if (a.equals("-R")) radius = args[++i];
        else if (a.equals("-t")) sam[0] = args[++i]; else if (a.equals("-i")) sam[1] = args[++i]; else if (a.equals("-G")) sam[2] = args[++i];
        else if (a.equals("-o")) out = args[++i];
        else throw new IOException("ERROR: unknown option " + a);
      }
      if (project == null || out == null || sam[0] == null || sam[1] == null || sam[2] == null || ip[9] == 0) throw new IOException("ERROR: -P, -o, -t, -i, -G and one of -g -e -j are required");
      ip[0] = sam[0].contains("-pe") ? 1 : 0;
      String ex = project + "/exon-pipeline-files", bed = ex + "/bed-files/" + base, map = ex + "/map-files/" + base;
      String[] files = {bed + "_" + radius + "_combined_exons.bed", bed + "_transcript_exons.bed", map + "_exon_gene.map", map + "_exon_exon.map",
                        gz(map + "_exon_junction.map"), gz(map + "_exon_junction_transcript.map"), gz(map + "_" + radius + "_junction_junction_coordinate.map"),
                        gz(map + "_equivalent_junctions.map"), ex + "/genome-files/genome.fa.fai"};
      final long ctx = EqpCountNative.create(device, files);
      long[] n = EqpCountNative.sizes(ctx);
      double[] gene = new double[(int) n[0]], exon = new double[(int) n[1]], junction = new double[(int) n[2]]; byte[] act = new byte[(int) n[3]];
      long[] sc = new long[EqpCountNative.N_SCALARS]; double[] sums = new double[3];
      // one inflate thread per input; the context is not thread-safe, so feed calls are serialised
      EqpCountNative.chunkBegin(ctx, 0, 0, 0);
      final Object lock = new Object(); final IOException[] failed = {null};
      // a .gz input smaller than 2 GiB (one direct buffer) goes to the device compressed and is inflated there; the others take the threads below
      final boolean[] onDevice = new boolean[3];
      for (int f = 0; f < 3; f++) {
        if (!sam[f].endsWith(".gz") || System.getenv("EQP_HOST_INFLATE") != null) continue;
        try (java.nio.channels.FileChannel ch = java.nio.channels.FileChannel.open(java.nio.file.Paths.get(sam[f]))) {
          if (ch.size() >= Integer.MAX_VALUE) continue;
          java.nio.ByteBuffer gz = java.nio.ByteBuffer.allocateDirect((int) ch.size());
          while (gz.hasRemaining() && ch.read(gz) > 0) { }
          onDevice[f] = EqpCountNative.chunkFeedGz(ctx, f, gz, ch.size());
        }
      }
      Thread[] th = new Thread[3];
      for (int f = 0; f < 3; f++) {
        final int input = f;
        th[f] = new Thread(() -> {
          if (onDevice[input]) return;
          byte[] buf = new byte[8 << 20];
          try (InputStream in = open(sam[input])) {
            for (int k; (k = in.read(buf)) > 0; ) synchronized (lock) { if (failed[0] != null) return; EqpCountNative.chunkFeed(ctx, input, buf, k); }
          } catch (IOException e) { synchronized (lock) { if (failed[0] == null) failed[0] = e; } }
        });
        th[f].start();
      }
      for (Thread t : th) t.join();
      if (failed[0] != null) throw failed[0];
      EqpCountNative.chunkEnd(ctx, ip, dp, gene, exon, junction, act, sc, sums);
      System.err.println("NUMBER_MAPPED_READS=" + sc[0]);                  // ConvertSamBed.java:266
      double[][] vec = {gene, exon, junction};
      for (int m = 0; m < 3; m++) {
        if ((ip[9] & (1 << m)) == 0) continue;
        String[] ids = EqpCountNative.rowIds(ctx, m); int[] tab = EqpCountNative.rowTable(ctx, m);
        boolean threeColumns = tab[3 * ids.length] == 3, outputZeroes = !(m == 2 && nonZeroJunctions);
        try (PrintWriter w = new PrintWriter(new BufferedWriter(new FileWriter(out + "-" + MODES[m] + ".cnt"), 1 << 20))) {
          for (int r = 0; r < ids.length; r++) {                           // ComputeCounts.java:875-950
            int feature = tab[3 * r], a = tab[3 * r + 1]; boolean always = tab[3 * r + 2] != 0;
            if (m == 2 && !(always || act[a] != 0)) continue;              // row is not in this chunk's filtered junction map
            double v = vec[m][feature];
            if (v != 0) w.println(ids[r] + "\t" + v);                      // Double.toString
            else if (outputZeroes) w.println(ids[r] + "\t" + (threeColumns ? "0.0" : "0"));
          }
        }
        System.err.format("Num weight objects in BED file: " + sc[2 + m] + ", num weight objects included: " + sc[5 + m] + " (%.2f%%), num weight objects excluded: " + sc[8 + m] + " (%.2f%%).%n",
                          sc[5 + m] * 100.0 / sc[2 + m], sc[8 + m] * 100.0 / sc[2 + m]);
        System.err.format("Sum of weights contributing to count objects: %.2f%n", sums[m]);
      }
      if ((ip[9] & 1) != 0) System.err.println("NUMBER_EXPRESSED_READS=" + sc[1]);   // ComputeCounts.java:991-993
      EqpCountNative.destroy(ctx);
    } catch (Exception e) {
      System.err.println(e.getMessage() == null ? "ERROR: " + e : e.getMessage());   // starts with ERROR: (getJobStatus.py:313-328)
      System.exit(1);
    }
  }
  static String gz(String plain) { return new File(plain).exists() || !new File(plain + ".gz").exists() ? plain : plain + ".gz"; }
}
