// SparkGolden.java — run the REAL Spark MLlib 2.4.6 RandomForestClassifier on this repository's synthetic inputs and
// dump everything needed to pin parity: toDebugString, every node of every tree (pre-order: prediction, impurity,
// impurityStats, gain, split) and the first 64 rows' rawPrediction / probability / prediction.
//
// SOURCE ONLY: the image this engine was built in has no JDK and no Spark jars (SURVEY.md §0.3), so this file has never
// been compiled here.  With Spark 2.4.6 (the version pinned by the reference, pom.xml:51-61) on the classpath:
//
//   javac -cp "$SPARK_HOME/jars/*" tools/SparkGolden.java -d /tmp/golden_cls
//   java  -cp "/tmp/golden_cls:$SPARK_HOME/jars/*" SparkGolden draws 1500 10 0 20 6 8 207336481 /tmp/golden_cfg1
//   java  -cp "/tmp/golden_cls:$SPARK_HOME/jars/*" SparkGolden flat 200000 64 50 6 8 96 207336481 /tmp/golden_flat
//   python tools/compare_spark_golden.py /tmp/golden_cfg1        # on a box with a B200: fits the same Dataset, diffs
//
// Inputs follow SURVEY.md Appendix C exactly as euromillioner_b200/synth.py does (index-addressable splitmix64):
//   draws <n_draws> <lag> <ball> <trees> <depth> <partitions> <seed> <out>     draw-history rows (cfg1 / cfg4)
//   flat  <rows> <features> <classes> <trees> <depth> <partitions> <seed> <out>  flat rows (cfg2 / cfg3 / cfg5)
// The partition layout is part of the contract (bagging and the split sample are seeded per partition): rows go through
// JavaSparkContext.parallelize(list, P), whose slices are [floor(i*N/P), floor((i+1)*N/P)) — the layout
// Dataset(num_partitions=P) describes on the engine's side.
import java.io.PrintWriter;
import java.util.ArrayList;
import java.util.List;

import org.apache.spark.api.java.JavaRDD;
import org.apache.spark.api.java.JavaSparkContext;
import org.apache.spark.ml.attribute.NominalAttribute;
import org.apache.spark.ml.classification.DecisionTreeClassificationModel;
import org.apache.spark.ml.classification.RandomForestClassificationModel;
import org.apache.spark.ml.classification.RandomForestClassifier;
import org.apache.spark.ml.linalg.SQLDataTypes;
import org.apache.spark.ml.linalg.Vector;
import org.apache.spark.ml.linalg.Vectors;
import org.apache.spark.ml.tree.ContinuousSplit;
import org.apache.spark.ml.tree.InternalNode;
import org.apache.spark.ml.tree.Node;
import org.apache.spark.sql.Dataset;
import org.apache.spark.sql.Row;
import org.apache.spark.sql.RowFactory;
import org.apache.spark.sql.SparkSession;
import org.apache.spark.sql.types.DataTypes;
import org.apache.spark.sql.types.Metadata;
import org.apache.spark.sql.types.StructField;
import org.apache.spark.sql.types.StructType;

public final class SparkGolden {
  // z = seed + 0x9E3779B97F4A7C15 * (idx + 1); two xorshift-multiply rounds (all arithmetic mod 2^64)
  static long splitmix64(long seed, long idx) {
    long z = seed + 0x9E3779B97F4A7C15L * (idx + 1);
    z = (z ^ (z >>> 30)) * 0xBF58476D1CE4E5B9L;
    z = (z ^ (z >>> 27)) * 0x94D049BB133111EBL;
    return z ^ (z >>> 31);
  }

  static long umod(long v, long m) { return Long.remainderUnsigned(v, m); }

  // synth.draws: 5 sorted distinct mains in 1..50 and 2 sorted distinct stars in 1..12 per draw (partial Fisher-Yates)
  static int[][] draws(int nDraws, long seed) {
    int[][] out = new int[nDraws][7];
    for (int d = 0; d < nDraws; d++) {
      int[] pool = new int[50];
      for (int i = 0; i < 50; i++) pool[i] = i + 1;
      for (int j = 0; j < 5; j++) {
        int k = j + (int) umod(splitmix64(seed, 16L * d + j), 50 - j);
        int t = pool[j]; pool[j] = pool[k]; pool[k] = t;
      }
      int[] m = java.util.Arrays.copyOf(pool, 5);
      java.util.Arrays.sort(m);
      int[] sp = new int[12];
      for (int i = 0; i < 12; i++) sp[i] = i + 1;
      for (int j = 0; j < 2; j++) {
        int k = j + (int) umod(splitmix64(seed, 16L * d + 8 + j), 12 - j);
        int t = sp[j]; sp[j] = sp[k]; sp[k] = t;
      }
      int[] s = java.util.Arrays.copyOf(sp, 2);
      java.util.Arrays.sort(s);
      System.arraycopy(m, 0, out[d], 0, 5);
      System.arraycopy(s, 0, out[d], 5, 2);
    }
    return out;
  }

  public static void main(String[] a) throws Exception {
    final String mode = a[0];
    final List<Row> rows = new ArrayList<>();
    int numClasses, trees, depth, parts;
    long seed;
    String out;
    if (mode.equals("draws")) {
      int nDraws = Integer.parseInt(a[1]), lag = Integer.parseInt(a[2]), ball = Integer.parseInt(a[3]);
      trees = Integer.parseInt(a[4]); depth = Integer.parseInt(a[5]); parts = Integer.parseInt(a[6]);
      seed = Long.parseLong(a[7]); out = a[8];
      int[][] b = draws(nDraws, 20200614L);
      numClasses = ball < 5 ? 50 : 12;
      for (int t = lag; t < nDraws; t++) {
        double[] x = new double[7 * lag];
        for (int l = 1; l <= lag; l++) for (int q = 0; q < 7; q++) x[(l - 1) * 7 + q] = b[t - l][q];
        rows.add(RowFactory.create((double) (b[t][ball] - 1), Vectors.dense(x)));
      }
    } else {
      int n = Integer.parseInt(a[1]), f = Integer.parseInt(a[2]);
      numClasses = Integer.parseInt(a[3]);
      trees = Integer.parseInt(a[4]); depth = Integer.parseInt(a[5]); parts = Integer.parseInt(a[6]);
      seed = Long.parseLong(a[7]); out = a[8];
      final long dataSeed = 42L;
      for (long i = 0; i < n; i++) {
        double[] x = new double[f];
        for (int j = 0; j < f; j++) x[j] = 1 + (int) umod(splitmix64(dataSeed, i * f + j), 50);
        long e = umod(splitmix64(dataSeed ^ 0xA5A5A5A5L, i), 5);
        double y = (double) (((long) x[0] + (f > 3 ? (long) x[3] : 0L) + e) % numClasses);
        rows.add(RowFactory.create(y, Vectors.dense(x)));
      }
    }
    SparkSession spark = SparkSession.builder().master("local[*]").appName("SparkGolden").getOrCreate();
    JavaSparkContext jsc = new JavaSparkContext(spark.sparkContext());
    JavaRDD<Row> rdd = jsc.parallelize(rows, parts);
    // the label column carries numClasses as nominal metadata, as a StringIndexer would have left it
    Metadata labelMeta = NominalAttribute.defaultAttr().withName("label").withNumValues(numClasses).toMetadata();
    StructType schema = new StructType(new StructField[] {
        new StructField("label", DataTypes.DoubleType, false, labelMeta),
        new StructField("features", SQLDataTypes.VectorType(), false, Metadata.empty())});
    Dataset<Row> df = spark.createDataFrame(rdd, schema);
    RandomForestClassifier rf = new RandomForestClassifier().setNumTrees(trees).setMaxDepth(depth).setMaxBins(32).setSeed(seed);
    RandomForestClassificationModel model = rf.fit(df);

    new java.io.File(out).mkdirs();
    try (PrintWriter w = new PrintWriter(out + "/debug_string.txt", "UTF-8")) { w.print(model.toDebugString()); }
    try (PrintWriter w = new PrintWriter(out + "/params.txt", "UTF-8")) {
      w.println(String.join(" ", a));
      w.println("uid " + model.uid());
      w.println("numClasses " + model.numClasses() + " numFeatures " + model.numFeatures() + " totalNumNodes " + model.totalNumNodes());
      w.println("sparkVersion " + spark.version());
    }
    // nodes, pre-order, one line each: tree index | L or I | prediction | impurity | gain | feature | threshold | stats...
    try (PrintWriter w = new PrintWriter(out + "/nodes.txt", "UTF-8")) {
      DecisionTreeClassificationModel[] ts = model.trees();
      for (int t = 0; t < ts.length; t++) dump(w, t, ts[t].rootNode());
    }
    try (PrintWriter w = new PrintWriter(out + "/predictions.txt", "UTF-8")) {
      for (Row r : model.transform(df).select("rawPrediction", "probability", "prediction").takeAsList(64)) {
        w.println(vec((Vector) r.get(0)) + " | " + vec((Vector) r.get(1)) + " | " + r.getDouble(2));
      }
    }
    try (PrintWriter w = new PrintWriter(out + "/feature_importances.txt", "UTF-8")) { w.println(vec(model.featureImportances())); }
    spark.stop();
  }

  static String vec(Vector v) {
    StringBuilder sb = new StringBuilder();
    for (double d : v.toArray()) sb.append(sb.length() == 0 ? "" : " ").append(Double.toString(d));
    return sb.toString();
  }

  // Node.impurityStats is private[ml] in Scala, i.e. a public method in the class file: callable from Java
  static void dump(PrintWriter w, int tree, Node n) {
    StringBuilder sb = new StringBuilder();
    boolean internal = n instanceof InternalNode;
    sb.append(tree).append(internal ? " I " : " L ").append(n.prediction()).append(' ').append(n.impurity()).append(' ');
    if (internal) {
      InternalNode in = (InternalNode) n;
      ContinuousSplit s = (ContinuousSplit) in.split();
      sb.append(in.gain()).append(' ').append(s.featureIndex()).append(' ').append(s.threshold());
    } else {
      sb.append("0.0 -1 0.0");
    }
    for (double d : n.impurityStats().stats()) sb.append(' ').append(d);
    w.println(sb);
    if (internal) {
      dump(w, tree, ((InternalNode) n).leftChild());
      dump(w, tree, ((InternalNode) n).rightChild());
    }
  }
}
