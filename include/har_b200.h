/*
 * har_b200.h — C ABI of the B200 (sm_100a) library behind the pyspark.ml surface.
 *
 * The reference (Lohitanvita/Activity-Recognition-Using-Apache-Spark) has no FFI of its own: its
 * hot path is reached from Main/main.py through py4j into Apache Spark MLlib 2.4 (JVM, not
 * vendored).  Each entry point below names the reference call site and the upstream MLlib
 * method it stands in for.  A maintainer binds these
 * with ctypes; no torch / C++ types cross the boundary.
 *
 * Conventions
 *   - every function returns 0 on success, a negative HAR_E_* code on failure; the message is
 *     available from har_last_error() (thread-local, valid until the next failing call).
 *   - pointers named d_* are DEVICE pointers on the current CUDA device; all others are host
 *     pointers.  The caller owns every buffer it passes in and keeps it alive until the call
 *     returns (calls that produce host-visible results synchronise the stream before returning;
 *     the others only enqueue work on `stream`).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - opaque handles are created / destroyed by paired calls; the library owns the device
 *     memory behind a handle.  Not re-entrant on one handle; thread-safe across handles.
 *   - integer outputs (bins, histogram counts, splits, predictions) are bit-exact w.r.t. the
 *     MLlib algorithm; fp64 gain arithmetic is evaluated without FMA contraction.
 */
#ifndef HAR_B200_H
#define HAR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HAR_OK 0
#define HAR_E_INVALID (-1)     /* bad argument (mirrors Spark's IllegalArgumentException cases) */
#define HAR_E_CUDA (-2)        /* CUDA runtime failure */
#define HAR_E_NOMEM (-3)
#define HAR_E_UNSUPPORTED (-4) /* valid in Spark, outside this library's limits (documented) */
#define HAR_E_COMM (-5)        /* the caller's all-reduce callback failed */

#define HAR_NUM_WINDOW_FEATURES 43  /* X0..9 Y0..9 Z0..9 XAVG YAVG ZAVG XPEAK YPEAK ZPEAK
                                       X/Y/ZABSDEV X/Y/ZSTDDEV RESULTANT: wisdm_data.csv:1 */
#define HAR_MAX_CLASSES 32
#define HAR_MAX_BINS 256

typedef void* har_stream_t;
typedef struct har_forest har_forest_t;

const char* har_last_error(void);
int har_version(void);
/* cudaGetDeviceCount / cudaSetDevice pass-throughs so a ctypes-only host needs no CUDA binding. */
int har_device_count(int32_t* count);
int har_set_device(int32_t device);
int har_stream_synchronize(har_stream_t stream);

/* har_forest_fit keeps its device scratch (row permutations, counters, node tables: ~3 GB at 5e5 rows x 100 trees,
 * ~14 GB at 5e6 rows) in a per-thread arena between calls, because re-allocating gigabytes per fit through the
 * stream-ordered allocator cost tens to hundreds of milliseconds.  This releases the calling thread's arena.
 * No reference counterpart (Spark's executors keep their own caches). */
void har_scratch_release(void);

/* ------------------------------------------------------------------------------------------------
 * (1)+(2) WISDM window transform.  No reference code exists for this stage: the
 * reference only ships its OUTPUT, Main/wisdm_main_ver_0.0/data/wisdm_data.csv (schema :1), read at
 * Main/main.py:18-20.  Definitions: SURVEY Appendix B / DESIGN.md.
 * ---------------------------------------------------------------------------------------------- */

/* a1 — segmentation.  A bout is a maximal run of equal (user, activity) (also broken where
 * ts[i]-ts[i-1] > gap_ns when gap_ns > 0 and d_ts != NULL); windows are consecutive non-overlapping
 * runs of `window_len` samples inside a bout, remainder dropped.  Writes the first-sample index of
 * every window in stream order to d_win_start (capacity n / window_len) and, when non-NULL, the
 * window's user / activity.  *num_windows is written on the host (the call synchronises). */
int har_window_segments(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                        int64_t n, int32_t window_len, int64_t gap_ns,
                        int64_t* d_win_start, int32_t* d_win_user, int32_t* d_win_activity,
                        int64_t* num_windows, har_stream_t stream);

/* har_window_segments without any allocation — this entry point has no counterpart in the reference
 * (or in Spark); it exists for hosts that are not allowed to allocate CUDA memory while running: a
 * process whose device memory is entirely owned by another allocator, a fixed-latency loop, a plug-in
 * sandbox limited to pre-approved buffers.  Same arguments and results as har_window_segments, but the
 * per-tile summaries live in a HAR_SEGMENTS_STATIC_BYTES array built into the library (held by every GPU
 * that loads it, for the life of the process) instead of coming from cudaMallocAsync, and the call
 * performs no cudaMalloc / cudaFree of any kind.
 *   - size limit: 512 + ceil(n / 2048) * (28 + 2 * (2048 / window_len + 1), rounded up to a multiple of 4)
 *     bytes must fit the array, i.e. 165,169,152 samples at window_len 200; larger inputs fail with
 *     HAR_E_UNSUPPORTED and a message that states the limit.
 *   - the first call on each device looks the array's address up; make that call at start-up, before
 *     any stream capture.
 *   - all calls on a device share the array: calls that can overlap on the GPU race and must be
 *     serialised by the caller.  The few words that must start at zero are cleared by each call (on the stream,
 *     without allocating); nothing else of an earlier call's contents is read. */
#define HAR_SEGMENTS_STATIC_BYTES (4u << 20)
int har_window_segments_static(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                               int64_t n, int32_t window_len, int64_t gap_ns,
                               int64_t* d_win_start, int32_t* d_win_user, int32_t* d_win_activity,
                               int64_t* num_windows, har_stream_t stream);

/* a2 — per-window features: 43 doubles per window, row-major with leading dimension ld_out
 * (>= 43), column order HAR_NUM_WINDOW_FEATURES above.  *PEAK is NaN when the axis has fewer than
 * three local maxima ('?' in the reference CSV).  n = samples in d_x/d_y/d_z/d_ts; every window must
 * lie inside [0, n).  window_len in [3, 256].  Asynchronous. */
int har_window_features(const double* d_x, const double* d_y, const double* d_z, const int64_t* d_ts,
                        int64_t n, const int64_t* d_win_start, int64_t num_windows, int32_t window_len,
                        double* d_out, int64_t ld_out, har_stream_t stream);

/* a1 + a2 in one call: har_window_segments followed by har_window_features over the windows it found, queued back
 * to back on `stream`.  With window_len 200 the feature kernel reads the window count from device memory and fills
 * d_win_user / d_win_activity itself, and the call returns as soon as the count has reached the host: *num_windows is
 * valid on return, the device outputs are complete in stream order (as with har_window_features).  d_win_start /
 * d_win_user / d_win_activity (the latter two may be NULL) and d_out must have room for n / window_len windows; rows
 * past *num_windows are left untouched.  Same definitions and limits as the two calls above.  No reference
 * counterpart beyond theirs (wisdm_data.csv:1 is this call's output). */
int har_window_transform(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                         const double* d_x, const double* d_y, const double* d_z, int64_t n, int32_t window_len,
                         int64_t gap_ns, int64_t* d_win_start, int32_t* d_win_user, int32_t* d_win_activity,
                         double* d_out, int64_t ld_out, int64_t* num_windows, har_stream_t stream);

/* Synthetic tri-axial stream of SURVEY §8(d) (bench / test input; counter-based so the CPU oracle
 * produces the same bits).  Bout table: host arrays, bout b covers global samples
 * [bout_start[b], bout_start[b+1]) of user bout_user[b] doing bout_activity[b]; `first_sample`
 * is the global index of local sample 0; bout_gain / bout_tilt scale the carrier and tilt the mean per
 * bout so that classes overlap.  tables: sine[1024], mu/amp/sigma[6*3], freq[6]. */
int har_synth_stream(int64_t n, int64_t first_sample, uint64_t seed,
                     const int64_t* bout_start, const int32_t* bout_user, const int32_t* bout_activity,
                     const int64_t* bout_ts0, const double* bout_gain, const double* bout_tilt, int64_t num_bouts,
                     const double* sine1024, const double* mu, const double* amp, const double* sigma,
                     const int32_t* freq,
                     int32_t* d_user, int32_t* d_activity, int64_t* d_ts,
                     double* d_x, double* d_y, double* d_z, har_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (3) Tree-ensemble fit: `dt.fit(train)` Main/main.py:297-300, `rf.fit(train)` :478-481 (and the
 * six fits inside each CrossValidator, :395, :576)  ->  ml.tree.impl.RandomForest.run.
 * ---------------------------------------------------------------------------------------------- */

typedef struct har_fit_params {
    int32_t num_trees;               /* 1 for DecisionTreeClassifier */
    int32_t max_depth;               /* 0..30 */
    int32_t max_bins;                /* 2..HAR_MAX_BINS */
    int32_t num_classes;             /* 2..HAR_MAX_CLASSES */
    int32_t min_instances_per_node;  /* >= 1 */
    int32_t impurity;                /* 0 gini, 1 entropy */
    int32_t features_per_node;       /* resolved featureSubsetStrategy; == num_features: all */
    int32_t reserved;
    double min_info_gain;
    double subsampling_rate;         /* (0, 1] */
    int64_t seed;
    int64_t max_memory_bytes;        /* maxMemoryInMB * 2^20 (node-group accounting) */
} har_fit_params;

/* Sum-all-reduce of `count` int32 values at device pointer d_buf, in place, across every rank,
 * ordered on `stream`.  Replaces `partitionAggregates.reduceByKey((a, b) => a.merge(b))` in
 * RandomForest.findBestSplits.  Return 0 on success. */
typedef int (*har_allreduce_i32_fn)(void* ctx, void* d_buf, int64_t count, har_stream_t stream);
/* With world_size > 1 the callback always carries the small exchanges (row counts, the findSplits sample).  For the
 * per-batch counters the library prefers a communicator of its own: if libnccl.so.2 is already loaded in every
 * rank's process it creates one over the same ranks (the unique id is exchanged once through this callback, and the
 * ranks agree through it that all of them got a communicator) and issues the NCCL calls on its own streams;
 * HAR_NCCL=0 in the environment, or no NCCL in some rank's process, keeps everything on the callback for every rank.
 * The callback must honour `stream` (the library may use streams of its own). */

typedef struct har_dist {
    int32_t rank, world_size;
    int64_t global_rows;             /* numExamples over all ranks */
    int32_t partition_offset;        /* global index of this rank's first partition */
    int32_t total_partitions;
    har_allreduce_i32_fn allreduce_i32;
    void* ctx;
    uint64_t group_id;               /* names the group of ranks behind the callback: fits that pass the same
                                        (group_id, world_size, rank) on one device share one transport decision and one
                                        library communicator; use a different value for a different set of processes */
} har_dist;

/* Which transport carries the per-batch exchanges of this group: 1 = the library's own NCCL communicator, 0 = the
 * caller's callback (the reason is copied to `why`, NUL-terminated, at most cap bytes), -1 = not decided yet (the
 * decision is made collectively inside the group's first exchange and is the same on every rank).
 * har_dist_release destroys the cached communicators (call it before the process group goes away).
 * No reference counterpart: Spark's executors own their shuffle connections. */
int har_dist_transport(const har_dist* dist, char* why, int32_t cap);
void har_dist_release(void);

/* a4 — RandomForest.findSplits / findSplitsForContinuousFeature: quantile thresholds of every
 * continuous feature (arity[f] <= 1) from the Bernoulli row sample MLlib draws.
 * d_X: row-major f64 [n][ld] (this rank's rows); part_off: host int64[num_parts+1] row offsets of
 * the Spark partitions inside this rank's rows.  thresholds: host f64 [F][max_bins] (row f holds
 * nsplits[f] ascending thresholds); nsplits: host int32 [F].  Synchronises. */
int har_find_splits(const double* d_X, int64_t n, int32_t num_features, int64_t ld,
                    const int32_t* arity, int32_t max_bins, int64_t seed,
                    const int64_t* part_off, int32_t num_parts, const har_dist* dist,
                    double* thresholds, int32_t* nsplits, har_stream_t stream);

/* a5 — TreePoint.convertToTreeRDD / findBin: bin index per (row, feature) as u8,
 * d_bins[i * ldb + f]; continuous = java.util.Arrays.binarySearch lower bound, categorical = value
 * (error HAR_E_INVALID with Spark's message when outside [0, arity)).  Synchronises. */
int har_bin_features(const double* d_X, int64_t n, int32_t num_features, int64_t ld,
                     const int32_t* arity, const double* thresholds, const int32_t* nsplits,
                     int32_t max_bins, uint8_t* d_bins, int64_t ldb, har_stream_t stream);

/* a6 — BaggedPoint.convertToBaggedRDD: per-(tree, row) sample weight, d_weights[t * n + i] (u8).
 * Poisson(rate) from commons-math3 Well19937c seeded seed + partitionIndex + 1 when num_trees > 1,
 * Bernoulli(rate) from XORShiftRandom otherwise, all ones when num_trees == 1 && rate == 1. */
int har_bag_weights(int64_t n, int32_t num_trees, double subsampling_rate, int64_t seed,
                    const int64_t* part_off, int32_t num_parts, int32_t partition_offset,
                    uint8_t* d_weights, har_stream_t stream);

/* a8 — one node's DTStatsAggregator: hist[hist_off[f] + bin * C + label] += weight over the listed
 * rows (d_rows == NULL: rows 0..nrows-1; d_weights == NULL: weight 1), parent[label] += weight.
 * hist_off: host int64 [F+1].  d_hist (uint32 [hist_off[F]]) and d_parent (uint32 [C]) are
 * accumulated into (zero them first).  Asynchronous. */
int har_node_histogram(const uint8_t* d_bins, int64_t ldb, const uint8_t* d_labels,
                       const uint8_t* d_weights, const int32_t* d_rows, int64_t nrows,
                       int32_t num_features, int32_t num_classes, const int64_t* hist_off,
                       uint32_t* d_hist, uint32_t* d_parent, har_stream_t stream);

/* a3..a11 — the whole of RandomForest.run on this rank's rows: metadata, split candidates, binning,
 * bagging, node-group loop (histograms -> all-reduce -> best split -> row re-partition), pruning.
 * d_X row-major f64 [n][ld]; d_y f64 [n] labels in [0, num_classes); arity: host int32 [F]
 * (0/1 continuous, k >= 2 categorical with k values; NULL = all continuous).
 * d_weights: optional explicit bag weights u8 [num_trees][n] (NULL: drawn as har_bag_weights).
 * dist == NULL: single GPU.  Synchronous from the host's view. */
int har_forest_fit(const double* d_X, int64_t n, int32_t num_features, int64_t ld, const double* d_y,
                   const int32_t* arity, const int64_t* part_off, int32_t num_parts,
                   const har_fit_params* params, const har_dist* dist, const uint8_t* d_weights,
                   har_stream_t stream, har_forest_t** out);

/* Model read-back (ml.tree.Node after LearningNode.toNode(prune = true)).  Flat,
 * pre-order per tree; indices in left/right are relative to the tree's first node; feature -1 =
 * leaf.  arity 0 = continuous split (threshold), k >= 2 = categorical (left_mask bit v = category v
 * goes left). */
int har_forest_shape(const har_forest_t* f, int32_t* num_trees, int32_t* num_classes,
                     int32_t* num_features, int64_t* total_nodes);
int har_forest_export(const har_forest_t* f, int64_t* tree_off, int32_t* feature, int32_t* left,
                      int32_t* right, int32_t* node_id, int32_t* arity, double* threshold,
                      uint64_t* left_mask, double* prediction, double* impurity, double* gain,
                      double* stats /* [total][C] */);
/* Training by-products for stage-level parity checks: thresholds [F][max_bins], nsplits [F],
 * number of node groups / device batches processed.  Any pointer may be NULL. */
int har_forest_fit_info(const har_forest_t* f, double* thresholds, int32_t* nsplits, int32_t max_bins,
                        int64_t* num_groups, int64_t* num_batches, int64_t* num_kernel_launches);
/* Rebuild a device-resident forest from the flat arrays (e.g. a model loaded from disk). */
int har_forest_import(int32_t num_trees, int32_t num_classes, int32_t num_features,
                      const int64_t* tree_off, const int32_t* feature, const int32_t* left,
                      const int32_t* right, const int32_t* arity, const double* threshold,
                      const uint64_t* left_mask, const double* stats, har_forest_t** out);
void har_forest_free(har_forest_t* f);

/* ------------------------------------------------------------------------------------------------
 * (4) Batched predict: `dtModel.transform(test_data)` Main/main.py:305, `rfModel.transform` :486
 * -> {DecisionTree,RandomForest}ClassificationModel.predictRaw / raw2probabilityInPlace /
 * raw2prediction.  One row per thread; d_raw / d_prob f64 [n][C], d_pred f64 [n];
 * any output may be NULL.  is_forest = 0: DT semantics (raw = leaf class counts), 1: RF (raw = sum
 * over trees, in tree order, of leaf counts / leaf total).  Asynchronous.
 * ---------------------------------------------------------------------------------------------- */
int har_forest_predict(const har_forest_t* f, const double* d_X, int64_t n, int64_t ld,
                       int32_t is_forest, double* d_raw, double* d_prob, double* d_pred,
                       har_stream_t stream);

/* f4 — raw WISDM text ingest (no call site in Main/main.py, which starts from the already
 * transformed wisdm_data.csv at main.py:18-26 — this is the step ahead of a1/a2 for the raw release
 * WISDM_ar_v1.1_raw.txt).  d_text: device bytes, records "user,activity,timestamp,x,y,z;" (one or several per line).
 * Blanks around records and fields are ignored, empty records vanish, a record that does not have six non-empty
 * fields of the right kinds (int user, one of `activity_names` -> its index, int64 timestamp, three decimals) is
 * dropped and counted in *num_dropped; text after the last ';' is one more record.  Decimals are converted exactly (what
 * strtod / Double.parseDouble return) when they have <= 15 significant digits and a decimal exponent in [-22, 22];
 * anything longer fails with HAR_E_UNSUPPORTED instead of being approximated.  Outputs: the six columns of
 * har_window_segments / har_window_features, *num_records of them (HAR_E_INVALID if > capacity; call with
 * capacity 0 and NULL columns to size).  Synchronises. */
int har_parse_wisdm_raw(const char* d_text, int64_t nbytes, const char* const* activity_names,
                        int32_t num_names, int64_t capacity, int32_t* d_user, int32_t* d_activity,
                        int64_t* d_ts, double* d_x, double* d_y, double* d_z, int64_t* num_records,
                        int64_t* num_dropped, har_stream_t stream);
/* Upper bound on the records of such a text (terminators + 1), to size the columns before parsing once.  Synchronises. */
int har_count_wisdm_raw(const char* d_text, int64_t nbytes, int64_t* max_records, har_stream_t stream);

/* CSR -> dense row-major f64 (the reference's `features` column is a 3100-d SparseVector,
 * result.txt:110).  d_out [n][ld] is fully overwritten.  Asynchronous. */
int har_csr_to_dense(const int64_t* d_indptr, const int32_t* d_indices, const double* d_values,
                     int64_t n, int32_t num_features, double* d_out, int64_t ld, har_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HAR_B200_H */
