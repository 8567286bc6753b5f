/*
 * har_oracle.c — CPU ORACLE.  TEST INFRASTRUCTURE ONLY: imported by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, never by
 * the product path (activity-recognition-using-apache-spark_b200/).
 *
 * Plain-C restatement of the hot path of Lohitanvita/Activity-Recognition-Using-Apache-Spark:
 * the tree-ensemble fit/predict that Main/main.py:297-305 and :478-486 delegate to Apache
 * Spark MLlib (org.apache.spark:spark-mllib_2.11:2.4.x, NOT vendored in the reference; + the
 * commons-math3 3.4.1 Poisson/Well19937c sampler), and of the WISDM window transform whose
 * output schema is Main/wisdm_main_ver_0.0/data/wisdm_data.csv:1.
 *
 * Parity status
 *   - tree path, T=1 / all features / gini / continuous + binary-unordered features:
 *     PINNED by the reference's own Spark output (result.txt:231 "depth 3 with 15 nodes",
 *     :268-270 1187 of 1625 correct, additional_param.csv:3 F1 0.6795556960736588, ...);
 *     see tests/test_oracle_golden.py.
 *   - bagging stream, feature-subset draws, ordered-categorical splits, entropy, pruning
 *     beyond the golden tree: restated from the published MLlib 2.4 algorithm, parity unpinned
 *     (no JVM on this box).
 *   - window features: parity unpinned (the reference ships neither code nor raw samples for
 *     that stage; definitions in SURVEY.md Appendix B / DESIGN.md are the spec).
 *
 * Each function cites the upstream file/method it follows (branch-2.4 paths; SURVEY.md
 * Appendix A section in brackets).  Everything is IEEE fp64 compiled with -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * JVM helpers
 * ---------------------------------------------------------------------------------------- */

/* java.lang.Double.compare: -0.0 < 0.0, NaN greatest and equal to itself */
static int jdcmp(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t ab, bb;
    if (a != a) ab = 0x7ff8000000000000LL; else memcpy(&ab, &a, 8);
    if (b != b) bb = 0x7ff8000000000000LL; else memcpy(&bb, &b, 8);
    return ab == bb ? 0 : (ab < bb ? -1 : 1);
}

/* java.util.Arrays.binarySearch(double[], key) folded with TreePoint.findBin's
 * `if (idx >= 0) idx else -idx - 1`  [A.5] */
static int java_find_bin(const double* thr, int n, double key) {
    int low = 0, high = n - 1;
    while (low <= high) {
        int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
        int c = jdcmp(thr[mid], key);
        if (c < 0) low = mid + 1;
        else if (c > 0) high = mid - 1;
        else return mid;
    }
    return low;
}

/* scala.util.hashing.MurmurHash3.bytesHash == MurmurHash3_x86_32  [A.4] */
static uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static uint32_t murmur3_bytes(const uint8_t* d, int len, uint32_t seed) {
    uint32_t h = seed;
    int i = 0;
    for (; i + 4 <= len; i += 4) {
        uint32_t k = (uint32_t)d[i] | ((uint32_t)d[i + 1] << 8) | ((uint32_t)d[i + 2] << 16) | ((uint32_t)d[i + 3] << 24);
        k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
        h ^= k; h = rotl32(h, 13); h = h * 5 + 0xe6546b64u;
    }
    uint32_t k = 0;
    int rem = len - i;
    if (rem == 3) k ^= (uint32_t)d[i + 2] << 16;
    if (rem >= 2) k ^= (uint32_t)d[i + 1] << 8;
    if (rem >= 1) { k ^= d[i]; k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u; h ^= k; }
    h ^= (uint32_t)len;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}

/* org.apache.spark.util.random.XORShiftRandom  [A.4]; hashSeed hashes a 64-BYTE buffer */
typedef struct { uint64_t s; } xorshift_t;
static void xs_seed(xorshift_t* r, int64_t seed) {
    uint8_t buf[64];
    memset(buf, 0, sizeof buf);
    for (int i = 0; i < 8; i++) buf[i] = (uint8_t)((uint64_t)seed >> (56 - 8 * i));
    uint32_t lo = murmur3_bytes(buf, 64, 0x3c074a61u);
    uint32_t hi = murmur3_bytes(buf, 64, lo);
    r->s = ((uint64_t)hi << 32) | lo;
}
static uint32_t xs_next(xorshift_t* r, int bits) {
    uint64_t s = r->s;
    s ^= s << 21; s ^= s >> 35; s ^= s << 4;
    r->s = s;
    return (uint32_t)(s & ((1ULL << bits) - 1));
}
static double xs_double(xorshift_t* r) {
    uint64_t a = xs_next(r, 26), b = xs_next(r, 27);
    return (double)((a << 27) + b) * 0x1.0p-53;
}
static int32_t xs_int(xorshift_t* r) { return (int32_t)xs_next(r, 32); }

/* java.util.Random  [A.7] */
typedef struct { uint64_t s; } jrand_t;
static void jr_seed(jrand_t* r, int64_t seed) { r->s = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
static int32_t jr_next(jrand_t* r, int bits) {
    r->s = (r->s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)(int64_t)(r->s >> (48 - bits));
}
static int64_t jr_long(jrand_t* r) {
    int64_t hi = jr_next(r, 32), lo = jr_next(r, 32);
    return (int64_t)((uint64_t)hi << 32) + lo;
}

/* org.apache.commons.math3.random.Well19937c (3.4.1)  [A.6] */
typedef struct { uint32_t v[624]; int index; } well_t;
static void well_seed(well_t* w, int64_t seed) {
    uint32_t s0 = (uint32_t)((uint64_t)seed >> 32), s1 = (uint32_t)((uint64_t)seed & 0xffffffffu);
    w->v[0] = s0; w->v[1] = s1;
    for (int i = 2; i < 624; i++) {
        int64_t l = (int64_t)(int32_t)w->v[i - 2];                 /* sign-extended, as in setSeed(int[]) */
        w->v[i] = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
    }
    w->index = 0;
}
static uint32_t well_next(well_t* w, int bits) {
    /* AbstractWell tables for K=19937: r=624, m1=70, m2=179, m3=449 */
    int idx = w->index;
    int iRm1 = (idx + 623) % 624, iRm2 = (idx + 622) % 624;
    int i1 = (idx + 70) % 624, i2 = (idx + 179) % 624, i3 = (idx + 449) % 624;
    uint32_t v0 = w->v[idx], vM1 = w->v[i1], vM2 = w->v[i2], vM3 = w->v[i3];
    uint32_t z0 = (0x80000000u & w->v[iRm1]) ^ (0x7FFFFFFFu & w->v[iRm2]);
    uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
    uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
    uint32_t z3 = z1 ^ z2;
    uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
    w->v[idx] = z3;
    w->v[iRm1] = z4;
    w->v[iRm2] &= 0x80000000u;
    w->index = iRm1;
    z4 ^= (z4 << 7) & 0xe46e1700u;
    z4 ^= (z4 << 15) & 0x9b868000u;
    return z4 >> (32 - bits);
}
static double well_double(well_t* w) {
    uint64_t hi = (uint64_t)well_next(w, 26) << 26;
    uint64_t lo = well_next(w, 26);
    return (double)(hi | lo) * 0x1.0p-52;
}
/* PoissonDistribution.sample() for mean < 40 (pivot)  [A.6] */
static int poisson_small(well_t* w, double mean, double p /* = exp(-mean) */) {
    long n = 0; double r = 1.0;
    while ((double)n < 1000.0 * mean) {
        r *= well_double(w);
        if (r >= p) n++; else return (int)n;
    }
    return (int)n;
}

/* ------------------------------------------------------------------------------------------
 * DecisionTreeMetadata.buildMetadata  [A.5]   (ml/tree/impl/DecisionTreeMetadata.scala)
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    int32_t num_trees, max_depth, max_bins, num_classes, min_instances_per_node, impurity;
    int32_t features_per_node;       /* resolved count; == F means no feature subsampling */
    int32_t reserved;
    double min_info_gain, subsampling_rate;
    int64_t seed, max_memory_bytes;
} orc_params;

typedef struct {
    int F, C, max_possible_bins;
    int* arity;        /* 0 continuous */
    int* unordered;    /* bool */
    int* num_bins;     /* Spark's numBins(f) */
    int* num_splits;   /* Spark's numSplits(f) */
    int* hist_bins;    /* bins the aggregator holds per feature: continuous numBins, categorical arity */
    int64_t* hist_off; /* prefix of hist_bins*C, F+1 */
} meta_t;

static int build_meta(meta_t* m, int F, int C, int64_t n, int max_bins, const int32_t* arity, char* err, int errlen) {
    m->F = F; m->C = C;
    m->max_possible_bins = (int)(max_bins < n ? max_bins : n);
    m->arity = calloc(F, sizeof(int)); m->unordered = calloc(F, sizeof(int));
    m->num_bins = calloc(F, sizeof(int)); m->num_splits = calloc(F, sizeof(int));
    m->hist_bins = calloc(F, sizeof(int)); m->hist_off = calloc(F + 1, sizeof(int64_t));
    int max_unordered = (int)floor(log((double)(m->max_possible_bins / 2 + 1)) / log(2.0) + 1.0);
    for (int f = 0; f < F; f++) {
        int k = arity ? arity[f] : 0;
        if (k <= 1) k = 0;                       /* arity-1 categorical behaves as continuous */
        m->arity[f] = k;
        if (k > m->max_possible_bins) {
            snprintf(err, errlen, "DecisionTree requires maxBins (= %d) to be at least as large as the number of "
                     "values in each categorical feature, but categorical feature %d has %d values. Considering "
                     "remove this and other categorical features with a large number of values, or add more "
                     "training examples.", m->max_possible_bins, f, k);
            return -1;
        }
        if (k == 0) { m->num_bins[f] = m->max_possible_bins; m->num_splits[f] = m->num_bins[f] - 1; }
        else if (C > 2 && k <= max_unordered) { m->unordered[f] = 1; m->num_bins[f] = (1 << (k - 1)) - 1; m->num_splits[f] = m->num_bins[f]; }
        else { m->num_bins[f] = k; m->num_splits[f] = k - 1; }
    }
    return 0;
}
static void finish_meta(meta_t* m) {
    m->hist_off[0] = 0;
    for (int f = 0; f < m->F; f++) {
        m->hist_bins[f] = m->arity[f] ? m->arity[f] : m->num_bins[f];
        m->hist_off[f + 1] = m->hist_off[f] + (int64_t)m->hist_bins[f] * m->C;
    }
}
static void free_meta(meta_t* m) {
    free(m->arity); free(m->unordered); free(m->num_bins); free(m->num_splits); free(m->hist_bins); free(m->hist_off);
}

/* ------------------------------------------------------------------------------------------
 * RandomForest.findSplits: row sample (RDD.sample -> BernoulliSampler / GapSampling)  [A.5]
 * (core/.../rdd/PartitionwiseSampledRDD.scala, util/random/RandomSampler.scala)
 * Writes the sampled row indices (ascending); returns their count.
 * ---------------------------------------------------------------------------------------- */
ORC_API int64_t orc_split_sample_rows(int64_t n, int32_t max_possible_bins, int64_t seed,
                                      const int64_t* part_off, int32_t num_parts, int64_t* out_idx) {
    int64_t required = (int64_t)max_possible_bins * max_possible_bins;
    if (required < 10000) required = 10000;
    if (required >= n) { for (int64_t i = 0; i < n; i++) out_idx[i] = i; return n; }
    double fraction = (double)required / (double)n;
    xorshift_t sx; xs_seed(&sx, seed);
    int32_t sample_seed = xs_int(&sx);
    jrand_t jr; jr_seed(&jr, (int64_t)sample_seed);
    int64_t cnt = 0;
    for (int p = 0; p < num_parts; p++) {
        int64_t pseed = jr_long(&jr);
        xorshift_t rng; xs_seed(&rng, pseed);
        if (fraction <= 0.4) {
            double lnq = log1p(-fraction);
            int drop;
            { double u = xs_double(&rng); if (u < 5e-11) u = 5e-11; drop = (int)(log(u) / lnq); }
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++) {
                if (drop > 0) { drop--; continue; }
                double u = xs_double(&rng); if (u < 5e-11) u = 5e-11;
                drop = (int)(log(u) / lnq);
                out_idx[cnt++] = i;
            }
        } else {
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++)
                if (xs_double(&rng) <= fraction) out_idx[cnt++] = i;
        }
    }
    return cnt;
}

/* RandomForest.findSplitsForContinuousFeature  [A.5] — thresholds of one feature from the sampled
 * values.  Returns the number of thresholds written (<= num_splits). */
static int cmp_jd(const void* a, const void* b) { return jdcmp(*(const double*)a, *(const double*)b); }
static int find_splits_cont(double* vals /* scratch, sorted in place */, int64_t ns, int num_splits, double* out) {
    if (ns == 0) return 0;
    qsort(vals, (size_t)ns, sizeof(double), cmp_jd);
    /* distinct values with counts */
    int64_t m = 0;
    int64_t* cnt = malloc(sizeof(int64_t) * (size_t)ns);
    for (int64_t i = 0; i < ns; i++) {
        if (m > 0 && vals[m - 1] == vals[i]) cnt[m - 1]++;
        else { vals[m] = vals[i]; cnt[m] = 1; m++; }
    }
    int64_t possible = m - 1;
    int k = 0;
    if (possible == 0) { free(cnt); return 0; }
    if (possible <= num_splits) {
        for (int64_t i = 1; i <= possible; i++) out[k++] = (vals[i - 1] + vals[i]) / 2.0;
    } else {
        double stride = (double)ns / (double)(num_splits + 1);
        int64_t cur = cnt[0];
        double target = stride;
        for (int64_t i = 1; i < m; i++) {
            int64_t prev = cur;
            cur += cnt[i];
            double pg = fabs((double)prev - target), cg = fabs((double)cur - target);
            if (pg < cg) {
                if (k < num_splits) out[k] = (vals[i - 1] + vals[i]) / 2.0;
                k++;
                target += stride;
            }
        }
        if (k > num_splits) k = num_splits;
    }
    free(cnt);
    return k;
}

/* thresholds: [F][max_bins] (only continuous rows used); nsplits: [F] */
ORC_API int orc_find_splits(const double* X, int64_t n, int32_t F, int64_t ld, const int32_t* arity,
                            int32_t max_bins, int64_t seed, const int64_t* part_off, int32_t num_parts,
                            double* thresholds, int32_t* nsplits) {
    int mpb = (int)(max_bins < n ? max_bins : n);
    int64_t* idx = malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    int64_t ns = orc_split_sample_rows(n, mpb, seed, part_off, num_parts, idx);
#pragma omp parallel
    {
        double* vals = malloc(sizeof(double) * (size_t)(ns > 0 ? ns : 1));
#pragma omp for schedule(dynamic, 1)
        for (int f = 0; f < F; f++) {
            if (arity && arity[f] > 1) { nsplits[f] = 0; continue; }
            for (int64_t i = 0; i < ns; i++) vals[i] = X[idx[i] * ld + f];
            nsplits[f] = find_splits_cont(vals, ns, mpb - 1, thresholds + (int64_t)f * max_bins);
        }
        free(vals);
    }
    free(idx);
    return 0;
}

/* TreePoint.convertToTreeRDD / findBin  [A.5] -> bins u8 [n][ldb].  Returns 0, or -(f+1)-style error. */
ORC_API int64_t orc_bin_features(const double* X, int64_t n, int32_t F, int64_t ld, const int32_t* arity,
                                 const double* thresholds, const int32_t* nsplits, int32_t max_bins,
                                 uint8_t* bins, int64_t ldb) {
    int64_t bad = 0;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        for (int f = 0; f < F; f++) {
            double x = X[i * ld + f];
            int b;
            if (arity && arity[f] > 1) {
                b = (int)x;
                if (x < 0 || b >= arity[f]) {
#pragma omp critical
                    { if (!bad) bad = i * (int64_t)F + f + 1; }
                    b = 0;
                }
            } else b = java_find_bin(thresholds + (int64_t)f * max_bins, nsplits[f], x);
            bins[i * ldb + f] = (uint8_t)b;
        }
    }
    return bad;   /* 0 ok; else 1 + flat index of an invalid categorical value */
}

/* BaggedPoint.convertToBaggedRDD  [A.6] -> weights u8 [T][n] */
ORC_API int orc_bag_weights(int64_t n, int32_t T, double rate, int64_t seed, const int64_t* part_off,
                            int32_t num_parts, uint8_t* w) {
    if (T == 1 && rate == 1.0) { memset(w, 1, (size_t)n); return 0; }
    int with_replacement = T > 1;
#pragma omp parallel for schedule(dynamic, 1)
    for (int p = 0; p < num_parts; p++) {
        if (with_replacement) {
            well_t* g = malloc(sizeof(well_t));
            well_seed(g, seed + p + 1);
            double pe = exp(-rate);
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++)
                for (int t = 0; t < T; t++) {
                    int k = poisson_small(g, rate, pe);
                    w[(int64_t)t * n + i] = (uint8_t)(k > 255 ? 255 : k);
                }
            free(g);
        } else {
            xorshift_t r; xs_seed(&r, seed + p + 1);
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++)
                for (int t = 0; t < T; t++)
                    w[(int64_t)t * n + i] = xs_double(&r) < rate ? 1 : 0;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Impurity + gain  [A.8]  (mllib/tree/impurity/{Gini,Entropy}.scala,
 * ml/tree/impl/RandomForest.calculateImpurityStats)
 * ---------------------------------------------------------------------------------------- */
static double stat_sum(const double* s, int C) { double t = 0.0; for (int c = 0; c < C; c++) t += s[c]; return t; }
static double impurity_of(const double* s, int C, int kind) {
    double total = stat_sum(s, C);
    if (total == 0) return 0.0;
    if (kind == 0) {
        double imp = 1.0;
        for (int c = 0; c < C; c++) { double f = s[c] / total; imp -= f * f; }
        return imp;
    }
    double imp = 0.0;
    for (int c = 0; c < C; c++) if (s[c] != 0) { double f = s[c] / total; imp -= f * (log(f) / log(2.0)); }
    return imp;
}
static int first_argmax(const double* s, int C) {
    int best = -1; double mv = -INFINITY;   /* indexOfLargestArrayElement: strict >, first wins */
    for (int c = 0; c < C; c++) if (best < 0 || s[c] > mv) { best = c; mv = s[c]; }
    return best;
}

typedef struct {
    double gain, impurity, left_imp, right_imp;
    int valid;
} gain_t;
#define JDBL_MIN (-1.7976931348623157e308)   /* scala Double.MinValue */

static gain_t calc_gain(const double* left, const double* right, int C, double parent_imp,
                        int min_inst, double min_gain, int kind) {
    gain_t g; g.impurity = parent_imp; g.left_imp = g.right_imp = 0; g.valid = 0; g.gain = JDBL_MIN;
    int64_t lc = (int64_t)stat_sum(left, C), rc = (int64_t)stat_sum(right, C);
    int64_t tc = lc + rc;
    if (lc < min_inst || rc < min_inst) return g;
    double li = impurity_of(left, C, kind), ri = impurity_of(right, C, kind);
    double lw = (double)lc / (double)tc, rw = (double)rc / (double)tc;
    double gain = parent_imp - lw * li - rw * ri;
    if (gain < min_gain) return g;
    g.gain = gain; g.left_imp = li; g.right_imp = ri; g.valid = 1;
    return g;
}

/* result of binsToBestSplit for one node */
typedef struct {
    int feature;           /* -1: no valid feature (dummy split) */
    int split_index;       /* continuous: threshold index; unordered: split number; ordered: prefix length-1 */
    int kind;              /* 0 continuous, 1 categorical */
    double threshold;
    uint64_t left_mask;    /* categorical: left categories */
    gain_t g;
    double* left;          /* [C] */
    double* right;         /* [C] */
} best_t;

/* binsToBestSplit  [A.8].  hist: per-feature [hist_bins][C] doubles (continuous: raw bin counts;
 * categorical: per-category counts); parent: [C].  feats/nfeats: evaluation order. */
static void best_split_for_node(const meta_t* m, const double* hist, const double* parent,
                                const int* feats, int nfeats, const double* thresholds, int max_bins,
                                const orc_params* p, best_t* out, double* scratch /* >= (maxbins+2)*C + 64 */) {
    int C = m->C;
    double parent_imp = impurity_of(parent, C, p->impurity);
    int have = 0;
    double* cum = scratch;                 /* [hist_bins][C] */
    double* L = out->left; double* R = out->right;
    double* tl = scratch + (size_t)(256 + 1) * C;   /* candidate left/right */
    double* tr = tl + C;
    out->feature = -1; out->split_index = 0; out->kind = 0; out->threshold = 0.0; out->left_mask = 0;
    out->g.gain = JDBL_MIN; out->g.impurity = parent_imp; out->g.left_imp = out->g.right_imp = 0; out->g.valid = 0;
    for (int c = 0; c < C; c++) { L[c] = 0; R[c] = 0; }
    for (int fi = 0; fi < nfeats; fi++) {
        int f = feats ? feats[fi] : fi;
        int ns = m->num_splits[f];
        if (ns == 0) continue;
        const double* h = hist + m->hist_off[f];
        int nb = m->hist_bins[f];
        int order[256];
        if (m->arity[f] == 0) {
            /* continuous: prefix sums over bins; left = cum[j], right = cum[ns] - left */
            for (int c = 0; c < C; c++) cum[c] = h[c];
            for (int b = 1; b < nb; b++) for (int c = 0; c < C; c++) cum[b * C + c] = cum[(b - 1) * C + c] + h[b * C + c];
        } else if (!m->unordered[f]) {
            /* ordered categorical: sort categories by centroid (stable), then prefix sums */
            double cen[256];
            for (int v = 0; v < nb; v++) {
                const double* s = h + v * C;
                int64_t cnt = (int64_t)stat_sum(s, C);
                if (cnt != 0) cen[v] = (C > 2) ? impurity_of(s, C, p->impurity) : s[1];
                else cen[v] = 1.7976931348623157e308;
                order[v] = v;
            }
            for (int i = 1; i < nb; i++) {           /* stable insertion sort == List.sortBy */
                int o = order[i]; int j = i - 1;
                while (j >= 0 && jdcmp(cen[order[j]], cen[o]) > 0) { order[j + 1] = order[j]; j--; }
                order[j + 1] = o;
            }
            for (int b = 0; b < nb; b++) for (int c = 0; c < C; c++)
                cum[b * C + c] = (b ? cum[(b - 1) * C + c] : 0.0) + h[order[b] * C + c];
        }
        for (int s = 0; s < ns; s++) {
            if (m->arity[f] && m->unordered[f]) {
                /* left = sum of the categories whose bit is set in s+1; right = parent - left */
                for (int c = 0; c < C; c++) tl[c] = 0.0;
                for (int v = 0; v < nb; v++) if (((s + 1) >> v) & 1) for (int c = 0; c < C; c++) tl[c] += h[v * C + c];
                for (int c = 0; c < C; c++) tr[c] = parent[c] - tl[c];
            } else {
                int last = (m->arity[f] == 0) ? ns : nb - 1;
                for (int c = 0; c < C; c++) { tl[c] = cum[s * C + c]; tr[c] = cum[last * C + c] - tl[c]; }
            }
            gain_t g = calc_gain(tl, tr, C, parent_imp, p->min_instances_per_node, p->min_info_gain, p->impurity);
            if (!have || g.gain > out->g.gain) {       /* maxBy: first maximum wins */
                have = 1;
                out->g = g; out->feature = f; out->split_index = s;
                memcpy(L, tl, sizeof(double) * C); memcpy(R, tr, sizeof(double) * C);
                if (m->arity[f] == 0) { out->kind = 0; out->threshold = thresholds[(int64_t)f * max_bins + s]; out->left_mask = 0; }
                else {
                    out->kind = 1; out->threshold = 0.0;
                    uint64_t mask = 0;
                    if (m->unordered[f]) mask = (uint64_t)(s + 1);
                    else for (int b = 0; b <= s; b++) mask |= 1ULL << order[b];
                    out->left_mask = mask;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Trees
 * ---------------------------------------------------------------------------------------- */
typedef struct lnode {
    int32_t id;                 /* Spark node index: root 1, children 2i / 2i+1 */
    int is_leaf;
    int has_stats;              /* node went through binsToBestSplit */
    int valid;                  /* stats.valid */
    int feature, kind, split_index;
    double threshold; uint64_t left_mask; int arity;
    double gain, impurity;
    double* stats;              /* [C] impurityCalculator.stats */
    struct lnode *left, *right;
    /* training state */
    int32_t* rows; int64_t nrows;   /* rows with weight > 0 currently at this node */
    int* feat_subset; int nsubset;
} lnode;

static lnode* new_node(int id, int C) {
    lnode* n = calloc(1, sizeof(lnode));
    n->id = id; n->stats = calloc(C, sizeof(double)); n->feature = -1; n->valid = 1;
    return n;
}
static void free_tree(lnode* n) {
    if (!n) return;
    free_tree(n->left); free_tree(n->right);
    free(n->stats); free(n->rows); free(n->feat_subset); free(n);
}
static int level_of(int id) { int l = 0; while (id > 1) { id >>= 1; l++; } return l; }

/* flat, pruned forest (Node.scala after LearningNode.toNode(prune = true))  [A.9] */
typedef struct {
    int T, C, F;
    int64_t total;
    int64_t* tree_off;          /* T+1 */
    int32_t *feature, *left, *right, *node_id, *arity;
    double *threshold, *prediction, *impurity, *gain, *stats;
    uint64_t* left_mask;
    /* stage outputs kept for stage-level parity tests */
    double* thresholds; int32_t* nsplits; int max_bins;
    int64_t num_groups;
} forest_t;

typedef struct { int64_t cap; forest_t* f; } fbuild_t;
static void fgrow(fbuild_t* b, int C) {
    forest_t* f = b->f;
    if (f->total < b->cap) return;
    b->cap = b->cap ? b->cap * 2 : 1024;
    f->feature = realloc(f->feature, 4 * b->cap); f->left = realloc(f->left, 4 * b->cap);
    f->right = realloc(f->right, 4 * b->cap); f->node_id = realloc(f->node_id, 4 * b->cap);
    f->arity = realloc(f->arity, 4 * b->cap);
    f->threshold = realloc(f->threshold, 8 * b->cap); f->prediction = realloc(f->prediction, 8 * b->cap);
    f->impurity = realloc(f->impurity, 8 * b->cap); f->gain = realloc(f->gain, 8 * b->cap);
    f->stats = realloc(f->stats, 8 * b->cap * C); f->left_mask = realloc(f->left_mask, 8 * b->cap);
}
/* returns index (relative to tree start) of the emitted node; *is_leaf_pred gets the prediction if leaf */
static int64_t emit_node(fbuild_t* b, const lnode* n, int64_t base, int C, int* leaf_out) {
    forest_t* f = b->f;
    fgrow(b, C);
    int64_t me = f->total++;
    memcpy(f->stats + me * C, n->stats, sizeof(double) * C);
    f->node_id[me] = n->id;
    f->prediction[me] = (double)first_argmax(n->stats, C);
    if (n->left || n->right) {
        int ll = 0, rl = 0;
        int64_t li = emit_node(b, n->left, base, C, &ll);
        int64_t ri = emit_node(b, n->right, base, C, &rl);
        if (ll && rl && f->prediction[base + li] == f->prediction[base + ri]) {
            /* prune: both children are leaves predicting the same class -> this node becomes a leaf
             * carrying its own stats */
            f->total = me + 1;
            f->feature[me] = -1; f->left[me] = f->right[me] = -1; f->arity[me] = 0;
            f->threshold[me] = 0; f->left_mask[me] = 0; f->gain[me] = 0;
            f->impurity[me] = n->impurity;
            *leaf_out = 1;
            return me - base;
        }
        f->feature[me] = n->feature; f->left[me] = (int32_t)li; f->right[me] = (int32_t)ri;
        f->arity[me] = n->kind ? n->arity : 0;
        f->threshold[me] = n->threshold; f->left_mask[me] = n->left_mask;
        f->gain[me] = n->gain; f->impurity[me] = n->impurity;
        *leaf_out = 0;
        return me - base;
    }
    f->feature[me] = -1; f->left[me] = f->right[me] = -1; f->arity[me] = 0;
    f->threshold[me] = 0; f->left_mask[me] = 0; f->gain[me] = 0;
    f->impurity[me] = n->valid ? n->impurity : -1.0;
    *leaf_out = 1;
    return me - base;
}

/* Scala 2.11 mutable.HashMap[Int, _] iteration order (HashTable.foreachEntry), used by
 * findBestSplits' `nodesForGroup.foreach`  [A.7, unpinned].  keys: insertion order.  Writes the
 * iteration order into out. */
static uint32_t bswap32(uint32_t x) { return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24); }
static int scala_index(uint32_t hcode, int table_len) {
    uint32_t i = hcode * 0x9e3775cdu; i = bswap32(i); i *= 0x9e3775cdu;
    int rotation = 4 % 32;                           /* seedvalue = bitCount(16 - 1), fixed at construction */
    uint32_t rotated = (i >> rotation) | (i << (32 - rotation));
    int ones = table_len - 1;
    int exponent = __builtin_clz((unsigned)ones);
    return (int)((rotated >> exponent) & (uint32_t)ones);
}
static void scala_hashmap_order(const int* keys, int nk, int* out) {
    int tlen = 16, thr = 12, size = 0;
    int* head = malloc(sizeof(int) * 16);
    int* next = malloc(sizeof(int) * (nk + 1));
    for (int i = 0; i < tlen; i++) head[i] = -1;
    for (int k = 0; k < nk; k++) {
        int h = scala_index((uint32_t)keys[k], tlen);
        next[k] = head[h]; head[h] = k; size++;
        if (size > thr) {
            int nlen = tlen * 2;
            int* nh = malloc(sizeof(int) * nlen);
            for (int i = 0; i < nlen; i++) nh[i] = -1;
            for (int i = tlen - 1; i >= 0; i--) {
                int e = head[i];
                while (e >= 0) {
                    int e1 = next[e];
                    int hh = scala_index((uint32_t)keys[e], nlen);
                    next[e] = nh[hh]; nh[hh] = e;
                    e = e1;
                }
            }
            free(head); head = nh; tlen = nlen;
            thr = (int)((int64_t)nlen * 750 / 1000);
        }
    }
    int o = 0;
    for (int i = tlen - 1; i >= 0; i--) for (int e = head[i]; e >= 0; e = next[e]) out[o++] = e;
    free(head); free(next);
}

/* SamplingUtils.reservoirSampleAndCount over 0 until F  [A.7] */
static void reservoir_features(int F, int k, int64_t seed, int* out) {
    int i = 0;
    for (; i < k && i < F; i++) out[i] = i;
    if (i < k) return;
    int64_t l = i;
    xorshift_t r; xs_seed(&r, seed);
    for (int item = i; item < F; item++) {
        l++;
        int64_t j = (int64_t)(xs_double(&r) * (double)l);
        if (j < k) out[j] = item;
    }
}

typedef struct { int tree; lnode* node; } sitem;

/* RandomForest.run  [A.5-A.9] */
ORC_API void* orc_fit_forest(const double* X, int64_t n, int32_t F, int64_t ld, const double* y,
                             const int32_t* arity, const int64_t* part_off, int32_t num_parts,
                             const orc_params* p, const uint8_t* weights_in /* optional [T][n] */,
                             char* err, int32_t errlen) {
    int C = p->num_classes, T = p->num_trees;
    meta_t m;
    if (err && errlen) err[0] = 0;
    if (p->max_depth < 0 || p->max_depth > 30) { snprintf(err, errlen, "DecisionTree currently only supports maxDepth <= 30, but was given maxDepth = %d.", p->max_depth); return NULL; }
    if (p->max_bins > 256) { snprintf(err, errlen, "oracle supports maxBins <= 256"); return NULL; }
    for (int64_t i = 0; i < n; i++) {
        double l = y[i];
        if (!(l >= 0) || l >= C || l != floor(l)) {
            snprintf(err, errlen, "Classifier given dataset with invalid label %g.  Labels must be integers in range [0, %d).", l, C);
            return NULL;
        }
    }
    if (build_meta(&m, F, C, n, p->max_bins, arity, err, errlen)) { free_meta(&m); return NULL; }
    int MB = p->max_bins;
    forest_t* fo = calloc(1, sizeof(forest_t));
    fo->T = T; fo->C = C; fo->F = F; fo->max_bins = MB;
    fo->thresholds = calloc((size_t)F * MB, sizeof(double));
    fo->nsplits = calloc(F, sizeof(int32_t));
    orc_find_splits(X, n, F, ld, arity, p->max_bins, p->seed, part_off, num_parts, fo->thresholds, fo->nsplits);
    for (int f = 0; f < F; f++) if (m.arity[f] == 0) { m.num_splits[f] = fo->nsplits[f]; m.num_bins[f] = fo->nsplits[f] + 1; }
    finish_meta(&m);
    int64_t ldb = F;
    uint8_t* bins = malloc((size_t)(n * ldb + 1));
    int64_t bad = orc_bin_features(X, n, F, ld, arity, fo->thresholds, fo->nsplits, MB, bins, ldb);
    if (bad) {
        int f = (int)((bad - 1) % F); int64_t i = (bad - 1) / F;
        snprintf(err, errlen, "DecisionTree given invalid data: Feature %d is categorical with values in {0,...,%d}, "
                 "but a data point gives it value %g.", f, m.arity[f] - 1, X[i * ld + f]);
        free(bins); free_meta(&m); free(fo->thresholds); free(fo->nsplits); free(fo); return NULL;
    }
    uint8_t* w = malloc((size_t)((int64_t)T * n + 1));
    if (weights_in) memcpy(w, weights_in, (size_t)((int64_t)T * n));
    else orc_bag_weights(n, T, p->subsampling_rate, p->seed, part_off, num_parts, w);
    uint8_t* lab = malloc((size_t)n + 1);
    for (int64_t i = 0; i < n; i++) lab[i] = (uint8_t)y[i];

    int subsample = p->features_per_node < F;
    int64_t total_bins_all = 0;
    for (int f = 0; f < F; f++) total_bins_all += m.num_bins[f];

    lnode** roots = malloc(sizeof(lnode*) * T);
    int64_t scap = 1024, sp = 0;
    sitem* stack = malloc(sizeof(sitem) * scap);
    for (int t = 0; t < T; t++) {
        lnode* r = new_node(1, C);
        int64_t k = 0;
        r->rows = malloc(sizeof(int32_t) * (size_t)(n + 1));
        for (int64_t i = 0; i < n; i++) if (w[(int64_t)t * n + i]) r->rows[k++] = (int32_t)i;
        r->nrows = k;
        roots[t] = r;
        stack[sp].tree = t; stack[sp].node = r; sp++;
        if (sp == scap) { scap *= 2; stack = realloc(stack, sizeof(sitem) * scap); }
    }
    jrand_t rng; jr_seed(&rng, p->seed);
    int64_t hist_len = m.hist_off[F];
    int maxthreads = 1;
#ifdef _OPENMP
    maxthreads = omp_get_max_threads();
#endif
    double* hist_pool = malloc(sizeof(double) * (size_t)(hist_len + 1) * maxthreads);
    double* scratch_pool = malloc(sizeof(double) * (size_t)((256 + 4) * C + 64) * maxthreads);

    sitem* group = NULL; int64_t gcap = 0;
    best_t* bests = NULL;
    while (sp > 0) {
        /* selectNodesToSplit */
        int64_t ng = 0, mem = 0;
        while (sp > 0) {
            sitem top = stack[sp - 1];
            int* subset = NULL;
            int64_t tb = total_bins_all;
            if (subsample) {
                subset = malloc(sizeof(int) * p->features_per_node);
                reservoir_features(F, p->features_per_node, jr_long(&rng), subset);
                tb = 0;
                for (int j = 0; j < p->features_per_node; j++) tb += m.num_bins[subset[j]];
            }
            int64_t cost = (int64_t)C * tb * 8;
            if (mem + cost <= p->max_memory_bytes || mem == 0) {
                sp--;
                if (ng == gcap) { gcap = gcap ? gcap * 2 : 256; group = realloc(group, sizeof(sitem) * gcap); }
                free(top.node->feat_subset);
                top.node->feat_subset = subset; top.node->nsubset = subset ? p->features_per_node : 0;
                group[ng++] = top;
                mem += cost;
            } else { free(subset); break; }
        }
        fo->num_groups++;
        bests = realloc(bests, sizeof(best_t) * ng);
        for (int64_t g = 0; g < ng; g++) { bests[g].left = calloc(C, sizeof(double)); bests[g].right = calloc(C, sizeof(double)); }
        /* findBestSplits: aggregate (binSeqOp) + binsToBestSplit per node */
#pragma omp parallel for schedule(dynamic, 1)
        for (int64_t g = 0; g < ng; g++) {
            int tid = 0;
#ifdef _OPENMP
            tid = omp_get_thread_num();
#endif
            double* hist = hist_pool + (size_t)(hist_len + 1) * tid;
            double* scratch = scratch_pool + (size_t)((256 + 4) * C + 64) * tid;
            lnode* nd = group[g].node; int t = group[g].tree;
            const uint8_t* wt = w + (int64_t)t * n;
            double parent[64];
            memset(hist, 0, sizeof(double) * (size_t)hist_len);
            for (int c = 0; c < C; c++) parent[c] = 0;
            int nf = nd->feat_subset ? nd->nsubset : F;
            for (int64_t r = 0; r < nd->nrows; r++) {
                int64_t i = nd->rows[r];
                double wi = (double)wt[i]; int l = lab[i];
                const uint8_t* b = bins + i * ldb;
                for (int fi = 0; fi < nf; fi++) {
                    int f = nd->feat_subset ? nd->feat_subset[fi] : fi;
                    hist[m.hist_off[f] + (int64_t)b[f] * C + l] += wi;
                }
                parent[l] += wi;
            }
            best_split_for_node(&m, hist, parent, nd->feat_subset, nf, fo->thresholds, MB, p, &bests[g], scratch);
            memcpy(nd->stats, parent, sizeof(double) * C);
        }
        /* driver side: Scala mutable.HashMap[treeIndex -> nodes] iteration order, nodes in pop order */
        int* tkeys = malloc(sizeof(int) * (ng + 1)); int nk = 0;
        for (int64_t g = 0; g < ng; g++) {
            int t = group[g].tree, seen = 0;
            for (int k = 0; k < nk; k++) if (tkeys[k] == t) { seen = 1; break; }
            if (!seen) tkeys[nk++] = t;
            if (nk > 4096 && 0) break;
        }
        int* torder = malloc(sizeof(int) * (nk + 1));
        scala_hashmap_order(tkeys, nk, torder);
        for (int oi = 0; oi < nk; oi++) {
            int t = tkeys[torder[oi]];
            for (int64_t g = 0; g < ng; g++) {
                if (group[g].tree != t) continue;
                lnode* nd = group[g].node; best_t* b = &bests[g];
                int level = level_of(nd->id);
                int is_leaf = (b->g.gain <= 0) || (level == p->max_depth);
                nd->is_leaf = is_leaf; nd->has_stats = 1; nd->valid = b->g.valid;
                nd->gain = b->g.gain; nd->impurity = b->g.impurity;
                if (!is_leaf) {
                    nd->feature = b->feature; nd->kind = b->kind; nd->split_index = b->split_index;
                    nd->threshold = b->threshold; nd->left_mask = b->left_mask; nd->arity = m.arity[b->feature];
                    int child_leaf = (level + 1) == p->max_depth;
                    int ll = child_leaf || (b->g.left_imp == 0.0), rl = child_leaf || (b->g.right_imp == 0.0);
                    lnode* L = new_node(nd->id * 2, C); lnode* R = new_node(nd->id * 2 + 1, C);
                    memcpy(L->stats, b->left, sizeof(double) * C); memcpy(R->stats, b->right, sizeof(double) * C);
                    L->is_leaf = ll; R->is_leaf = rl; L->impurity = b->g.left_imp; R->impurity = b->g.right_imp;
                    nd->left = L; nd->right = R;
                    /* route rows (LearningNode.predictImpl on binned features) */
                    if (!ll || !rl) {
                        L->rows = malloc(sizeof(int32_t) * (size_t)(nd->nrows + 1));
                        R->rows = malloc(sizeof(int32_t) * (size_t)(nd->nrows + 1));
                        int f = b->feature;
                        for (int64_t r = 0; r < nd->nrows; r++) {
                            int32_t i = nd->rows[r];
                            int bv = bins[(int64_t)i * ldb + f];
                            int go_left = b->kind == 0 ? (bv <= b->split_index) : (int)((b->left_mask >> bv) & 1);
                            if (go_left) L->rows[L->nrows++] = i; else R->rows[R->nrows++] = i;
                        }
                    }
                    if (!ll) { if (sp == scap) { scap *= 2; stack = realloc(stack, sizeof(sitem) * scap); } stack[sp].tree = t; stack[sp].node = L; sp++; }
                    if (!rl) { if (sp == scap) { scap *= 2; stack = realloc(stack, sizeof(sitem) * scap); } stack[sp].tree = t; stack[sp].node = R; sp++; }
                }
                free(nd->rows); nd->rows = NULL; nd->nrows = 0;
            }
        }
        free(tkeys); free(torder);
        for (int64_t g = 0; g < ng; g++) { free(bests[g].left); free(bests[g].right); }
    }
    /* toNode(prune = true) into the flat layout (pre-order, left child = next index) */
    fbuild_t fb = {0, fo};
    fo->tree_off = calloc(T + 1, sizeof(int64_t));
    for (int t = 0; t < T; t++) {
        int leaf = 0;
        fo->tree_off[t] = fo->total;
        emit_node(&fb, roots[t], fo->total, C, &leaf);
        free_tree(roots[t]);
    }
    fo->tree_off[T] = fo->total;
    free(roots); free(stack); free(group); free(bests); free(hist_pool); free(scratch_pool);
    free(bins); free(w); free(lab); free_meta(&m);
    return fo;
}

ORC_API int64_t orc_forest_total_nodes(void* h) { return ((forest_t*)h)->total; }
ORC_API int64_t orc_forest_num_groups(void* h) { return ((forest_t*)h)->num_groups; }
ORC_API void orc_forest_export(void* h, int64_t* tree_off, int32_t* feature, int32_t* left, int32_t* right,
                               int32_t* node_id, int32_t* arity, double* threshold, uint64_t* left_mask,
                               double* prediction, double* impurity, double* gain, double* stats,
                               double* thresholds, int32_t* nsplits) {
    forest_t* f = h; int64_t N = f->total;
    memcpy(tree_off, f->tree_off, 8 * (f->T + 1));
    memcpy(feature, f->feature, 4 * N); memcpy(left, f->left, 4 * N); memcpy(right, f->right, 4 * N);
    memcpy(node_id, f->node_id, 4 * N); memcpy(arity, f->arity, 4 * N);
    memcpy(threshold, f->threshold, 8 * N); memcpy(left_mask, f->left_mask, 8 * N);
    memcpy(prediction, f->prediction, 8 * N); memcpy(impurity, f->impurity, 8 * N);
    memcpy(gain, f->gain, 8 * N); memcpy(stats, f->stats, 8 * N * f->C);
    if (thresholds) memcpy(thresholds, f->thresholds, 8 * (size_t)f->F * f->max_bins);
    if (nsplits) memcpy(nsplits, f->nsplits, 4 * (size_t)f->F);
}
ORC_API void orc_forest_free(void* h) {
    forest_t* f = h; if (!f) return;
    free(f->tree_off); free(f->feature); free(f->left); free(f->right); free(f->node_id); free(f->arity);
    free(f->threshold); free(f->prediction); free(f->impurity); free(f->gain); free(f->stats); free(f->left_mask);
    free(f->thresholds); free(f->nsplits); free(f);
}

/* One node's aggregate (DTStatsAggregator layout restated as [feature][bin|category][class]) from an
 * explicit row list — the unit the GPU histogram kernel is checked against.  hist: int64 [hist_len] */
ORC_API void orc_node_histogram(const uint8_t* bins, int64_t ldb, const uint8_t* labels, const uint8_t* w,
                                const int32_t* rows, int64_t nrows, int32_t F, int32_t C,
                                const int64_t* hist_off, int64_t* hist, int64_t* parent) {
    for (int64_t r = 0; r < nrows; r++) {
        int64_t i = rows ? rows[r] : r;
        int64_t wi = w ? w[i] : 1; int l = labels[i];
        for (int f = 0; f < F; f++) hist[hist_off[f] + (int64_t)bins[i * ldb + f] * C + l] += wi;
        parent[l] += wi;
    }
}

/* ------------------------------------------------------------------------------------------
 * Prediction  [A.10]  (ml/classification/{DecisionTree,RandomForest}Classifier.scala,
 * ProbabilisticClassifier.scala, ml/tree/Node.scala predictImpl, Split.scala shouldGoLeft)
 * ---------------------------------------------------------------------------------------- */
static int cat_goes_left(double x, uint64_t left_mask, int arity) {
    /* CategoricalSplit keeps the smaller side: isLeft = |left| <= arity/2 */
    int nl = __builtin_popcountll(left_mask);
    int is_left = nl <= arity / 2;
    int integral = (x == floor(x)) && x >= 0 && x < 64;
    if (is_left) return integral && ((left_mask >> (int)x) & 1);
    uint64_t all = arity >= 64 ? ~0ULL : ((1ULL << arity) - 1);
    uint64_t right = all & ~left_mask;
    return !(integral && ((right >> (int)x) & 1));
}

ORC_API void orc_predict(int32_t T, int32_t C, const int64_t* tree_off, const int32_t* feature,
                         const int32_t* left, const int32_t* right, const int32_t* arity,
                         const double* threshold, const uint64_t* left_mask, const double* stats,
                         const double* X, int64_t n, int64_t ld, int32_t is_forest,
                         double* raw, double* prob, double* pred) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        const double* x = X + i * ld;
        double* r = raw + i * C;
        for (int c = 0; c < C; c++) r[c] = 0.0;
        for (int t = 0; t < T; t++) {
            int64_t base = tree_off[t], k = base;
            while (feature[k] >= 0) {
                double v = x[feature[k]];
                int go_left = arity[k] ? cat_goes_left(v, left_mask[k], arity[k]) : (v <= threshold[k]);
                k = base + (go_left ? left[k] : right[k]);
            }
            const double* s = stats + k * C;
            if (!is_forest) { for (int c = 0; c < C; c++) r[c] = s[c]; }
            else {
                double total = stat_sum(s, C);
                if (total != 0) for (int c = 0; c < C; c++) r[c] += s[c] / total;
            }
        }
        double sum = stat_sum(r, C);
        for (int c = 0; c < C; c++) prob[i * C + c] = (sum != 0) ? r[c] / sum : r[c];
        pred[i] = (double)first_argmax(r, C);
    }
}

/* ------------------------------------------------------------------------------------------
 * WISDM window transform — SURVEY.md Appendix B (parity unpinned; this file is the arbiter)
 * ---------------------------------------------------------------------------------------- */

/* B.1 segmentation: a bout is a maximal run of equal (user, activity) (optionally also broken where
 * ts[i]-ts[i-1] > gap_ns); windows are consecutive non-overlapping runs of L samples inside a bout,
 * remainder dropped.  win_start may be NULL to count only. */
ORC_API int64_t orc_window_segments(const int32_t* user, const int32_t* act, const int64_t* ts, int64_t n,
                                    int32_t L, int64_t gap_ns, int64_t* win_start) {
    int64_t nw = 0, bstart = 0;
    for (int64_t i = 1; i <= n; i++) {
        int boundary = (i == n) || user[i] != user[i - 1] || act[i] != act[i - 1] ||
                       (gap_ns > 0 && ts && ts[i] - ts[i - 1] > gap_ns);
        if (boundary) {
            int64_t k = (i - bstart) / L;
            for (int64_t j = 0; j < k; j++) { if (win_start) win_start[nw] = bstart + j * L; nw++; }
            bstart = i;
        }
    }
    return nw;
}

/* time between peaks of one axis (ms); NaN when fewer than 3 local maxima */
static double peak_interval_ms(const double* a, const int64_t* ts, int L, double lo, double hi) {
    int npk = 0;
    double p1 = -INFINITY, p2 = -INFINITY, p3 = -INFINITY;     /* three highest peak heights */
    for (int i = 1; i + 1 < L; i++) {
        if (a[i] > a[i - 1] && a[i] >= a[i + 1]) {
            double h = a[i]; npk++;
            if (h > p1) { p3 = p2; p2 = p1; p1 = h; }
            else if (h > p2) { p3 = p2; p2 = h; }
            else if (h > p3) { p3 = h; }
        }
    }
    if (npk < 3) return NAN;
    /* threshold ladder tau_k = P - ((k+1)*0.1)*(P - lo), k = 0..9; first k admitting >= 3 peaks */
    double P = p1, R = P - lo, tau = lo;
    for (int k = 0; k < 10; k++) {
        double t = P - ((double)(k + 1) * 0.1) * R;
        if (k == 9) t = lo;
        if (t <= p3) { tau = t; break; }
    }
    (void)hi;
    int first = -1, last = -1, cnt = 0;
    for (int i = 1; i + 1 < L; i++)
        if (a[i] > a[i - 1] && a[i] >= a[i + 1] && a[i] >= tau) { if (first < 0) first = i; last = i; cnt++; }
    return ((double)(ts[last] - ts[first]) / (double)(cnt - 1)) / 1.0e6;
}

/* B.2 per-window features, column order of wisdm_data.csv:1 minus UID/USER/ACTIVITY:
 * X0..X9 Y0..Y9 Z0..Z9 XAVG YAVG ZAVG XPEAK YPEAK ZPEAK XABSDEV YABSDEV ZABSDEV XSTDDEV YSTDDEV ZSTDDEV
 * RESULTANT  (43 doubles per window, row-major) */
ORC_API void orc_window_features(const double* x, const double* y, const double* z, const int64_t* ts,
                                 const int64_t* win_start, int64_t nw, int32_t L, double* out) {
#pragma omp parallel for schedule(static)
    for (int64_t w = 0; w < nw; w++) {
        int64_t s = win_start[w];
        const double* ax[3] = {x + s, y + s, z + s};
        double* o = out + w * 43;
        for (int a = 0; a < 3; a++) {
            const double* v = ax[a];
            double sum = 0, lo = v[0], hi = v[0];
            for (int i = 0; i < L; i++) { sum += v[i]; if (v[i] < lo) lo = v[i]; if (v[i] > hi) hi = v[i]; }
            double mean = sum / (double)L;
            double sq = 0, ab = 0;
            int cnt[10] = {0};
            double scale = hi > lo ? 10.0 / (hi - lo) : 0.0;
            for (int i = 0; i < L; i++) {
                double d = v[i] - mean;
                sq += d * d; ab += fabs(d);
                int b = 0;
                { double q = floor((v[i] - lo) * scale); b = q >= 9.0 ? 9 : (int)q; }
                cnt[b]++;
            }
            for (int b = 0; b < 10; b++) o[a * 10 + b] = (double)cnt[b] / (double)L;
            o[30 + a] = mean;
            o[33 + a] = peak_interval_ms(v, ts + s, L, lo, hi);
            o[36 + a] = ab / (double)L;
            o[39 + a] = sqrt(sq / (double)L);
        }
        double rs = 0;
        for (int i = 0; i < L; i++) rs += sqrt(ax[0][i] * ax[0][i] + ax[1][i] * ax[1][i] + ax[2][i] * ax[2][i]);
        o[42] = rs / (double)L;
    }
}

/* ------------------------------------------------------------------------------------------
 * Synthetic tri-axial stream: counter-based, so CPU and GPU inputs are identical.
 * ---------------------------------------------------------------------------------------- */
static uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
/* bouts: user u owns samples [u*per_user, (u+1)*per_user); inside, consecutive bouts of length
 * 600 + h % 5400 with activity h' % 6 (never equal to the previous one) */
ORC_API void orc_synth_stream(int64_t n, int64_t first_sample, int64_t per_user, uint64_t seed,
                              const double* sine1024, const double* mu /*[6][3]*/, const double* amp /*[6][3]*/,
                              const double* sigma /*[6][3]*/, const int32_t* freq /*[6]*/,
                              int32_t* user, int32_t* act, int64_t* ts, double* x, double* y, double* z) {
    /* sequential over users; bout layout depends only on (seed, user) */
    int64_t i = 0;
    while (i < n) {
        int64_t g = first_sample + i;
        int64_t u = g / per_user, ubeg = u * per_user, uend = ubeg + per_user;
        /* walk the user's bouts up to g */
        int64_t b0 = ubeg; int64_t bi = 0; int prev_act = -1;
        for (;;) {
            uint64_t h = splitmix64(seed ^ ((uint64_t)u * 0xD1B54A32D192ED03ULL) ^ ((uint64_t)bi * 0x9E3779B97F4A7C15ULL));
            int64_t len = 600 + (int64_t)(h % 5400);
            int a = (int)((h >> 32) % 6);
            if (a == prev_act) a = (a + 1) % 6;
            const double gain = 0.35 + 1.3 * ((double)((h >> 8) & 0xffff) / 65535.0);
            const double tilt = (double)((h >> 24) & 0xffff) / 65535.0 - 0.5;
            static const double tiltv[3] = {3.0, -4.0, 3.5};
            int64_t b1 = b0 + len; if (b1 > uend) b1 = uend;
            if (g < b1) {
                int64_t stop = b1 - first_sample; if (stop > n) stop = n;
                for (; i < stop; i++) {
                    int64_t gi = first_sample + i, k = gi - b0;
                    user[i] = (int32_t)u; act[i] = a;
                    ts[i] = (ubeg * 50000000LL) + (int64_t)bi * 1000000000LL + gi * 50000000LL;
                    double* outp[3] = {x, y, z};
                    for (int ax = 0; ax < 3; ax++) {
                        uint64_t r = splitmix64(seed ^ ((uint64_t)gi * 0x9E3779B97F4A7C15ULL) ^ (uint64_t)(ax + 1));
                        int64_t nz = (int64_t)(r & 0xffff) + (int64_t)((r >> 16) & 0xffff) +
                                     (int64_t)((r >> 32) & 0xffff) + (int64_t)((r >> 48) & 0xffff) - 2 * 65535;
                        double noise = (double)nz * (sigma[a * 3 + ax] * (1.0 / 37837.0));
                        /* per-bout gain / tilt (from the bout hash) make the six classes overlap */
                        double car = (amp[a * 3 + ax] * gain) * sine1024[(uint64_t)(k * freq[a] + ax * 85) & 1023];
                        outp[ax][i] = ((mu[a * 3 + ax] + tilt * tiltv[ax]) + car) + noise;
                    }
                }
                break;
            }
            prev_act = a; b0 = b1; bi++;
        }
    }
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its children; the timed CPU baseline sets its thread count explicitly */
ORC_API void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- raw WISDM text ingest (no reference code: the rules are those of
 * include/har_b200.h har_parse_wisdm_raw, restated with strtod / strtoll).  Returns the number of valid records
 * (written up to `capacity`), *dropped = malformed records. ------------------------------------------------------ */
static int orc_blank(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

static int orc_field_number(const char* p, const char* q, int want_int, long long* iv, double* dv) {
    char buf[64];
    size_t L = (size_t)(q - p);
    if (L == 0 || L >= sizeof buf) return 1;
    for (size_t i = 0; i < L; i++) {
        char c = p[i];
        int ok = (c >= '0' && c <= '9') || c == '-' || c == '+' || (!want_int && (c == '.' || c == 'e' || c == 'E'));
        if (!ok) return 1;
    }
    memcpy(buf, p, L); buf[L] = 0;
    char* end = NULL;
    if (want_int) {
        size_t digits = L - ((buf[0] == '-' || buf[0] == '+') ? 1 : 0);
        if (digits == 0 || digits > 18) return 1;
        *iv = strtoll(buf, &end, 10);
    } else *dv = strtod(buf, &end);
    return (end == buf + L) ? 0 : 1;
}

ORC_API int64_t orc_parse_wisdm_raw(const char* text, int64_t nbytes, const char* const* names, int32_t num_names, int64_t capacity,
                            int32_t* user, int32_t* act, int64_t* ts, double* x, double* y, double* z, int64_t* dropped) {
    int64_t nvalid = 0, ndrop = 0, a = 0;
    for (int64_t i = 0; i <= nbytes; i++) {
        if (i < nbytes && text[i] != ';') continue;
        int64_t s = a, e = i;
        a = i + 1;
        while (s < e && orc_blank((unsigned char)text[s])) s++;
        while (e > s && orc_blank((unsigned char)text[e - 1])) e--;
        if (s >= e) continue;                                  /* empty record (or blank tail) */
        const char* f0[6]; const char* f1[6];
        int nf = 0, bad = 0;
        const char* st = text + s;
        for (const char* p = text + s;; p++) {
            if (p == text + e || *p == ',') {
                if (nf == 6) { bad = 1; break; }
                f0[nf] = st; f1[nf] = p; nf++;
                if (p == text + e) break;
                st = p + 1;
            }
        }
        if (bad || nf != 6) { ndrop++; continue; }
        for (int k = 0; k < 6 && !bad; k++) {
            while (f0[k] < f1[k] && orc_blank((unsigned char)*f0[k])) f0[k]++;
            while (f1[k] > f0[k] && orc_blank((unsigned char)*(f1[k] - 1))) f1[k]--;
            if (f0[k] >= f1[k]) bad = 1;
        }
        long long u = 0, t = 0;
        double v[3] = {0, 0, 0};
        int ai = -1;
        if (!bad) bad = orc_field_number(f0[0], f1[0], 1, &u, NULL) || u < -2147483647LL || u > 2147483647LL;
        if (!bad) {
            size_t L = (size_t)(f1[1] - f0[1]);
            for (int k = 0; k < num_names && ai < 0; k++) if (strlen(names[k]) == L && memcmp(names[k], f0[1], L) == 0) ai = k;
            bad = ai < 0;
        }
        if (!bad) bad = orc_field_number(f0[2], f1[2], 1, &t, NULL);
        for (int k = 0; k < 3 && !bad; k++) {
            /* a decimal needs at least one digit in its mantissa ("." or "-" alone are malformed) */
            int has_digit = 0;
            for (const char* p = f0[3 + k]; p < f1[3 + k] && *p != 'e' && *p != 'E'; p++) has_digit |= (*p >= '0' && *p <= '9');
            bad = !has_digit || orc_field_number(f0[3 + k], f1[3 + k], 0, NULL, &v[k]);
        }
        if (bad) { ndrop++; continue; }
        if (nvalid < capacity) { user[nvalid] = (int32_t)u; act[nvalid] = ai; ts[nvalid] = t; x[nvalid] = v[0]; y[nvalid] = v[1]; z[nvalid] = v[2]; }
        nvalid++;
    }
    if (dropped) *dropped = ndrop;
    return nvalid;
}
