/*
 * oracle_match.c - CPU restatement of the matcher and the pose-error reductions.
 * TEST INFRASTRUCTURE ONLY - see pose_oracle.h for the parity statement.
 * Compiled with -ffp-contract=off so distances are reproducible bit for bit.
 *
 * Restated third-party semantics (libraries not vendored by the reference):
 *   OpenCV 3.1 BFMatcher(NORM_L2)::knnMatch  - call sites src/networkevaluator.cpp:193-197,
 *       263-267 and src/networksolver.cpp:466-472: dist = sqrt(sum_d (q_d - t_d)^2) in fp32,
 *       k smallest ascending; equal distances keep the lower train index (batchDistance
 *       scans templates in order and only replaces on strictly smaller).
 *   Eigen Quaternionf::angularDistance       - src/networkevaluator.cpp:297,331.  ASSUMED VERSION: Eigen 3.2.x (what
 *       `libeigen3-dev` installs on the Ubuntu 14.04 / 16.04 + CUDA 7.5 / 8.0 systems of README.md:10-14), whose
 *       Quaternion.h computes  d = |a.b|; if (d >= 1) return 0; return 2*acos(d)  =  2*acos(min(1,|a.b|)).
 *       Eigen 3.3+ instead returns 2*atan2(|(a * conj(b)).vec()|, |(a * conj(b)).w()|): the same angle, better conditioned
 *       near 0 degrees.  That form is restated too (orc_angular_distance_atan2, selected with orc_set_angular_form(1));
 *       tests/test_oracle_golden.py bounds the disagreement of the two (<= 0.07 deg near 0, <= 1e-3 deg beyond 5 deg,
 *       a handful of bin / exact-match flips per 3,000 queries).  The CUDA path implements the 3.2 form.
 *       std::acos(float) is restated as the correctly rounded value (float)acos((double)x):
 *       any libm acosf is within 1 ulp of it, and it is reproducible bit for bit on the GPU,
 *       so the strict `<` of the exact-match scan (:331) agrees between oracle and CUDA.
 *   Eigen toRotationMatrix()(2,2)            - src/networkevaluator.cpp:299:
 *       tx=2x, ty=2y, txx=tx*x, tyy=ty*y, R22 = 1-(txx+tyy).
 * Own glue logic restated from src/networkevaluator.cpp:153-175,199-245,269-355 and
 * src/networksolver.cpp:474-504.
 */
#include "pose_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------ L2 distance --- */
/* Squared distance with an 8-lane blocked sum (the shape of a SIMD normL2Sqr):
 * lane l accumulates dims l, l+8, l+16 ...; lanes are then combined by the fixed
 * tree ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)).  Multiply and add are rounded
 * separately.  The CUDA refine step reproduces exactly this expression. */
static inline float l2_sq(const float* a, const float* b, int dim) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= dim; i += 8)
        for (int l = 0; l < 8; ++l) {
            const float d = a[i + l] - b[i + l];
            acc[l] = acc[l] + d * d;
        }
    for (int l = 0; i < dim; ++i, ++l) {
        const float d = a[i] - b[i];
        acc[l] = acc[l] + d * d;
    }
    return ((acc[0] + acc[4]) + (acc[2] + acc[6])) + ((acc[1] + acc[5]) + (acc[3] + acc[7]));
}

float orc_l2_distance(const float* a, const float* b, int dim) { return sqrtf(l2_sq(a, b, dim)); }

#define QB 4 /* queries scanned together so a template row is read once per block */

void orc_knn_l2(const float* q, int64_t nq, const float* t, int64_t nt, int dim, int k,
                int32_t* idx, float* dist, float* gap, int threads) {
    const int kk = k + 1; /* track one extra neighbour for the gap */
#ifdef _OPENMP
    const int nth = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel num_threads(nth)
#endif
    {
        float* bd = (float*)malloc(sizeof(float) * QB * kk);
        int32_t* bi = (int32_t*)malloc(sizeof(int32_t) * QB * kk);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t q0 = 0; q0 < nq; q0 += QB) {
            const int nb = (int)(nq - q0 < QB ? nq - q0 : QB);
            for (int j = 0; j < nb * kk; ++j) { bd[j] = INFINITY; bi[j] = -1; }
            for (int64_t ti = 0; ti < nt; ++ti) {
                const float* tr = t + (size_t)ti * dim;
                for (int b = 0; b < nb; ++b) {
                    const float d = sqrtf(l2_sq(q + (size_t)(q0 + b) * dim, tr, dim));
                    float* ld = bd + b * kk;
                    int32_t* li = bi + b * kk;
                    if (!(d < ld[kk - 1])) continue; /* ties keep the earlier (lower) index */
                    int p = kk - 1;
                    while (p > 0 && d < ld[p - 1]) { ld[p] = ld[p - 1]; li[p] = li[p - 1]; --p; }
                    ld[p] = d; li[p] = (int32_t)ti;
                }
            }
            for (int b = 0; b < nb; ++b) {
                for (int j = 0; j < k; ++j) {
                    idx[(size_t)(q0 + b) * k + j] = bi[b * kk + j];
                    dist[(size_t)(q0 + b) * k + j] = bd[b * kk + j];
                }
                if (gap) {
                    const float dk = bd[b * kk + k - 1], dn = bd[b * kk + k];
                    gap[q0 + b] = (isinf(dn) || dn <= 0.f) ? INFINITY : (dn - dk) / dn;
                }
            }
        }
        free(bd); free(bi);
    }
    (void)threads;
}

/* ---------------------------------------------------------------- Hamming --- */
void orc_binarize(const float* desc, int64_t n, int dim, uint32_t* bits) {
    const int words = (dim + 31) / 32;
    for (int64_t i = 0; i < n; ++i)
        for (int w = 0; w < words; ++w) {
            uint32_t v = 0;
            for (int j = 0; j < 32 && w * 32 + j < dim; ++j)
                if (desc[(size_t)i * dim + w * 32 + j] > 0.f) v |= (1u << j);
            bits[(size_t)i * words + w] = v;
        }
}

void orc_knn_hamming(const uint32_t* q, int64_t nq, const uint32_t* t, int64_t nt, int words, int k,
                     int32_t* idx, int32_t* dist, int threads) {
#ifdef _OPENMP
    const int nth = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 8) num_threads(nth)
#endif
    for (int64_t qi = 0; qi < nq; ++qi) {
        int32_t* li = idx + (size_t)qi * k;
        int32_t* ld = dist + (size_t)qi * k;
        for (int j = 0; j < k; ++j) { li[j] = -1; ld[j] = INT32_MAX; }
        const uint32_t* qr = q + (size_t)qi * words;
        for (int64_t ti = 0; ti < nt; ++ti) {
            const uint32_t* tr = t + (size_t)ti * words;
            int32_t d = 0;
            for (int w = 0; w < words; ++w) d += __builtin_popcount(qr[w] ^ tr[w]);
            if (!(d < ld[k - 1])) continue;
            int p = k - 1;
            while (p > 0 && d < ld[p - 1]) { ld[p] = ld[p - 1]; li[p] = li[p - 1]; --p; }
            ld[p] = d; li[p] = (int32_t)ti;
        }
    }
    (void)threads;
}

/* ------------------------------------------------------------ shard merge --- */
/* candidates of shard s for query q live at [(s*nq + q)*k ...]; order by
 * (distance, global index); -1 entries are padding */
#define MERGE_BODY(DT, BIG)                                                              \
    for (int64_t qi = 0; qi < nq; ++qi) {                                                \
        int32_t* oi = out_idx + (size_t)qi * k;                                          \
        DT* od = out_dist + (size_t)qi * k;                                              \
        for (int j = 0; j < k; ++j) { oi[j] = -1; od[j] = BIG; }                         \
        for (int s = 0; s < n_shards; ++s)                                               \
            for (int j = 0; j < k; ++j) {                                                \
                const int32_t ci = idx[((size_t)s * nq + qi) * k + j];                   \
                const DT cd = dist[((size_t)s * nq + qi) * k + j];                       \
                if (ci < 0) continue;                                                    \
                int p = k;                                                               \
                while (p > 0 && (oi[p - 1] < 0 || cd < od[p - 1] ||                      \
                                 (cd == od[p - 1] && ci < oi[p - 1]))) --p;              \
                if (p == k) continue;                                                    \
                for (int m = k - 1; m > p; --m) { oi[m] = oi[m - 1]; od[m] = od[m - 1]; } \
                oi[p] = ci; od[p] = cd;                                                  \
            }                                                                            \
    }

void orc_merge_topk_f32(const int32_t* idx, const float* dist, int n_shards, int64_t nq, int k,
                        int32_t* out_idx, float* out_dist) {
    MERGE_BODY(float, INFINITY)
}
void orc_merge_topk_i32(const int32_t* idx, const int32_t* dist, int n_shards, int64_t nq, int k,
                        int32_t* out_idx, int32_t* out_dist) {
    MERGE_BODY(int32_t, INT32_MAX)
}

/* ------------------------------------------------------------------ Eigen --- */
static inline float acos_cr(float x) { return (float)acos((double)x); }

static int g_angular_form = 0;      /* 0: Eigen 3.2 (acos of the dot product), 1: Eigen 3.3+ (atan2 of q * conj(k)) */
void orc_set_angular_form(int form) { g_angular_form = form ? 1 : 0; }

/* Eigen 3.3+ Quaternion.h angularDistance: d = a * conj(b); 2 * atan2(d.vec().norm(), abs(d.w())), all in float */
float orc_angular_distance_atan2(const float* a, const float* b) {
    /* coefficients are (x, y, z, w); conj(b) = (-bx, -by, -bz, bw); quaternion product a * conj(b) */
    const float ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = -b[0], by = -b[1], bz = -b[2], bw = b[3];
    const float w = aw * bw - ax * bx - ay * by - az * bz;
    const float x = aw * bx + ax * bw + ay * bz - az * by;
    const float y = aw * by + ay * bw + az * bx - ax * bz;
    const float z = aw * bz + az * bw + ax * by - ay * bx;
    return 2.0f * atan2f(sqrtf(x * x + y * y + z * z), fabsf(w));
}

float orc_angular_distance(const float* a, const float* b) {
    if (g_angular_form) return orc_angular_distance_atan2(a, b);
    float d = a[0] * b[0];
    d = d + a[1] * b[1];
    d = d + a[2] * b[2];
    d = d + a[3] * b[3];
    d = fabsf(d);
    if (d > 1.0f) d = 1.0f;
    return 2.0f * acos_cr(d);
}

float orc_rot22(const float* q) {
    const float tx = 2.0f * q[0], ty = 2.0f * q[1];
    const float txx = tx * q[0], tyy = ty * q[1];
    return 1.0f - (txx + tyy);
}

/* `x*180.f/M_PI` as the reference writes it: float*float, then a double division,
 * then narrowed on assignment to a float (src/networkevaluator.cpp:297,299) */
static inline float to_degrees(float rad) { return (float)((double)(rad * 180.f) / M_PI); }

/* ---------------------------------------------------- computeHistogram body --- */
static int cmp_float(const void* a, const void* b) {
    const float x = *(const float*)a, y = *(const float*)b;
    return (x > y) - (x < y);
}

void orc_histogram(const int32_t* knn_idx, int64_t nq, int k, const float* q_labels,
                   const float* t_labels, int nr_tmpl_poses, const int* rot_inv, const float* bins,
                   int n_bins, float* out, float* per_query_err) {
    float* histo = out;
    for (int b = 0; b < n_bins; ++b) histo[b] = 0.f;
    float mean_angle = 0.f;
    float* diffs = (float*)malloc(sizeof(float) * (size_t)(nq > 0 ? nq : 1));
    int64_t n_diffs = 0;

    for (int64_t qi = 0; qi < nq; ++qi) {
        const float* ql = q_labels + (size_t)qi * 8;
        const float* qquat = ql + 1;
        float best = FLT_MAX;
        int64_t best_t = -1;
        for (int nn = 0; nn < k; ++nn) {
            const int32_t ti = knn_idx[(size_t)qi * k + nn];
            if (ti < 0) continue;
            const float* tl = t_labels + (size_t)ti * 8;
            if (ql[0] != tl[0]) continue; /* :288 class test on the float labels */
            float d;
            if (rot_inv[(int)ql[0]] == 0) {
                d = to_degrees(orc_angular_distance(qquat, tl + 1)); /* :297 */
            } else {
                d = to_degrees(fabsf(acos_cr(orc_rot22(qquat)) - acos_cr(orc_rot22(tl + 1)))); /* :299 */
                if (isnan(d)) d = FLT_MAX;                                                 /* :300 */
            }
            if (d >= best) continue; /* :303 keeps the earlier neighbour on ties */
            best = d;
            best_t = ti;
        }
        if (per_query_err) per_query_err[qi] = best_t < 0 ? -1.f : best;
        if (best_t < 0) { histo[0] += 1.f; continue; } /* :311-313 */

        mean_angle += best;        /* :321 */
        diffs[n_diffs++] = best;   /* :322 */

        /* exact match: no template of the matched object is strictly closer (:326-334) */
        const int64_t obj0 = (best_t / nr_tmpl_poses) * (int64_t)nr_tmpl_poses;
        const float ref = orc_angular_distance(qquat, t_labels + (size_t)best_t * 8 + 1);
        int exact = 1;
        for (int p = 0; p < nr_tmpl_poses; ++p)
            if (orc_angular_distance(qquat, t_labels + (size_t)(obj0 + p) * 8 + 1) < ref) exact = 0;
        if (exact) histo[1] += 1.f;

        for (int b = 2; b < n_bins; ++b)
            if (best < bins[b]) histo[b] += 1.f; /* :337-338 cumulative */
    }

    const float total = histo[0] + histo[n_bins - 1]; /* :342 */
    for (int b = 0; b < n_bins; ++b) histo[b] /= total;
    mean_angle /= (float)nq; /* :346 all queries in the denominator */
    float median = NAN;      /* reference indexes an empty vector here; defined as NaN */
    if (n_diffs > 0) {
        qsort(diffs, (size_t)n_diffs, sizeof(float), cmp_float);
        median = diffs[(n_diffs - 1) / 2]; /* :350 */
    }
    out[n_bins] = mean_angle;
    out[n_bins + 1] = median;
    free(diffs);
}

/* ------------------------------------------------ computeConfusionMatrix body --- */
void orc_confusion(const int32_t* knn_idx, int64_t nq, int k, const float* q_labels,
                   const float* t_labels, const int* class_to_local, int n_local,
                   int nr_test_poses, float* out) {
    for (int i = 0; i < n_local * n_local; ++i) out[i] = 0.f;
    for (int64_t qi = 0; qi < nq; ++qi) {
        const float qc = q_labels[(size_t)qi * 8];
        int32_t pick = -1;
        for (int nn = 0; nn < k; ++nn) { /* first same-class neighbour wins (:212-219) */
            const int32_t ti = knn_idx[(size_t)qi * k + nn];
            if (ti >= 0 && t_labels[(size_t)ti * 8] == qc) { pick = ti; break; }
        }
        if (pick < 0) pick = knn_idx[(size_t)qi * k]; /* else the 1st neighbour (:222-231) */
        if (pick < 0) continue;
        const int r = class_to_local[(int)qc];
        const int c = class_to_local[(int)t_labels[(size_t)pick * 8]];
        out[r * n_local + c] += 1.f;
    }
    for (int i = 0; i < n_local * n_local; ++i) out[i] /= (float)nr_test_poses; /* :242-244 */
}

/* ------------------------------------------------ networkSolver::computeKNN body --- */
void orc_knn_bootstrap(const int32_t* knn3_idx, int64_t nq, int nr_query_poses, int nr_tmpl_poses,
                       int32_t* out, int32_t* out_len) {
    const int knn = 3;
    for (int64_t qi = 0; qi < nq; ++qi) {
        const int query_object = (int)(qi / nr_query_poses);
        int32_t* o = out + (size_t)qi * 4;
        int len = 0;
        o[0] = o[1] = o[2] = o[3] = -1;
        for (int nn = 0; nn < knn; ++nn) {
            const int32_t ti = knn3_idx[(size_t)qi * knn + nn];
            const int obj = ti / nr_tmpl_poses, pose = ti % nr_tmpl_poses;
            if (nn == 0) { o[len++] = obj; o[len++] = pose; }                 /* :492-495 */
            else if (len < 4 && obj != query_object) { o[len++] = obj; o[len++] = pose; } /* :497-501 */
        }
        out_len[qi] = len;
    }
}

void orc_knn_accuracy(const int32_t* max_sim_tmpl0, const int32_t* bootstrap, int64_t nq,
                      int nr_query_poses, float* intra_pct, float* inter_pct) {
    int intra = 0, inter = 0;
    for (int64_t qi = 0; qi < nq; ++qi) {
        const int object = (int)(qi / nr_query_poses);
        if (max_sim_tmpl0[qi] == bootstrap[(size_t)qi * 4 + 1]) intra++; /* :169 */
        if (bootstrap[(size_t)qi * 4 + 0] == object) inter++;           /* :170 */
    }
    *intra_pct = intra / (float)nq * 100;
    *inter_pct = inter / (float)nq * 100;
}
