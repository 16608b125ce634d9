/*
 * oracle_miner.c - CPU restatement of the training-side consumer of the k-NN: the most-similar-template table and the
 * bootstrapping triplet miner (TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may
 * load it; see pose_oracle.h).
 *
 *   orc_max_sim_tmpl      src/datasetmanager.cpp:687-734 (inplane: two smallest quaternion angular distances) and
 *                        :736-783 (otherwise: two largest translation dot products, started from
 *                        numeric_limits<float>::min(), i.e. the smallest POSITIVE float, as the reference writes it)
 *   orc_triplet_indices   src/networksolver.cpp:52-186 (buildBatch): anchor / puller / pusher0-2 selection
 *   orc_gather_batch      src/networksolver.cpp:32-40 (HWC -> Caffe CHW batch) + bgfill::randomColorFill
 *                        (src/bgfill.cpp:136-158) for input.random_background = 1
 *
 * Random numbers.  The reference draws from std::random_device (include/networksolver.h:53), so no two of its runs
 * agree and only the distribution can be restated.  Both sides of the parity check (this file and csrc/miner.cu) draw
 * from the same counter-based generator instead: value j of triplet `linearId` under `seed` is
 * mix64(mix64(seed ^ linearId * K1) + j * K2) with the splitmix64 finaliser; an index below n is the high half of the
 * 128-bit product with n, a uniform float is the top 24 bits * 2^-24.  The draws are consumed in the order the
 * reference's statements make theirs.  Rejection loops the reference would spin in forever (every template inside the
 * elevation band, a single object) stop after ORC_MAX_DRAWS and raise the triplet's flag.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#include "pose_oracle.h"

#define K1 0x9E3779B97F4A7C15ull
#define K2 0xD1B54A32D192ED03ull
#define K3 0x8CB92BA72F3D8DD7ull

static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t orc_miner_key(uint64_t seed, uint64_t linear_id) { return mix64(seed ^ (linear_id * K1)); }
uint64_t orc_miner_draw(uint64_t key, uint64_t j) { return mix64(key + j * K2); }
static inline uint64_t below(uint64_t r, uint64_t n) { return (uint64_t)(((unsigned __int128)r * n) >> 64); }

static inline float acos_cr(float x) { return (float)acos((double)x); }

/* src/datasetmanager.cpp:687-783.  labels: [n_objects * poses][8] = {class, qx, qy, qz, qw, tx, ty, tz};
 * out: [n_objects * nr_training_poses][2], -1 where the reference would push an uninitialised int */
void orc_max_sim_tmpl(const float* tmpl_labels, const float* train_labels, int n_objects, int nr_template_poses,
                      int nr_training_poses, int inplane, int32_t* out) {
    for (int object = 0; object < n_objects; ++object)
        for (int tp = 0; tp < nr_training_poses; ++tp) {
            const float* a = train_labels + ((size_t)object * nr_training_poses + tp) * 8;
            const float* tl = tmpl_labels + (size_t)object * nr_template_poses * 8;
            int s1 = -1, s2 = -1;
            if (inplane) {
                float best = FLT_MAX, best2 = FLT_MAX;
                for (int t = 0; t < nr_template_poses; ++t) {
                    const float d = orc_angular_distance(a + 1, tl + (size_t)t * 8 + 1);
                    if (d >= best) continue;
                    best = d;
                    s1 = t;
                }
                for (int t = 0; t < nr_template_poses; ++t) {
                    const float d = orc_angular_distance(a + 1, tl + (size_t)t * 8 + 1);
                    if (d >= best2 || d == best) continue;
                    best2 = d;
                    s2 = t;
                }
            } else {
                float best = FLT_MIN, best2 = FLT_MIN;
                for (int t = 0; t < nr_template_poses; ++t) {
                    const float* b = tl + (size_t)t * 8;
                    float d = a[5] * b[5];
                    d = d + a[6] * b[6];
                    d = d + a[7] * b[7];
                    if (d <= best) continue;
                    best = d;
                    s1 = t;
                }
                for (int t = 0; t < nr_template_poses; ++t) {
                    const float* b = tl + (size_t)t * 8;
                    float d = a[5] * b[5];
                    d = d + a[6] * b[6];
                    d = d + a[7] * b[7];
                    if (d <= best2 || d == best) continue;
                    best2 = d;
                    s2 = t;
                }
            }
            out[((size_t)object * nr_training_poses + tp) * 2 + 0] = s1;
            out[((size_t)object * nr_training_poses + tp) * 2 + 1] = s2;
        }
}

/* elevation of a template as buildBatch compares it: acos(R(2,2)) of its quaternion (:83, :118, :148) */
float orc_elevation(const float* label) { return acos_cr(orc_rot22(label + 1)); }

typedef struct {
    uint64_t key, j;
} draws_t;
static inline uint64_t next(draws_t* d) { return orc_miner_draw(d->key, d->j++); }

/* random template of `object` whose elevation differs from the puller's by >= 0.2 (:81-84) */
static int pick_elevation(draws_t* d, const float* tmpl_labels, int object, int nr_template_poses, int puller, int* flag) {
    const float* tl = tmpl_labels + (size_t)object * nr_template_poses * 8;
    const float ep = orc_elevation(tl + (size_t)puller * 8);
    int p = (int)below(next(d), nr_template_poses);
    int tries = 1;
    while ((double)fabsf(orc_elevation(tl + (size_t)p * 8) - ep) < 0.2) {
        if (tries++ >= ORC_MAX_DRAWS) { *flag = 1; break; }
        p = (int)below(next(d), nr_template_poses);
    }
    return p;
}

/* src/networksolver.cpp:52-186.  max_sim: [n_objects * nr_training_poses][2]; knn_table / knn_len: maxSimKNNTmpl as
 * orc_knn_bootstrap writes it, or NULL for bootstrapping == false; rot_inv[object] = rotInv[model_index[used_models[
 * object]]]; neg_offsets: n_objects + 1 prefix offsets into the hard-negative pool or NULL.
 * out_idx: [batch_size / triplet_size][5] = {anchor (training linear index), puller, pusher0, pusher1, pusher2
 * (template linear indices)}; a hard negative is written as ~index into the negative pool.  out_flag[t] != 0 when a
 * rejection loop gave up. */
void orc_triplet_indices(const float* tmpl_labels, const int32_t* max_sim, const int32_t* knn_table, const int32_t* knn_len,
                         const int* rot_inv, const int64_t* neg_offsets, int n_objects, int nr_template_poses,
                         int nr_training_poses, int batch_size, int triplet_size, int iter, uint64_t seed,
                         int32_t* out_idx, int32_t* out_flag) {
    const int per = batch_size / triplet_size;
    for (int i = 0; i < per; ++i) {
        const int64_t linearId = (int64_t)iter * batch_size / triplet_size + i;
        const int training_pose = (int)(linearId / n_objects), object = (int)(linearId % n_objects);
        draws_t d = {orc_miner_key(seed, (uint64_t)linearId), 0};
        int flag = 0;
        const size_t q = (size_t)object * nr_training_poses + training_pose;
        int puller = max_sim[q * 2];
        if (puller < 0) { puller = 0; flag = 2; }   /* the reference would index with an uninitialised int */
        const int tbase = object * nr_template_poses;
        int pusher0, pusher1, pusher2;
        int32_t p0_out, p1_out, p2_out;
        if (rot_inv[object] != 0) {
            pusher0 = pick_elevation(&d, tmpl_labels, object, nr_template_poses, puller, &flag);
        } else {
            pusher0 = (int)below(next(&d), nr_template_poses);
            int tries = 1;
            while (pusher0 == puller) {
                if (tries++ >= ORC_MAX_DRAWS) { flag = 1; break; }
                pusher0 = (int)below(next(&d), nr_template_poses);
            }
        }
        p0_out = tbase + pusher0;
        if (neg_offsets && neg_offsets[object + 1] > neg_offsets[object]) {     /* :95-100 */
            const int64_t n = neg_offsets[object + 1] - neg_offsets[object];
            p0_out = ~(int32_t)(neg_offsets[object] + (int64_t)below(next(&d), (uint64_t)n));
        }
        int knn_object = -1, knn_pose = -1, missed = 0;
        if (knn_table) {
            knn_object = knn_table[q * 4];
            knn_pose = knn_table[q * 4 + 1];
            missed = knn_object != object || knn_pose != puller;
        }
        if (missed) {
            p1_out = knn_object * nr_template_poses + knn_pose;                 /* :109-111 */
        } else {
            if (rot_inv[object] != 0) {
                pusher1 = pick_elevation(&d, tmpl_labels, object, nr_template_poses, puller, &flag);
            } else {
                pusher1 = (int)below(next(&d), nr_template_poses);
                int tries = 1;
                while (pusher1 == puller && pusher1 == pusher0) {              /* `&&` as the reference has it */
                    if (tries++ >= ORC_MAX_DRAWS) { flag = 1; break; }
                    pusher1 = (int)below(next(&d), nr_template_poses);
                }
            }
            p1_out = tbase + pusher1;
        }
        if (knn_table && knn_len[q] > 2 && missed) {                            /* :132-136 */
            p2_out = knn_table[q * 4 + 2] * nr_template_poses + knn_table[q * 4 + 3];
        } else {
            pusher2 = (int)below(next(&d), n_objects);
            int tries = 1;
            while (pusher2 == object) {
                if (tries++ >= ORC_MAX_DRAWS) { flag = 1; break; }
                pusher2 = (int)below(next(&d), n_objects);
            }
            p2_out = pusher2 * nr_template_poses + (int)below(next(&d), nr_template_poses);
        }
        out_idx[i * 5 + 0] = (int32_t)q;
        out_idx[i * 5 + 1] = tbase + puller;
        out_idx[i * 5 + 2] = p0_out;
        out_idx[i * 5 + 3] = p1_out;
        out_idx[i * 5 + 4] = p2_out;
        out_flag[i] = flag;
    }
}

/* src/networksolver.cpp:32-40 over the five samples of every triplet (+ :169-171 with random_background == 1:
 * src/bgfill.cpp:136-158 on the anchor - mask = depth channel == 0, dilated 3x3, every masked pixel gets seven uniform
 * floats).  Pools are HWC [*][64][64][C]; out is [n_triplets * 5][C][64][64].  The fill's value for (row, col, ch) of
 * triplet linearId is draw ((row * 64 + col) * 7 + ch) under key mix64(seed ^ linearId * K1 ^ K3). */
void orc_gather_batch(const float* tmpl_pool, const float* train_pool, const float* neg_pool, const int32_t* idx,
                      int n_triplets, int channels, int64_t first_linear_id, uint64_t seed, int random_background, float* out) {
    const size_t img = (size_t)64 * 64 * channels;
    for (int t = 0; t < n_triplets; ++t)
        for (int s = 0; s < 5; ++s) {
            const int32_t id = idx[t * 5 + s];
            const float* src = s == 0 ? train_pool + (size_t)id * img : (id < 0 ? neg_pool + (size_t)(~id) * img : tmpl_pool + (size_t)id * img);
            float* dst = out + ((size_t)t * 5 + s) * img;
            for (int h = 0; h < 64; ++h)
                for (int w = 0; w < 64; ++w)
                    for (int c = 0; c < channels; ++c) dst[((size_t)c * 64 + h) * 64 + w] = src[((size_t)h * 64 + w) * channels + c];
            if (s == 0 && random_background == 1 && channels == 7) {
                const uint64_t key = mix64(seed ^ ((uint64_t)(first_linear_id + t) * K1) ^ K3);
                for (int h = 0; h < 64; ++h)
                    for (int w = 0; w < 64; ++w) {
                        int m = 0;
                        for (int dy = -1; dy <= 1 && !m; ++dy)
                            for (int dx = -1; dx <= 1; ++dx) {
                                const int y = h + dy, x = w + dx;
                                if (y < 0 || y > 63 || x < 0 || x > 63) continue;
                                if (src[((size_t)y * 64 + x) * 7 + 3] == 0.f) { m = 1; break; }
                            }
                        if (!m) continue;
                        for (int c = 0; c < 7; ++c)
                            dst[((size_t)c * 64 + h) * 64 + w] =
                                (float)(orc_miner_draw(key, (uint64_t)((h * 64 + w) * 7 + c)) >> 40) * (1.0f / 16777216.0f);
                    }
            }
        }
}
