/*
 * pose.cu - pose-error reductions over the kNN result on sm_100a.
 *
 * Replaces the per-query loops that follow knnMatch in the reference:
 *   pose_err_kernel    networkEvaluator::computeHistogram body (src/networkevaluator.cpp:273-339):
 *                      best same-class neighbour by angular error (Eigen angularDistance, or the
 *                      elevation difference |acos(Rq22)-acos(Rk22)| for rotInv classes, :296-301),
 *                      exact-match scan over the matched object's templates (:326-334),
 *                      cumulative bins (:337-338).  One warp per eight queries, the scan shared by queries
 *                      that matched the same object.
 *   median_select      the sort + middle element of :349-350 as a 4-pass radix select.
 *   confusion_kernel   computeConfusionMatrix body (:201-239).
 *   bootstrap_kernel   networkSolver::computeKNN body (src/networksolver.cpp:474-504).
 *   accuracy_kernel    computeKNNAccuracy (src/networkevaluator.cpp:153-175).
 * acos is evaluated as the correctly rounded float of a double acos on both sides of the parity
 * check (CPU oracle and here), so `<` comparisons between angles agree bit for bit.
 */
#include <cfloat>
#include <cmath>

#include <mutex>
#include <vector>

#include "common.h"

namespace pe {

namespace {

__device__ __forceinline__ float acos_cr(float x) { return __double2float_rn(acos((double)x)); }

/* |a.b| clamped to 1, products and sums rounded separately in the order Eigen's dot() of a
 * 4-vector evaluates them (oracle_match.c orc_angular_distance) */
__device__ __forceinline__ float quat_absdot(const float* __restrict__ a, const float* __restrict__ b) {
    float d = __fmul_rn(a[0], b[0]);
    d = __fadd_rn(d, __fmul_rn(a[1], b[1]));
    d = __fadd_rn(d, __fmul_rn(a[2], b[2]));
    d = __fadd_rn(d, __fmul_rn(a[3], b[3]));
    d = fabsf(d);
    return d > 1.0f ? 1.0f : d;
}

__device__ __forceinline__ float rot22(const float* __restrict__ q) {
    const float tx = __fmul_rn(2.0f, q[0]), ty = __fmul_rn(2.0f, q[1]);
    const float txx = __fmul_rn(tx, q[0]), tyy = __fmul_rn(ty, q[1]);
    return __fsub_rn(1.0f, __fadd_rn(txx, tyy));
}

/* `x*180.f/M_PI` as the reference writes it: float product, double division, narrowed to float */
__device__ __forceinline__ float to_degrees(float rad) {
    return __double2float_rn(__ddiv_rn((double)__fmul_rn(rad, 180.f), 3.14159265358979323846));
}

constexpr int kPoseWarps = 4;
constexpr int kPoseQ = 8;      /* consecutive queries per warp */
constexpr int kMaxBins = 16;

/* One warp per kPoseQ consecutive queries.  Step 1 (per query): lane nn < k scores neighbour nn, a warp reduction picks the
 * best same-class one.  Step 2: the exact-match scan over the matched object's templates - the expensive part: 32 bytes
 * of label per template, 3,334 templates per object in config 5 - runs once per DISTINCT matched object among the warp's
 * queries: a lane loads a template's quaternion once and scores it against every query of the group (query sets are laid
 * out [object][pose], so the eight queries normally share one object and the label traffic per query drops eightfold;
 * queries in any other order still work, a group is then a single query).  max() over |dot| is exact in any order, so the
 * grouping does not change a bit of the result. */
__global__ void __launch_bounds__(kPoseWarps * 32)
pose_err_kernel(const int* __restrict__ knn_idx, long long nq, int k, const float* __restrict__ q_labels,
                const float* __restrict__ t_labels, long long t_total, int nr_tmpl_poses,
                const int* __restrict__ rot_inv, int n_rot, const float* __restrict__ bins, int n_bins,
                float* __restrict__ err_out, unsigned long long* __restrict__ counts,
                double* __restrict__ sum_out) {
    __shared__ unsigned int s_cnt[kMaxBins];
    __shared__ double s_sum[kPoseWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < kMaxBins) s_cnt[threadIdx.x] = 0;
    if (lane == 0) s_sum[warp] = 0.0;
    __syncthreads();
    const long long q0 = ((long long)blockIdx.x * kPoseWarps + warp) * kPoseQ;
    float qv[kPoseQ][4], refv[kPoseQ], bdv[kPoseQ];        /* warp-uniform, indexed by unrolled loops only */
    long long objv[kPoseQ];
    unsigned pending = 0;
    double wsum = 0.0;
#pragma unroll
    for (int j = 0; j < kPoseQ; ++j) {
        const long long qi = q0 + j;
        objv[j] = -1;
        refv[j] = bdv[j] = 0.f;
        qv[j][0] = qv[j][1] = qv[j][2] = qv[j][3] = 0.f;
        if (qi >= nq) continue;
        const float* ql = q_labels + qi * 8;
        const float qc = ql[0];
        const float qq[4] = {ql[1], ql[2], ql[3], ql[4]};
        const int cls = (int)qc;
        const int inv = (cls >= 0 && cls < n_rot) ? rot_inv[cls] : 0;
        float d = FLT_MAX;
        int ti = -1;
        if (lane < k) {
            ti = knn_idx[qi * k + lane];
            if (ti >= 0 && ti < t_total) {
                const float* tl = t_labels + (long long)ti * 8;
                if (tl[0] == qc) {
                    if (inv == 0) {
                        d = to_degrees(__fmul_rn(2.0f, acos_cr(quat_absdot(qq, tl + 1))));
                    } else {
                        d = to_degrees(fabsf(__fsub_rn(acos_cr(rot22(qq)), acos_cr(rot22(tl + 1)))));
                        if (isnan(d)) d = FLT_MAX;
                    }
                }
            }
        }
        /* `if (knn_dist >= best_dist) continue` with best_dist starting at FLT_MAX:
         * smallest error wins, earlier neighbour on ties, FLT_MAX never wins */
        float bd = d;
        int bn = (d < FLT_MAX) ? lane : 64;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, off);
            const int on = __shfl_xor_sync(0xffffffffu, bn, off);
            if (od < bd || (od == bd && on < bn)) { bd = od; bn = on; }
        }
        if (bn >= 64) { /* no same-class neighbour */
            if (lane == 0) {
                err_out[qi] = -1.f;
                atomicAdd(&s_cnt[0], 1u);
            }
        } else {
            const int best_t = __shfl_sync(0xffffffffu, ti, bn);
            objv[j] = (long long)(best_t / nr_tmpl_poses) * nr_tmpl_poses;
            refv[j] = quat_absdot(qq, t_labels + (long long)best_t * 8 + 1);
            bdv[j] = bd;
            qv[j][0] = qq[0]; qv[j][1] = qq[1]; qv[j][2] = qq[2]; qv[j][3] = qq[3];
            pending |= 1u << j;
        }
    }
    /* exact match: no template of the matched object is strictly closer */
#pragma unroll
    for (int j0 = 0; j0 < kPoseQ; ++j0) {
        if (!((pending >> j0) & 1u)) continue;
        const long long obj0 = objv[j0];
        unsigned grp = 0;
#pragma unroll
        for (int j = 0; j < kPoseQ; ++j)
            if (((pending >> j) & 1u) && objv[j] == obj0) grp |= 1u << j;
        float mx[kPoseQ];
#pragma unroll
        for (int j = 0; j < kPoseQ; ++j) mx[j] = 0.f;
        for (int p = lane; p < nr_tmpl_poses; p += 32) {
            const long long t = obj0 + p;
            if (t >= t_total) break;
            const float4 a = __ldg(reinterpret_cast<const float4*>(t_labels + t * 8));
            const float4 b = __ldg(reinterpret_cast<const float4*>(t_labels + t * 8) + 1);
            const float tq[4] = {a.y, a.z, a.w, b.x};
#pragma unroll
            for (int j = 0; j < kPoseQ; ++j)
                if ((grp >> j) & 1u) mx[j] = fmaxf(mx[j], quat_absdot(qv[j], tq));
        }
#pragma unroll
        for (int j = 0; j < kPoseQ; ++j) {
            if (!((grp >> j) & 1u)) continue;
            float m = mx[j];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
            if (lane == 0) {
                const float ref = refv[j], bd = bdv[j];
                const bool exact = !(m > ref && acos_cr(m) < acos_cr(ref));
                err_out[q0 + j] = bd;
                if (exact) atomicAdd(&s_cnt[1], 1u);
                for (int b2 = 2; b2 < n_bins; ++b2)
                    if (bd < bins[b2]) atomicAdd(&s_cnt[b2], 1u);
                wsum += (double)bd;
            }
        }
        pending &= ~grp;
    }
    if (lane == 0) s_sum[warp] = wsum;
    __syncthreads();
    if (threadIdx.x < n_bins && s_cnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < kPoseWarps; ++w) s += s_sum[w];
        if (s != 0.0) atomicAdd(sum_out, s);
    }
}

/* element of rank (n_valid-1)/2 among the non-negative entries of err (radix select on the
 * float bit patterns, which order like the floats for values >= 0); NaN when none */
__global__ void __launch_bounds__(1024)
median_select_kernel(const float* __restrict__ err, long long n, float* __restrict__ out) {
    __shared__ unsigned int hist[256];
    __shared__ unsigned int s_prefix, s_mask;
    __shared__ unsigned long long s_rank, s_valid;
    const int tid = threadIdx.x;
    if (tid == 0) { s_valid = 0; s_prefix = 0; s_mask = 0; }
    __syncthreads();
    unsigned long long mine = 0;
    for (long long i = tid; i < n; i += blockDim.x) mine += err[i] >= 0.f;
    atomicAdd(&s_valid, mine);
    __syncthreads();
    if (s_valid == 0) {
        if (tid == 0) *out = nanf("");
        return;
    }
    if (tid == 0) s_rank = (s_valid - 1) / 2;
    for (int pass = 3; pass >= 0; --pass) {
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        const unsigned int prefix = s_prefix, mask = s_mask;
        const int shift = pass * 8;
        for (long long i = tid; i < n; i += blockDim.x) {
            const float v = err[i];
            if (!(v >= 0.f)) continue;
            const unsigned int u = __float_as_uint(v);
            if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            unsigned long long r = s_rank, cum = 0;
            int b = 0;
            for (; b < 256; ++b) {
                if (cum + hist[b] > r) break;
                cum += hist[b];
            }
            s_rank = r - cum;
            s_prefix = prefix | ((unsigned int)b << shift);
            s_mask = mask | (255u << shift);
        }
        __syncthreads();
    }
    if (tid == 0) *out = __uint_as_float(s_prefix);
}

__global__ void confusion_kernel(const int* __restrict__ knn_idx, long long nq, int k,
                                 const float* __restrict__ q_labels, const float* __restrict__ t_labels,
                                 long long t_total, const int* __restrict__ class_to_local, int n_classes,
                                 int n_local, unsigned int* __restrict__ mat, int* __restrict__ bad) {
    const long long qi = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    const float qc = q_labels[qi * 8];
    int pick = -1;
    for (int nn = 0; nn < k; ++nn) { /* first same-class neighbour wins */
        const int ti = knn_idx[qi * k + nn];
        if (ti >= 0 && ti < t_total && t_labels[(long long)ti * 8] == qc) { pick = ti; break; }
    }
    if (pick < 0) pick = knn_idx[qi * k]; /* else the 1st neighbour */
    if (pick < 0 || pick >= t_total) return;
    const int qcls = (int)qc, tcls = (int)t_labels[(long long)pick * 8];
    if (qcls < 0 || qcls >= n_classes || tcls < 0 || tcls >= n_classes) { atomicExch(bad, 1); return; }
    const int r = class_to_local[qcls], c = class_to_local[tcls];
    if (r < 0 || r >= n_local || c < 0 || c >= n_local) { atomicExch(bad, 1); return; }
    atomicAdd(&mat[r * n_local + c], 1u);
}

__global__ void bootstrap_kernel(const int* __restrict__ knn3, long long nq, int nr_query_poses,
                                 int nr_tmpl_poses, int* __restrict__ out, int* __restrict__ out_len) {
    const long long qi = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    const int query_object = (int)(qi / nr_query_poses);
    int o[4] = {-1, -1, -1, -1};
    int len = 0;
    for (int nn = 0; nn < 3; ++nn) {
        const int ti = knn3[qi * 3 + nn];
        const int obj = ti / nr_tmpl_poses, pose = ti % nr_tmpl_poses;
        if (nn == 0) { o[0] = obj; o[1] = pose; len = 2; }
        else if (len < 4 && obj != query_object) { o[len] = obj; o[len + 1] = pose; len += 2; }
    }
    out[qi * 4 + 0] = o[0]; out[qi * 4 + 1] = o[1]; out[qi * 4 + 2] = o[2]; out[qi * 4 + 3] = o[3];
    out_len[qi] = len;
}

__global__ void accuracy_kernel(const int* __restrict__ max_sim0, const int* __restrict__ boot, long long nq,
                                int nr_query_poses, unsigned long long* __restrict__ counts) {
    const long long qi = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    int intra = 0, inter = 0;
    if (qi < nq) {
        intra = max_sim0[qi] == boot[qi * 4 + 1];
        inter = boot[qi * 4 + 0] == (int)(qi / nr_query_poses);
    }
    const unsigned bi = __ballot_sync(0xffffffffu, intra), be = __ballot_sync(0xffffffffu, inter);
    if ((threadIdx.x & 31) == 0) {
        if (bi) atomicAdd(&counts[0], (unsigned long long)__popc(bi));
        if (be) atomicAdd(&counts[1], (unsigned long long)__popc(be));
    }
}

}  // namespace

/* first half of the histogram: per-query errors, bin counts and the error sum of nq queries, left on the device in
 * db->scratch ([counts: 16 x u64][sum: f64][median: f32 (+pad)][rot_inv][bins][err: err_cap x f32]); err_cap >= nq leaves
 * room for the errors of other ranks (sharded evaluation) */
int pose_hist_local(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k, const int* rot_inv, int n_rot,
                    const float* bins, int n_bins, int64_t err_cap, unsigned long long** d_counts_out, double** d_sum_out,
                    float** d_med_out, float** d_err_out) {
    PE_REQUIRE(db->d_labels != nullptr, "DB was created without labels");
    PE_REQUIRE(n_bins >= 2 && n_bins <= kMaxBins, "n_bins=%d outside 2..%d", n_bins, kMaxBins);
    PE_REQUIRE(k >= 1 && k <= 32, "k=%d outside 1..32", k);
    PE_REQUIRE(db->poses_per_object > 0, "poses_per_object not set");
    PE_CUDA(cudaSetDevice(db->device));
    const size_t off_sum = kMaxBins * 8, off_med = off_sum + 8, off_rot = off_med + 8;
    const size_t off_bins = off_rot + (((size_t)n_rot * 4 + 15) & ~(size_t)15);
    const size_t off_err = off_bins + kMaxBins * 4;
    if (err_cap < nq) err_cap = nq;
    PE_CHECK(db->scratch.reserve(off_err + (size_t)(err_cap > 0 ? err_cap : 1) * 4));
    char* base = db->scratch.as<char>();
    PE_CUDA(cudaMemsetAsync(base, 0, off_rot, db->stream));
    if (n_rot > 0) PE_CUDA(cudaMemcpyAsync(base + off_rot, rot_inv, (size_t)n_rot * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CUDA(cudaMemcpyAsync(base + off_bins, bins, (size_t)n_bins * 4, cudaMemcpyHostToDevice, db->stream));
    unsigned long long* d_counts = reinterpret_cast<unsigned long long*>(base);
    double* d_sum = reinterpret_cast<double*>(base + off_sum);
    float* d_err = reinterpret_cast<float*>(base + off_err);
    if (nq > 0) {
        pose_err_kernel<<<(unsigned)((nq + kPoseWarps * kPoseQ - 1) / (kPoseWarps * kPoseQ)), kPoseWarps * 32, 0, db->stream>>>(
            d_knn_idx, nq, k, d_q_labels, db->d_labels, db->t_total, db->poses_per_object,
            reinterpret_cast<const int*>(base + off_rot), n_rot, reinterpret_cast<const float*>(base + off_bins),
            n_bins, d_err, d_counts, d_sum);
        PE_LAUNCHED();
    }
    *d_counts_out = d_counts;
    *d_sum_out = d_sum;
    *d_med_out = reinterpret_cast<float*>(base + off_med);
    *d_err_out = d_err;
    return PE_OK;
}

/* second half: median over n_err per-query errors (entries < 0 = "no same-class neighbour" are skipped), then the
 * normalisation of src/networkevaluator.cpp:342-350 over nq_total queries */
int pose_hist_finish(pe_db* db, const unsigned long long* d_counts, const double* d_sum, float* d_med, const float* d_err,
                     int64_t n_err, int64_t nq_total, int n_bins, float* out_host) {
    median_select_kernel<<<1, 1024, 0, db->stream>>>(d_err, n_err, d_med);
    PE_LAUNCHED();
    /* counts, sum and median are neighbours in the scratch (pose_hist_local's layout): one copy back */
    struct { unsigned long long counts[kMaxBins]; double sum; float med; float pad; } h;
    static_assert(sizeof(h) == kMaxBins * 8 + 16, "layout of the head of db->scratch");
    PE_REQUIRE(reinterpret_cast<const char*>(d_sum) == reinterpret_cast<const char*>(d_counts) + kMaxBins * 8 &&
               reinterpret_cast<const char*>(d_med) == reinterpret_cast<const char*>(d_sum) + 8, "histogram scratch layout");
    PE_CUDA(cudaMemcpyAsync(&h, d_counts, sizeof(h), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    const unsigned long long* h_counts = h.counts;
    const double h_sum = h.sum;
    const float h_med = h.med;
    /* normalisation of src/networkevaluator.cpp:342-346 */
    const float total = (float)h_counts[0] + (float)h_counts[n_bins - 1];
    for (int b = 0; b < n_bins; ++b) out_host[b] = (float)h_counts[b] / total;
    out_host[n_bins] = (float)(h_sum / (double)nq_total);
    out_host[n_bins + 1] = h_med;
    return PE_OK;
}

/* all pointers device-resident; out_host: n_bins + 2 floats */
int pose_histogram_dev(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                       const int* rot_inv, int n_rot, const float* bins, int n_bins, float* out_host) {
    unsigned long long* d_counts = nullptr;
    double* d_sum = nullptr;
    float *d_med = nullptr, *d_err = nullptr;
    PE_CHECK(pose_hist_local(db, d_knn_idx, d_q_labels, nq, k, rot_inv, n_rot, bins, n_bins, nq, &d_counts, &d_sum, &d_med, &d_err));
    return pose_hist_finish(db, d_counts, d_sum, d_med, d_err, nq, nq, n_bins, out_host);
}

/* confusion counts of nq queries, left on the device: db->scratch = [n_local^2 x u32][bad flag][class map] */
int pose_confusion_local(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k, const int* class_to_local,
                         int n_classes, int n_local, unsigned int** d_mat_out) {
    PE_REQUIRE(db->d_labels != nullptr, "DB was created without labels");
    PE_REQUIRE(n_classes > 0 && n_local > 0 && k >= 1, "bad confusion-matrix arguments");
    PE_CUDA(cudaSetDevice(db->device));
    const size_t n_mat = (size_t)n_local * n_local;
    const size_t off_map = (n_mat * 4 + 16 + 15) & ~(size_t)15;
    PE_CHECK(db->scratch.reserve(off_map + (size_t)n_classes * 4));
    char* base = db->scratch.as<char>();
    PE_CUDA(cudaMemsetAsync(base, 0, off_map, db->stream));
    PE_CUDA(cudaMemcpyAsync(base + off_map, class_to_local, (size_t)n_classes * 4, cudaMemcpyHostToDevice, db->stream));
    unsigned int* d_mat = reinterpret_cast<unsigned int*>(base);
    int* d_bad = reinterpret_cast<int*>(base + n_mat * 4);
    if (nq > 0) {
        confusion_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, db->stream>>>(
            d_knn_idx, nq, k, d_q_labels, db->d_labels, db->t_total, reinterpret_cast<const int*>(base + off_map),
            n_classes, n_local, d_mat, d_bad);
        PE_LAUNCHED();
    }
    *d_mat_out = d_mat;
    return PE_OK;
}

/* counts (+ the bad flag behind them) -> row-normalised matrix (src/networkevaluator.cpp:242-244) */
int pose_confusion_finish(pe_db* db, const unsigned int* d_mat, int n_local, int nr_test_poses, float* out_host) {
    const size_t n_mat = (size_t)n_local * n_local;
    std::vector<unsigned int> h(n_mat + 1);
    PE_CUDA(cudaMemcpyAsync(h.data(), d_mat, (n_mat + 1) * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    PE_REQUIRE(h[n_mat] == 0, "class label outside class_to_local");
    for (size_t i = 0; i < n_mat; ++i) out_host[i] = (float)h[i] / (float)nr_test_poses;
    return PE_OK;
}

int pose_confusion_dev(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                       const int* class_to_local, int n_classes, int n_local, int nr_test_poses,
                       float* out_host) {
    unsigned int* d_mat = nullptr;
    PE_CHECK(pose_confusion_local(db, d_knn_idx, d_q_labels, nq, k, class_to_local, n_classes, n_local, &d_mat));
    return pose_confusion_finish(db, d_mat, n_local, nr_test_poses, out_host);
}

int pose_bootstrap_dev(pe_db* db, const int* d_knn3, int64_t nq, int nr_query_poses, int* d_out, int* d_len) {
    PE_REQUIRE(nr_query_poses > 0 && db->poses_per_object > 0, "pose counts not set");
    PE_CUDA(cudaSetDevice(db->device));
    if (nq == 0) return PE_OK;
    bootstrap_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, db->stream>>>(d_knn3, nq, nr_query_poses,
                                                                         db->poses_per_object, d_out, d_len);
    PE_LAUNCHED();
    return PE_OK;
}

int pose_accuracy_dev(pe_db* db, const int* d_max_sim0, const int* d_boot, int64_t nq, int nr_query_poses,
                      float* intra_pct, float* inter_pct) {
    PE_REQUIRE(nr_query_poses > 0, "nr_query_poses not set");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->scratch2.reserve(16));
    unsigned long long* d_c = db->scratch2.as<unsigned long long>();
    PE_CUDA(cudaMemsetAsync(d_c, 0, 16, db->stream));
    if (nq > 0) {
        accuracy_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, db->stream>>>(d_max_sim0, d_boot, nq, nr_query_poses, d_c);
        PE_LAUNCHED();
    }
    unsigned long long h[2];
    PE_CUDA(cudaMemcpyAsync(h, d_c, 16, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    *intra_pct = (float)h[0] / (float)nq * 100; /* src/networkevaluator.cpp:173-174 */
    *inter_pct = (float)h[1] / (float)nq * 100;
    return PE_OK;
}

}  // namespace pe

/* ------------------------------------------------------------------------------------------------
 * pe_histogram_dev_noalloc - the histogram reduction for callers that may not allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernels and arithmetic as pe_histogram_dev; the only difference is where the scratch lives:
 * a file-scope device array of fixed size instead of the DB's growable buffer.
 *
 * Every GPU that loads the library holds these kNaScratchBytes (260.2 KiB) until the process exits.  In HOST code the
 * array's name is only a stand-in for the symbol, never a device address: its address is resolved once per device with
 * cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
 * All calls on a device share the array and it keeps the last call's contents: the counters that must start at zero are
 * cleared on the stream at the start of every call, the rest is overwritten before it is read.  Calls that overlap on the
 * GPU (two DBs / two streams of one device) race on it: serialise them.
 * ------------------------------------------------------------------------------------------------ */
namespace pe {
namespace {

constexpr int kNaMaxQueries = 65536;    /* per call */
constexpr int kNaMaxClasses = 1024;     /* rot_inv entries */
/* [counts: 16 x u64][sum: f64][median: f32 + pad][rot_inv: kNaMaxClasses x i32][bins: 16 x f32][err: kNaMaxQueries x f32] */
constexpr size_t kNaOffSum = kMaxBins * 8, kNaOffMed = kNaOffSum + 8, kNaOffRot = kNaOffMed + 8;
constexpr size_t kNaOffBins = kNaOffRot + (size_t)kNaMaxClasses * 4, kNaOffErr = kNaOffBins + kMaxBins * 4;
constexpr size_t kNaScratchBytes = kNaOffErr + (size_t)kNaMaxQueries * 4;      /* 266,448 B */

static __device__ __align__(256) unsigned char g_hist_noalloc_scratch[kNaScratchBytes];

constexpr int kNaMaxDevices = 64;
std::mutex g_na_mutex;
unsigned char* g_na_ptr[kNaMaxDevices] = {};

/* device address of the array on `device` (which must be current); resolved on the first call per device */
int noalloc_scratch(int device, unsigned char** out) {
    PE_REQUIRE(device >= 0 && device < kNaMaxDevices, "device %d outside 0..%d", device, kNaMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[device]) {
        void* p = nullptr;
        PE_CUDA(cudaGetSymbolAddress(&p, g_hist_noalloc_scratch));
        PE_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address for the scratch array");
        g_na_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_na_ptr[device];
    return PE_OK;
}

}  // namespace
}  // namespace pe

extern "C" int pe_histogram_dev_noalloc(pe_db* db, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                                        const int* rot_inv, int n_rot, const float* bins, int n_bins, float* out) {
    using namespace pe;
    PE_REQUIRE(db && bins && out && (rot_inv || n_rot == 0), "null pointer");
    PE_REQUIRE(nq >= 0 && (nq == 0 || (d_knn_idx && d_q_labels)), "null pointer");
    PE_REQUIRE(nq <= kNaMaxQueries, "pe_histogram_dev_noalloc: %lld queries exceed the fixed scratch (at most %d per call; "
               "use pe_histogram_dev, which allocates)", (long long)nq, kNaMaxQueries);
    PE_REQUIRE(n_rot >= 0 && n_rot <= kNaMaxClasses, "pe_histogram_dev_noalloc: %d rotInv classes exceed the fixed scratch (at most %d; "
               "use pe_histogram_dev, which allocates)", n_rot, kNaMaxClasses);
    PE_REQUIRE(db->d_labels != nullptr, "DB was created without labels");
    PE_REQUIRE(n_bins >= 2 && n_bins <= kMaxBins, "n_bins=%d outside 2..%d", n_bins, kMaxBins);
    PE_REQUIRE(k >= 1 && k <= 32, "k=%d outside 1..32", k);
    PE_REQUIRE(db->poses_per_object > 0, "poses_per_object not set");
    PE_CUDA(cudaSetDevice(db->device));
    unsigned char* base = nullptr;
    PE_CHECK(noalloc_scratch(db->device, &base));
    unsigned long long* d_counts = reinterpret_cast<unsigned long long*>(base);
    double* d_sum = reinterpret_cast<double*>(base + kNaOffSum);
    float* d_med = reinterpret_cast<float*>(base + kNaOffMed);
    int* d_rot = reinterpret_cast<int*>(base + kNaOffRot);
    float* d_bins = reinterpret_cast<float*>(base + kNaOffBins);
    float* d_err = reinterpret_cast<float*>(base + kNaOffErr);
    /* the array keeps the previous call's contents: counts, sum and median start at zero, the tables are rewritten,
     * err[0, nq) is written by pose_err_kernel before median_select_kernel reads exactly that range */
    PE_CUDA(cudaMemsetAsync(base, 0, kNaOffRot, db->stream));
    if (n_rot > 0) PE_CUDA(cudaMemcpyAsync(d_rot, rot_inv, (size_t)n_rot * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CUDA(cudaMemcpyAsync(d_bins, bins, (size_t)n_bins * 4, cudaMemcpyHostToDevice, db->stream));
    if (nq > 0) {
        pose_err_kernel<<<(unsigned)((nq + kPoseWarps * kPoseQ - 1) / (kPoseWarps * kPoseQ)), kPoseWarps * 32, 0, db->stream>>>(
            d_knn_idx, nq, k, d_q_labels, db->d_labels, db->t_total, db->poses_per_object, d_rot, n_rot, d_bins, n_bins,
            d_err, d_counts, d_sum);
        PE_LAUNCHED();
    }
    median_select_kernel<<<1, 1024, 0, db->stream>>>(d_err, nq, d_med);
    PE_LAUNCHED();
    unsigned long long h_counts[kMaxBins];
    double h_sum = 0.0;
    float h_med = 0.f;
    PE_CUDA(cudaMemcpyAsync(h_counts, d_counts, sizeof(h_counts), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(&h_sum, d_sum, 8, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(&h_med, d_med, 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    /* normalisation of src/networkevaluator.cpp:342-346, as pose_hist_finish */
    const float total = (float)h_counts[0] + (float)h_counts[n_bins - 1];
    for (int b = 0; b < n_bins; ++b) out[b] = (float)h_counts[b] / total;
    out[n_bins] = (float)(h_sum / (double)nq);
    out[n_bins + 1] = h_med;
    return PE_OK;
}
