/*
 * miner.cu - the training-side consumer of the k-NN: most-similar-template table, bootstrapping
 * triplet selection and the Caffe-layout batch, with the template / training / hard-negative pools resident in HBM.
 *
 *   max_sim_kernel    datasetManager::computeMaxSimTmpl[Inplane] (reference src/datasetmanager.cpp:687-783): one CTA per
 *                     training sample scans its object's templates twice (best, then best of a different value);
 *                     ties keep the lower template index as the reference's strict compares do.
 *   elevation_kernel  acos(R(2,2)) per template, the quantity buildBatch compares for rotation-invariant objects
 *                     (src/networksolver.cpp:83,118,148), computed once.
 *   triplet_kernel    networkSolver::buildBatch (src/networksolver.cpp:52-186): one thread per triplet picks anchor,
 *                     puller, pusher0-2.  The reference draws from std::random_device; here every draw is a counter-
 *                     based hash of (seed, linearId, draw number) - the same generator oracle_miner.c restates, so
 *                     the selection is reproducible and checked bit for bit.
 *   gather_kernel     buildBatchQueue's batch_caffe fill (src/networksolver.cpp:32-40): HWC pool rows -> CHW batch.
 *                     HBM bound: 2 x 64 x 64 x C x 4 bytes per sample; one CTA moves 16 image rows through shared
 *                     memory (16-byte loads, pixel stride padded to an odd word count, coalesced plane stores).
 *   colorfill_kernel  bgfill::randomColorFill (src/bgfill.cpp:136-158) on the anchors for random_background = 1.
 *
 * pe_miner_compute_knn is networkSolver::computeKNN (src/networksolver.cpp:456-505) with everything left on the
 * device: forward over both pools, k = 3 matcher, bootstrap table.
 */
#include <cfloat>
#include <cmath>

#include "common.h"

struct pe_miner {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    int channels = 0, n_objects = 0, n_tmpl = 0, n_train = 0, inplane = 0;
    int64_t n_neg = 0;
    float *d_tmpl = nullptr, *d_train = nullptr, *d_neg = nullptr;       /* HWC pools */
    float *d_tmpl_labels = nullptr, *d_train_labels = nullptr, *d_elev = nullptr;
    std::vector<float> h_tmpl_labels;
    int* d_rot_inv = nullptr;
    long long* d_neg_off = nullptr;
    int* d_max_sim = nullptr;      /* [n_objects * n_train][2] */
    bool have_max_sim = false;
    int* d_knn = nullptr;          /* [n_objects * n_train][4] + [n_objects * n_train] lengths */
    bool bootstrapping = false;
    pe::DevBuf idx, flag, batch, desc_t, desc_q, knn_i, knn_d;
};

namespace pe {
namespace {

constexpr int kMaxDraws = 4096;
constexpr unsigned long long K1 = 0x9E3779B97F4A7C15ull, K2 = 0xD1B54A32D192ED03ull, K3 = 0x8CB92BA72F3D8DD7ull;

__host__ __device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ unsigned long long draw(unsigned long long key, unsigned long long j) { return mix64(key + j * K2); }
__device__ __forceinline__ int below(unsigned long long r, int n) { return (int)__umul64hi(r, (unsigned long long)n); }

__device__ __forceinline__ float acos_cr(float x) { return __double2float_rn(acos((double)x)); }
__device__ __forceinline__ float quat_angle(const float* __restrict__ a, const float* __restrict__ b) {
    float d = __fmul_rn(a[0], b[0]);
    d = __fadd_rn(d, __fmul_rn(a[1], b[1]));
    d = __fadd_rn(d, __fmul_rn(a[2], b[2]));
    d = __fadd_rn(d, __fmul_rn(a[3], b[3]));
    d = fabsf(d);
    d = d > 1.0f ? 1.0f : d;
    return __fmul_rn(2.0f, acos_cr(d));
}
__device__ __forceinline__ float trans_dot(const float* __restrict__ a, const float* __restrict__ b) {
    float d = __fmul_rn(a[5], b[5]);
    d = __fadd_rn(d, __fmul_rn(a[6], b[6]));
    return __fadd_rn(d, __fmul_rn(a[7], b[7]));
}

/* lexicographic (key, index) minimum across the CTA; every thread returns the winner */
__device__ void block_argmin(float& key, int& idx, float* s_k, int* s_i) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const float ok = __shfl_xor_sync(0xffffffffu, key, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oi >= 0 && (idx < 0 || ok < key || (ok == key && oi < idx))) { key = ok; idx = oi; }
    }
    __syncthreads();
    if (lane == 0) { s_k[warp] = key; s_i[warp] = idx; }
    __syncthreads();
    key = s_k[0]; idx = s_i[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
        const float ok = s_k[w];
        const int oi = s_i[w];
        if (oi >= 0 && (idx < 0 || ok < key || (ok == key && oi < idx))) { key = ok; idx = oi; }
    }
}

/* key to minimise: the angular distance (inplane) or minus the translation dot product; `ok` = passes the
 * reference's start value (FLT_MAX resp. numeric_limits<float>::min()) */
__global__ void __launch_bounds__(128) max_sim_kernel(const float* __restrict__ tmpl_labels, const float* __restrict__ train_labels,
                                                      int n_tmpl, int n_train, int inplane, int* __restrict__ out) {
    __shared__ float s_k[4];
    __shared__ int s_i[4];
    __shared__ float s_a[8];
    const long long q = blockIdx.x;
    const int object = (int)(q / n_train);
    if (threadIdx.x < 8) s_a[threadIdx.x] = train_labels[q * 8 + threadIdx.x];
    __syncthreads();
    const float* tl = tmpl_labels + (long long)object * n_tmpl * 8;
    auto sim = [&](int t, bool* ok) {
        if (inplane) {
            const float d = quat_angle(s_a + 1, tl + (long long)t * 8 + 1);
            *ok = d < FLT_MAX;
            return d;
        }
        const float d = trans_dot(s_a, tl + (long long)t * 8);
        *ok = d > FLT_MIN;
        return -d;
    };
    float bk = 0.f;
    int bi = -1;
    for (int t = threadIdx.x; t < n_tmpl; t += blockDim.x) {
        bool ok;
        const float k = sim(t, &ok);
        if (ok && (bi < 0 || k < bk)) { bk = k; bi = t; }
    }
    block_argmin(bk, bi, s_k, s_i);
    float ck = 0.f;
    int ci = -1;
    for (int t = threadIdx.x; t < n_tmpl; t += blockDim.x) {
        bool ok;
        const float k = sim(t, &ok);
        if (ok && !(bi >= 0 && k == bk) && (ci < 0 || k < ck)) { ck = k; ci = t; }
    }
    block_argmin(ck, ci, s_k, s_i);
    if (threadIdx.x == 0) { out[q * 2] = bi; out[q * 2 + 1] = ci; }
}

__global__ void elevation_kernel(const float* __restrict__ labels, long long n, float* __restrict__ elev) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* q = labels + i * 8 + 1;
    const float tx = __fmul_rn(2.0f, q[0]), ty = __fmul_rn(2.0f, q[1]);
    elev[i] = acos_cr(__fsub_rn(1.0f, __fadd_rn(__fmul_rn(tx, q[0]), __fmul_rn(ty, q[1]))));
}

struct Draws {
    unsigned long long key, j;
    __device__ unsigned long long next() { return draw(key, j++); }
};

__device__ int pick_elevation(Draws& d, const float* __restrict__ elev, int n_tmpl, int puller, int* flag) {
    const float ep = elev[puller];
    int p = below(d.next(), n_tmpl);
    int tries = 1;
    while ((double)fabsf(__fsub_rn(elev[p], ep)) < 0.2) {
        if (tries++ >= kMaxDraws) { *flag = 1; break; }
        p = below(d.next(), n_tmpl);
    }
    return p;
}

__global__ void triplet_kernel(const float* __restrict__ elev, const int* __restrict__ max_sim, const int* __restrict__ knn_table,
                               const int* __restrict__ knn_len, const int* __restrict__ rot_inv, const long long* __restrict__ neg_off,
                               int n_objects, int n_tmpl, int n_train, long long first_id, int n_triplets, unsigned long long seed,
                               int* __restrict__ out_idx, int* __restrict__ out_flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_triplets) return;
    const long long linearId = first_id + i;
    const int training_pose = (int)(linearId / n_objects), object = (int)(linearId % n_objects);
    Draws d{mix64(seed ^ ((unsigned long long)linearId * K1)), 0};
    int flag = 0;
    const long long q = (long long)object * n_train + training_pose;
    int puller = max_sim[q * 2];
    if (puller < 0) { puller = 0; flag = 2; }       /* the reference would index with an uninitialised int */
    const int tbase = object * n_tmpl;
    const float* el = elev + tbase;
    const bool rinv = rot_inv[object] != 0;
    int pusher0;
    if (rinv) {
        pusher0 = pick_elevation(d, el, n_tmpl, puller, &flag);
    } else {
        pusher0 = below(d.next(), n_tmpl);
        int tries = 1;
        while (pusher0 == puller) {
            if (tries++ >= kMaxDraws) { flag = 1; break; }
            pusher0 = below(d.next(), n_tmpl);
        }
    }
    int p0 = tbase + pusher0;
    if (neg_off != nullptr && neg_off[object + 1] > neg_off[object]) {
        const long long n = neg_off[object + 1] - neg_off[object];
        p0 = ~(int)(neg_off[object] + (long long)__umul64hi(d.next(), (unsigned long long)n));
    }
    int knn_object = -1, knn_pose = -1;
    bool missed = false;
    if (knn_table != nullptr) {
        knn_object = knn_table[q * 4];
        knn_pose = knn_table[q * 4 + 1];
        missed = knn_object != object || knn_pose != puller;
    }
    int p1, p2;
    if (missed) {
        p1 = knn_object * n_tmpl + knn_pose;
    } else {
        int pusher1;
        if (rinv) {
            pusher1 = pick_elevation(d, el, n_tmpl, puller, &flag);
        } else {
            pusher1 = below(d.next(), n_tmpl);
            int tries = 1;
            while (pusher1 == puller && pusher1 == pusher0) {
                if (tries++ >= kMaxDraws) { flag = 1; break; }
                pusher1 = below(d.next(), n_tmpl);
            }
        }
        p1 = tbase + pusher1;
    }
    if (knn_table != nullptr && knn_len[q] > 2 && missed) {
        p2 = knn_table[q * 4 + 2] * n_tmpl + knn_table[q * 4 + 3];
    } else {
        int pusher2 = below(d.next(), n_objects);
        int tries = 1;
        while (pusher2 == object) {
            if (tries++ >= kMaxDraws) { flag = 1; break; }
            pusher2 = below(d.next(), n_objects);
        }
        p2 = pusher2 * n_tmpl + below(d.next(), n_tmpl);
    }
    out_idx[i * 5 + 0] = (int)q;
    out_idx[i * 5 + 1] = tbase + puller;
    out_idx[i * 5 + 2] = p0;
    out_idx[i * 5 + 3] = p1;
    out_idx[i * 5 + 4] = p2;
    out_flag[i] = flag;
}

constexpr int kGatherRows = 16;       /* image rows per CTA */
constexpr int kGatherThreads = 256;

/* grid = (64 / kGatherRows, n_samples).  Sample s of the batch is slot s % 5 of triplet s / 5. */
template <int C>
__global__ void __launch_bounds__(kGatherThreads) gather_kernel(const float* __restrict__ tmpl_pool, const float* __restrict__ train_pool,
                                                               const float* __restrict__ neg_pool, const int* __restrict__ idx,
                                                               float* __restrict__ out) {
    constexpr int CP = C | 1;                          /* odd pixel stride: conflict-free plane reads */
    constexpr int PIX = kGatherRows * 64;
    __shared__ float s[PIX * CP];
    const int sample = blockIdx.y, slot = sample % 5;
    const int id = idx[sample];
    const size_t img = (size_t)64 * 64 * C;
    const float* src = slot == 0 ? train_pool + (size_t)id * img : (id < 0 ? neg_pool + (size_t)(~id) * img : tmpl_pool + (size_t)id * img);
    src += (size_t)blockIdx.x * PIX * C;
    const float4* src4 = reinterpret_cast<const float4*>(src);
#pragma unroll
    for (int v = threadIdx.x; v < PIX * C / 4; v += kGatherThreads) {
        const float4 x = __ldg(src4 + v);
        const float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int el = v * 4 + j, pix = el / C, c = el - pix * C;
            s[pix * CP + c] = e[j];
        }
    }
    __syncthreads();
    float* dst = out + (size_t)sample * img + (size_t)blockIdx.x * PIX;
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int p = threadIdx.x; p < PIX; p += kGatherThreads) __stcs(dst + (size_t)c * 4096 + p, s[p * CP + c]);
}

/* one CTA per anchor: mask = depth plane == 0 dilated by the 3x3 box cv::dilate(mask, mask, Mat()) uses (pixels
 * outside the image do not count), every masked pixel gets 7 uniform floats */
__global__ void __launch_bounds__(256) colorfill_kernel(float* __restrict__ batch, long long first_id, unsigned long long seed) {
    __shared__ unsigned char zero[64 * 64];
    float* img = batch + (size_t)blockIdx.x * 5 * 7 * 4096;
    for (int p = threadIdx.x; p < 4096; p += 256) zero[p] = img[3 * 4096 + p] == 0.f;
    __syncthreads();
    const unsigned long long key = mix64(seed ^ ((unsigned long long)(first_id + blockIdx.x) * K1) ^ K3);
    for (int p = threadIdx.x; p < 4096; p += 256) {
        const int h = p >> 6, w = p & 63;
        bool m = false;
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
                const int y = h + dy, x = w + dx;
                if (y >= 0 && y < 64 && x >= 0 && x < 64 && zero[y * 64 + x]) m = true;
            }
        if (!m) continue;
#pragma unroll
        for (int c = 0; c < 7; ++c)
            img[c * 4096 + p] = __fmul_rn((float)(draw(key, (unsigned long long)(p * 7 + c)) >> 40), 1.0f / 16777216.0f);
    }
}

int miner_alloc(void** p, size_t bytes) {
    PE_CUDA(cudaMalloc(p, bytes > 0 ? bytes : 16));
    return PE_OK;
}

int miner_max_sim(pe_miner* m) {
    if (m->have_max_sim) return PE_OK;
    const long long nq = (long long)m->n_objects * m->n_train;
    max_sim_kernel<<<(unsigned)nq, 128, 0, m->stream>>>(m->d_tmpl_labels, m->d_train_labels, m->n_tmpl, m->n_train, m->inplane, m->d_max_sim);
    PE_LAUNCHED();
    m->have_max_sim = true;
    return PE_OK;
}

int launch_gather(pe_miner* m, const int* d_idx, int n_samples, float* d_out) {
    const dim3 grid(64 / kGatherRows, (unsigned)n_samples);
#define PE_GATHER(C_)                                                                                                      \
    case C_:                                                                                                               \
        gather_kernel<C_><<<grid, kGatherThreads, 0, m->stream>>>(m->d_tmpl, m->d_train, m->d_neg, d_idx, d_out);          \
        break;
    switch (m->channels) {
        PE_GATHER(1) PE_GATHER(2) PE_GATHER(3) PE_GATHER(4) PE_GATHER(5) PE_GATHER(6) PE_GATHER(7) PE_GATHER(8)
        default: set_error("miner: %d channels", m->channels); return PE_ERR_INVALID;
    }
#undef PE_GATHER
    PE_LAUNCHED();
    return PE_OK;
}

}  // namespace
}  // namespace pe

using namespace pe;

extern "C" {

int pe_miner_create(int device, int channels, int n_objects, int nr_template_poses, int nr_training_poses,
                    const float* tmpl_patches_hwc, const float* tmpl_labels, const float* train_patches_hwc,
                    const float* train_labels, const int* rot_inv, int inplane, pe_miner** out) {
    PE_REQUIRE(out != nullptr, "null out");
    *out = nullptr;
    PE_REQUIRE(channels >= 1 && channels <= 8, "channels must be 1..8");
    PE_REQUIRE(n_objects >= 1 && nr_template_poses >= 1 && nr_training_poses >= 1, "empty sets");
    PE_REQUIRE((int64_t)n_objects * nr_template_poses < INT32_MAX && (int64_t)n_objects * nr_training_poses < INT32_MAX, "sets too large");
    PE_REQUIRE(tmpl_labels && train_labels && rot_inv, "null labels / rot_inv");
    PE_CUDA(cudaSetDevice(device));
    pe_miner* m = new pe_miner();
    m->device = device;
    m->channels = channels;
    m->n_objects = n_objects;
    m->n_tmpl = nr_template_poses;
    m->n_train = nr_training_poses;
    m->inplane = inplane != 0;
    *out = m;                                       /* so that a failure below can be cleaned up by the caller */
    const size_t img = (size_t)64 * 64 * channels * 4;
    const size_t nt = (size_t)n_objects * nr_template_poses, nq = (size_t)n_objects * nr_training_poses;
    auto fail = [&](int st) { pe_miner_destroy(m); *out = nullptr; return st; };
#define PE_TRY(expr)                                \
    do {                                            \
        const int st_ = (expr);                     \
        if (st_ != PE_OK) return fail(st_);         \
    } while (0)
#define PE_TRY_CUDA(expr)                                                                             \
    do {                                                                                              \
        const cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) {                                                                      \
            set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_));     \
            return fail(PE_ERR_CUDA);                                                                 \
        }                                                                                             \
    } while (0)
    PE_TRY_CUDA(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
    PE_TRY(miner_alloc((void**)&m->d_tmpl, nt * img));
    PE_TRY(miner_alloc((void**)&m->d_train, nq * img));
    PE_TRY(miner_alloc((void**)&m->d_tmpl_labels, nt * 32));
    PE_TRY(miner_alloc((void**)&m->d_train_labels, nq * 32));
    PE_TRY(miner_alloc((void**)&m->d_elev, nt * 4));
    PE_TRY(miner_alloc((void**)&m->d_rot_inv, (size_t)n_objects * 4));
    PE_TRY(miner_alloc((void**)&m->d_max_sim, nq * 8));
    PE_TRY(miner_alloc((void**)&m->d_knn, nq * 20));
    m->h_tmpl_labels.assign(tmpl_labels, tmpl_labels + nt * 8);
    if (tmpl_patches_hwc) PE_TRY_CUDA(cudaMemcpyAsync(m->d_tmpl, tmpl_patches_hwc, nt * img, cudaMemcpyHostToDevice, m->stream));
    if (train_patches_hwc) PE_TRY_CUDA(cudaMemcpyAsync(m->d_train, train_patches_hwc, nq * img, cudaMemcpyHostToDevice, m->stream));
    PE_TRY_CUDA(cudaMemcpyAsync(m->d_tmpl_labels, tmpl_labels, nt * 32, cudaMemcpyHostToDevice, m->stream));
    PE_TRY_CUDA(cudaMemcpyAsync(m->d_train_labels, train_labels, nq * 32, cudaMemcpyHostToDevice, m->stream));
    PE_TRY_CUDA(cudaMemcpyAsync(m->d_rot_inv, rot_inv, (size_t)n_objects * 4, cudaMemcpyHostToDevice, m->stream));
    elevation_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, m->stream>>>(m->d_tmpl_labels, (long long)nt, m->d_elev);
    count_launch();
    PE_TRY_CUDA(cudaGetLastError());
    PE_TRY_CUDA(cudaStreamSynchronize(m->stream));
#undef PE_TRY
#undef PE_TRY_CUDA
    return PE_OK;
}

void pe_miner_destroy(pe_miner* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    void* ptrs[] = {m->d_tmpl, m->d_train, m->d_neg, m->d_tmpl_labels, m->d_train_labels, m->d_elev, m->d_rot_inv, m->d_neg_off,
                    m->d_max_sim, m->d_knn};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    m->idx.release(); m->flag.release(); m->batch.release(); m->desc_t.release(); m->desc_q.release(); m->knn_i.release(); m->knn_d.release();
    if (m->stream && m->own_stream) cudaStreamDestroy(m->stream);
    delete m;
}

int pe_miner_pools_dev(pe_miner* m, float** d_tmpl_patches, float** d_train_patches) {
    PE_REQUIRE(m != nullptr, "null miner");
    if (d_tmpl_patches) *d_tmpl_patches = m->d_tmpl;
    if (d_train_patches) *d_train_patches = m->d_train;
    return PE_OK;
}

int pe_miner_share_stream(pe_miner* m, pe_net* net) {
    PE_REQUIRE(m != nullptr && net != nullptr, "null miner / net");
    PE_REQUIRE(net->device == m->device, "net and miner live on different devices");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaSetDevice(m->device));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    if (m->own_stream) PE_CUDA(cudaStreamDestroy(m->stream));
    m->stream = net->stream;
    m->own_stream = false;
    return PE_OK;
}

int pe_miner_sync(pe_miner* m) {
    PE_REQUIRE(m != nullptr, "null miner");
    PE_CUDA(cudaSetDevice(m->device));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    return PE_OK;
}

int pe_miner_set_hard_negatives(pe_miner* m, const float* patches_hwc, const int64_t* offsets) {
    PE_REQUIRE(m != nullptr, "null miner");
    PE_CUDA(cudaSetDevice(m->device));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    if (m->d_neg) { PE_CUDA(cudaFree(m->d_neg)); m->d_neg = nullptr; }
    if (m->d_neg_off) { PE_CUDA(cudaFree(m->d_neg_off)); m->d_neg_off = nullptr; }
    m->n_neg = 0;
    if (offsets == nullptr) return PE_OK;
    PE_REQUIRE(offsets[0] == 0, "offsets[0] must be 0");
    for (int o = 0; o < m->n_objects; ++o) PE_REQUIRE(offsets[o + 1] >= offsets[o], "offsets must not decrease");
    const int64_t n = offsets[m->n_objects];
    PE_REQUIRE(n < INT32_MAX, "too many hard negatives");
    PE_REQUIRE(n == 0 || patches_hwc != nullptr, "null patches");
    const size_t img = (size_t)64 * 64 * m->channels * 4;
    PE_CHECK(miner_alloc((void**)&m->d_neg, (size_t)n * img));
    PE_CHECK(miner_alloc((void**)&m->d_neg_off, (size_t)(m->n_objects + 1) * 8));
    if (n > 0) PE_CUDA(cudaMemcpyAsync(m->d_neg, patches_hwc, (size_t)n * img, cudaMemcpyHostToDevice, m->stream));
    PE_CUDA(cudaMemcpyAsync(m->d_neg_off, offsets, (size_t)(m->n_objects + 1) * 8, cudaMemcpyHostToDevice, m->stream));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    m->n_neg = n;
    return PE_OK;
}

int pe_miner_max_sim(pe_miner* m, int32_t* out) {
    PE_REQUIRE(m != nullptr, "null miner");
    PE_CUDA(cudaSetDevice(m->device));
    PE_CHECK(miner_max_sim(m));
    if (out) {
        PE_CUDA(cudaMemcpyAsync(out, m->d_max_sim, (size_t)m->n_objects * m->n_train * 8, cudaMemcpyDeviceToHost, m->stream));
        PE_CUDA(cudaStreamSynchronize(m->stream));
    }
    return PE_OK;
}

int pe_miner_set_knn(pe_miner* m, const int32_t* table, const int32_t* len) {
    PE_REQUIRE(m != nullptr, "null miner");
    PE_CUDA(cudaSetDevice(m->device));
    if (table == nullptr) { m->bootstrapping = false; return PE_OK; }
    PE_REQUIRE(len != nullptr, "null len");
    const size_t nq = (size_t)m->n_objects * m->n_train;
    PE_CUDA(cudaMemcpyAsync(m->d_knn, table, nq * 16, cudaMemcpyHostToDevice, m->stream));
    PE_CUDA(cudaMemcpyAsync(m->d_knn + nq * 4, len, nq * 4, cudaMemcpyHostToDevice, m->stream));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    m->bootstrapping = true;
    return PE_OK;
}

int pe_miner_compute_knn(pe_miner* m, pe_net* net, int32_t* out_table, int32_t* out_len) {
    PE_REQUIRE(m != nullptr && net != nullptr, "null miner / net");
    PE_REQUIRE(net->device == m->device, "net and miner live on different devices");
    int batch, channels, height, width, dim;
    PE_CHECK(pe_net_info(net, &batch, &channels, &height, &width, &dim));
    PE_REQUIRE(height == 64 && width == 64, "net input is not 64x64");
    PE_REQUIRE(channels == m->channels, "pools have %d channels, net input has %d", m->channels, channels);
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaSetDevice(m->device));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    const int64_t nt = (int64_t)m->n_objects * m->n_tmpl, nq = (int64_t)m->n_objects * m->n_train;
    PE_CHECK(m->desc_t.reserve((size_t)nt * dim * 4));
    PE_CHECK(m->desc_q.reserve((size_t)nq * dim * 4));
    PE_CHECK(m->knn_i.reserve((size_t)nq * 3 * 4));
    PE_CHECK(m->knn_d.reserve((size_t)nq * 3 * 4));
    PE_CHECK(net_forward_dev(net, m->d_tmpl, nt, m->channels, m->desc_t.as<float>(), dim));
    PE_CHECK(net_forward_dev(net, m->d_train, nq, m->channels, m->desc_q.as<float>(), dim));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    PE_CHECK(forward_tc_check(net));
    pe_db* db = nullptr;
    PE_CHECK(pe_db_create_dev(m->desc_t.as<float>(), nt, dim, 0, m->h_tmpl_labels.data(), nt, m->n_tmpl, m->device, &db));
    int st = knn_l2_local(db, m->desc_q.as<float>(), nq, 3, m->knn_i.as<int32_t>(), m->knn_d.as<float>());
    if (st == PE_OK) st = pose_bootstrap_dev(db, m->knn_i.as<int>(), nq, m->n_train, m->d_knn, m->d_knn + nq * 4);
    if (st == PE_OK && cudaStreamSynchronize(db->stream) != cudaSuccess) { set_error("miner: k-NN pass failed"); st = PE_ERR_CUDA; }
    if (st == PE_OK) st = knn_tc_check(db);
    pe_db_destroy(db);
    PE_CHECK(st);
    m->bootstrapping = true;
    if (out_table) PE_CUDA(cudaMemcpy(out_table, m->d_knn, (size_t)nq * 16, cudaMemcpyDeviceToHost));
    if (out_len) PE_CUDA(cudaMemcpy(out_len, m->d_knn + nq * 4, (size_t)nq * 4, cudaMemcpyDeviceToHost));
    return PE_OK;
}

int pe_miner_build_batch_dev(pe_miner* m, int batch_size, int triplet_size, int iter, uint64_t seed, int random_background,
                             float* d_out_batch, int32_t* d_out_idx) {
    PE_REQUIRE(m != nullptr, "null miner");
    PE_REQUIRE(triplet_size == 5, "buildBatch stores five samples per triplet (triplet_size must be 5)");
    PE_REQUIRE(batch_size >= 0 && iter >= 0, "bad batch_size / iter");
    PE_REQUIRE(random_background == 0 || random_background == 1,
               "random_background: only 0 (off) and 1 (colour) are built; shapes / fractals / real images are bgfill's");
    PE_REQUIRE(random_background == 0 || m->channels == 7, "randomColorFill writes seven channels per pixel");
    const int n_triplets = batch_size / triplet_size;
    if (n_triplets == 0) return PE_OK;
    PE_REQUIRE(m->n_objects >= 2, "pusher2 needs a second object");
    const long long first_id = (long long)iter * batch_size / triplet_size;
    PE_REQUIRE(first_id + n_triplets <= (long long)m->n_objects * m->n_train, "iter runs past the training set");
    PE_CUDA(cudaSetDevice(m->device));
    PE_CHECK(miner_max_sim(m));
    PE_CHECK(m->idx.reserve((size_t)n_triplets * 5 * 4));
    PE_CHECK(m->flag.reserve((size_t)n_triplets * 4));
    int* d_idx = d_out_idx ? d_out_idx : m->idx.as<int>();
    const size_t nq = (size_t)m->n_objects * m->n_train;
    triplet_kernel<<<(n_triplets + 127) / 128, 128, 0, m->stream>>>(
        m->d_elev, m->d_max_sim, m->bootstrapping ? m->d_knn : nullptr, m->bootstrapping ? m->d_knn + nq * 4 : nullptr, m->d_rot_inv,
        m->n_neg > 0 ? m->d_neg_off : nullptr, m->n_objects, m->n_tmpl, m->n_train, first_id, n_triplets, seed, d_idx, m->flag.as<int>());
    PE_LAUNCHED();
    if (d_out_batch) {
        PE_CHECK(launch_gather(m, d_idx, n_triplets * 5, d_out_batch));
        if (random_background == 1) {
            colorfill_kernel<<<n_triplets, 256, 0, m->stream>>>(d_out_batch, first_id, seed);
            PE_LAUNCHED();
        }
    }
    return PE_OK;
}

int pe_miner_flags(pe_miner* m, int n_triplets, int32_t* out_flags) {
    PE_REQUIRE(m != nullptr && out_flags != nullptr && n_triplets >= 0, "bad arguments");
    PE_REQUIRE((size_t)n_triplets * 4 <= m->flag.cap, "no batch of that size has been built");
    PE_CUDA(cudaSetDevice(m->device));
    PE_CUDA(cudaMemcpyAsync(out_flags, m->flag.p, (size_t)n_triplets * 4, cudaMemcpyDeviceToHost, m->stream));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    return PE_OK;
}

int pe_miner_build_batch(pe_miner* m, int batch_size, int triplet_size, int iter, uint64_t seed, int random_background,
                         float* out_batch, int32_t* out_idx) {
    PE_REQUIRE(m != nullptr, "null miner");
    const int n_triplets = triplet_size > 0 ? batch_size / triplet_size : 0;
    const size_t bytes = (size_t)n_triplets * 5 * 64 * 64 * m->channels * 4;
    PE_CUDA(cudaSetDevice(m->device));
    if (out_batch) PE_CHECK(m->batch.reserve(bytes));
    PE_CHECK(pe_miner_build_batch_dev(m, batch_size, triplet_size, iter, seed, random_background,
                                      out_batch ? m->batch.as<float>() : nullptr, nullptr));
    if (n_triplets == 0) return PE_OK;
    if (out_batch) PE_CUDA(cudaMemcpyAsync(out_batch, m->batch.p, bytes, cudaMemcpyDeviceToHost, m->stream));
    if (out_idx) PE_CUDA(cudaMemcpyAsync(out_idx, m->idx.p, (size_t)n_triplets * 20, cudaMemcpyDeviceToHost, m->stream));
    std::vector<int32_t> flags(n_triplets);
    PE_CUDA(cudaMemcpyAsync(flags.data(), m->flag.p, (size_t)n_triplets * 4, cudaMemcpyDeviceToHost, m->stream));
    PE_CUDA(cudaStreamSynchronize(m->stream));
    for (int i = 0; i < n_triplets; ++i)
        if (flags[i]) {
            set_error(flags[i] == 2 ? "buildBatch: triplet %d has no most-similar template (maxSimTmpl entry unset)"
                                    : "buildBatch: triplet %d found no admissible pusher in %d draws (the reference would not return)",
                      i, kMaxDraws);
            return PE_ERR_INVALID;
        }
    return PE_OK;
}

}  // extern "C"
