/*
 * knn.cu - brute-force k-nearest-neighbour matching of query descriptors against the
 * template-descriptor DB on sm_100a.
 *
 * Replaces `DescriptorMatcher::create("BruteForce")` + `add(DBfeats)` + `knnMatch(DBtest, matches, k)`
 * (reference src/networkevaluator.cpp:193-197, 263-267; src/networksolver.cpp:466-472), i.e.
 * OpenCV BFMatcher(NORM_L2): dist = sqrt(sum (q-t)^2), k smallest ascending, ties -> lower index.
 *
 * L2 path, two kernels
 *   knn_l2_scan     grid (query block, DB chunk).  A thread owns RQ queries (descriptors held in
 *                   registers, pre-scaled by -2); the CTA streams its DB chunk through a
 *                   cp.async double-buffered shared-memory tile and every template row is read
 *                   by a warp-broadcast LDS.  Ranking value |t|^2 - 2 q.t (one FFMA per
 *                   dimension); per-query sorted candidate lists of k+2 entries live in shared
 *                   memory, only the admission threshold is in a register.
 *   knn_select      one warp per query: recomputes the reference's exact fp32 expression
 *                   sqrt(sum (q-t)^2) for every candidate of every chunk and extracts the k
 *                   smallest by (distance, index).  Also merges per-shard lists after the NCCL
 *                   all-gather (distances taken as given).
 * Hamming path (new behaviour): same scan with XOR + POPC on packed codes; exact.
 */
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>

#include "common.h"
#include "knn_exact.cuh"

namespace pe {

namespace {

constexpr int kScanThreads = 128;
constexpr int kTile = 256; /* templates per shared-memory tile */

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc));
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

/* sorted insert into the calling thread's list j (entries strided by RQ*threads in shared memory);
 * strict '>' keeps the earlier (lower-index) entry first among equals.  Returns the new threshold. */
template <typename T, int RQ>
__device__ __noinline__ T list_insert(T* ld, int* li, int kc, int j, T d, int idx) {
    const int stride = RQ * kScanThreads;
    const int base = j * kScanThreads + threadIdx.x;
    int p = kc - 1;
    while (p > 0 && ld[(p - 1) * stride + base] > d) {
        ld[p * stride + base] = ld[(p - 1) * stride + base];
        li[p * stride + base] = li[(p - 1) * stride + base];
        --p;
    }
    ld[p * stride + base] = d;
    li[p * stride + base] = idx;
    return ld[(kc - 1) * stride + base];
}

/* ---------------------------------------------------------------- L2 scan ---- */
/* templates per shared-memory tile of the L2 scan: 256 rows up to 16 dims, fewer for wide rows so that the
 * double-buffered tile stays well under the 227 KB limit (dim 128: 2 x 64 x 129 x 4 B = 66 KB) */
__host__ __device__ constexpr int l2_tile_rows(int dp) { return dp <= 16 ? 256 : (dp <= 64 ? 128 : 64); }

template <int DP, int RQ>
__global__ void __launch_bounds__(kScanThreads)
knn_l2_scan_kernel(const float* __restrict__ q, int nq, const float* __restrict__ db,
                   const float* __restrict__ norm, const float* __restrict__ ctr, long long t_local, long long chunk_len,
                   int n_chunks, int kc, long long row_offset, float* __restrict__ cand_d,
                   int* __restrict__ cand_i) {
    constexpr int kT = l2_tile_rows(DP);
    extern __shared__ __align__(16) float smem[];
    float* tile = smem;                                  /* [2][kT][DP] */
    float* tnorm = tile + 2 * kT * DP;                /* [2][kT] */
    float* ld = tnorm + 2 * kT;                       /* [kc][RQ][threads] */
    int* li = reinterpret_cast<int*>(ld + kc * RQ * kScanThreads);
    const int tid = threadIdx.x;
    const int chunk = blockIdx.y;
    const long long t_begin = chunk * chunk_len;
    long long t_end = t_begin + chunk_len;
    if (t_end > t_local) t_end = t_local;

    float qr[RQ][DP];
    float thr[RQ];
    int qi[RQ];
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        qi[j] = (blockIdx.x * RQ + j) * kScanThreads + tid;
        const bool ok = qi[j] < nq;
        thr[j] = ok ? INFINITY : -INFINITY; /* idle slots never admit a candidate */
#pragma unroll
        for (int d4 = 0; d4 < DP / 4; ++d4) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok) {      /* centred on the DB mean, like the rows and their norms (L2 is translation invariant) */
                const float4 c = __ldg(reinterpret_cast<const float4*>(ctr) + d4);
                v = __ldg(reinterpret_cast<const float4*>(q + (size_t)qi[j] * DP) + d4);
                v.x -= c.x; v.y -= c.y; v.z -= c.z; v.w -= c.w;
            }
            qr[j][4 * d4 + 0] = -2.f * v.x; qr[j][4 * d4 + 1] = -2.f * v.y;
            qr[j][4 * d4 + 2] = -2.f * v.z; qr[j][4 * d4 + 3] = -2.f * v.w;
        }
        for (int p = 0; p < kc; ++p) {
            ld[(p * RQ + j) * kScanThreads + tid] = INFINITY;
            li[(p * RQ + j) * kScanThreads + tid] = -1;
        }
    }

    const int n_tiles = (int)((t_end - t_begin + kT - 1) / kT);
    auto issue = [&](int it) {
        const long long t0 = t_begin + (long long)it * kT;
        const int valid = (int)((t_end - t0) < kT ? (t_end - t0) : kT);
        float* dst = tile + (it & 1) * kT * DP;
        const float* src = db + (size_t)t0 * DP;
        for (int i = tid; i < valid * (DP / 4); i += kScanThreads) cp_async16(dst + 4 * i, src + 4 * i);
        float* ndst = tnorm + (it & 1) * kT;
        for (int i = tid; i < valid; i += kScanThreads) cp_async4(ndst + i, norm + t0 + i);
    };
    if (n_tiles > 0) issue(0);
    cp_async_commit();
    for (int it = 0; it < n_tiles; ++it) {
        if (it + 1 < n_tiles) issue(it + 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const long long t0 = t_begin + (long long)it * kT;
        const int valid = (int)((t_end - t0) < kT ? (t_end - t0) : kT);
        float* tl = tile + (it & 1) * kT * DP;
        const float* tn = tnorm + (it & 1) * kT;
        /* centre the landed rows in place: valid x DP subtractions per tile against RQ x valid x DP FMAs per thread */
        for (int i = tid; i < valid * (DP / 4); i += kScanThreads) {
            float4 v = reinterpret_cast<float4*>(tl)[i];
            const float4 c = __ldg(reinterpret_cast<const float4*>(ctr) + (i % (DP / 4)));
            v.x -= c.x; v.y -= c.y; v.z -= c.z; v.w -= c.w;
            reinterpret_cast<float4*>(tl)[i] = v;
        }
        __syncthreads();
#pragma unroll 2
        for (int t = 0; t < valid; ++t) {
            float tv[DP];
#pragma unroll
            for (int d4 = 0; d4 < DP / 4; ++d4) {
                const float4 v = *reinterpret_cast<const float4*>(tl + t * DP + 4 * d4);
                tv[4 * d4 + 0] = v.x; tv[4 * d4 + 1] = v.y; tv[4 * d4 + 2] = v.z; tv[4 * d4 + 3] = v.w;
            }
            const float n2 = tn[t];
            float s[RQ];
            bool any = false;
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                float a = n2;
#pragma unroll
                for (int d = 0; d < DP; ++d) a = fmaf(qr[j][d], tv[d], a);
                s[j] = a;
                any |= a < thr[j];
            }
            if (any) {
#pragma unroll
                for (int j = 0; j < RQ; ++j)
                    if (s[j] < thr[j]) thr[j] = list_insert<float, RQ>(ld, li, kc, j, s[j], (int)(t0 + t + row_offset));
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        if (qi[j] >= nq) continue;
        const size_t o = ((size_t)qi[j] * n_chunks + chunk) * kc;
        for (int p = 0; p < kc; ++p) {
            cand_d[o + p] = ld[(p * RQ + j) * kScanThreads + tid];
            cand_i[o + p] = li[(p * RQ + j) * kScanThreads + tid];
        }
    }
}

/* ----------------------------------------------------------- Hamming scan ---- */
template <int W, int RQ>
__global__ void __launch_bounds__(kScanThreads)
knn_hamming_scan_kernel(const uint32_t* __restrict__ q, int nq, const uint32_t* __restrict__ db,
                        long long t_local, long long chunk_len, int n_chunks, int kc,
                        long long row_offset, int* __restrict__ cand_d, int* __restrict__ cand_i) {
    extern __shared__ __align__(16) float smem[];
    uint32_t* tile = reinterpret_cast<uint32_t*>(smem);          /* [2][kTile][W] */
    int* ld = reinterpret_cast<int*>(tile + 2 * kTile * W);      /* [kc][RQ][threads] */
    int* li = ld + kc * RQ * kScanThreads;
    const int tid = threadIdx.x;
    const int chunk = blockIdx.y;
    const long long t_begin = chunk * chunk_len;
    long long t_end = t_begin + chunk_len;
    if (t_end > t_local) t_end = t_local;

    uint32_t qr[RQ][W];
    int thr[RQ], qi[RQ];
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        qi[j] = (blockIdx.x * RQ + j) * kScanThreads + tid;
        const bool ok = qi[j] < nq;
        thr[j] = ok ? INT_MAX : INT_MIN;
#pragma unroll
        for (int w = 0; w < W; ++w) qr[j][w] = ok ? __ldg(q + (size_t)qi[j] * W + w) : 0u;
        for (int p = 0; p < kc; ++p) {
            ld[(p * RQ + j) * kScanThreads + tid] = INT_MAX;
            li[(p * RQ + j) * kScanThreads + tid] = -1;
        }
    }
    const int n_tiles = (int)((t_end - t_begin + kTile - 1) / kTile);
    auto issue = [&](int it) {
        const long long t0 = t_begin + (long long)it * kTile;
        const int valid = (int)((t_end - t0) < kTile ? (t_end - t0) : kTile);
        uint32_t* dst = tile + (it & 1) * kTile * W;
        const uint32_t* src = db + (size_t)t0 * W;
        for (int i = tid; i < valid * W; i += kScanThreads) cp_async4(dst + i, src + i);
    };
    if (n_tiles > 0) issue(0);
    cp_async_commit();
    for (int it = 0; it < n_tiles; ++it) {
        if (it + 1 < n_tiles) issue(it + 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const long long t0 = t_begin + (long long)it * kTile;
        const int valid = (int)((t_end - t0) < kTile ? (t_end - t0) : kTile);
        const uint32_t* tl = tile + (it & 1) * kTile * W;
#pragma unroll 2
        for (int t = 0; t < valid; ++t) {
            uint32_t tv[W];
#pragma unroll
            for (int w = 0; w < W; ++w) tv[w] = tl[t * W + w];
            int s[RQ];
            bool any = false;
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                int a = 0;
#pragma unroll
                for (int w = 0; w < W; ++w) a += __popc(qr[j][w] ^ tv[w]);
                s[j] = a;
                any |= a < thr[j];
            }
            if (any) {
#pragma unroll
                for (int j = 0; j < RQ; ++j)
                    if (s[j] < thr[j]) thr[j] = list_insert<int, RQ>(ld, li, kc, j, s[j], (int)(t0 + t + row_offset));
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        if (qi[j] >= nq) continue;
        const size_t o = ((size_t)qi[j] * n_chunks + chunk) * kc;
        for (int p = 0; p < kc; ++p) {
            cand_d[o + p] = ld[(p * RQ + j) * kScanThreads + tid];
            cand_i[o + p] = li[(p * RQ + j) * kScanThreads + tid];
        }
    }
}

/* ------------------------------------------------------------------ select ---- */
template <typename T>
__device__ __forceinline__ T big();
template <>
__device__ __forceinline__ float big<float>() { return INFINITY; }
template <>
__device__ __forceinline__ int big<int>() { return INT_MAX; }

/* candidate (q, l, p) lives at q*stride_q + l*stride_l + p, l < n_lists, p < kc.
 * REFINE: recompute the distance of each candidate from the DB rows (L2 only). */
template <typename T, bool REFINE>
__global__ void __launch_bounds__(32)
knn_select_kernel(const T* __restrict__ cand_d, const int* __restrict__ cand_i, long long stride_q,
                  long long stride_l, int n_lists, int kc, const float* __restrict__ q,
                  const float* __restrict__ db, int dp, long long row_offset, int k,
                  int* __restrict__ out_idx, T* __restrict__ out_dist,
                  const float* __restrict__ ctr = nullptr, float max_norm2 = 0.f, int* __restrict__ flag = nullptr) {
    extern __shared__ __align__(16) float smem[];
    const int n_cand = n_lists * kc;
    float* sq = smem; /* REFINE: the query row (dp floats, keeps the lists 16-byte aligned) */
    T* sd = reinterpret_cast<T*>(smem + (REFINE ? dp : 0));
    int* si = reinterpret_cast<int*>(sd) + n_cand;
    const int lane = threadIdx.x;
    const long long qi = blockIdx.x;
    float qn2 = 0.f, b_key = INFINITY;          /* certificate (REFINE && flag): |q - c|^2, smallest key a list dropped */
    if (REFINE) {
        for (int d = lane; d < dp; d += 32) {
            const float v = q[(size_t)qi * dp + d];
            sq[d] = v;
            if (flag) { const float c = v - ctr[d]; qn2 = fmaf(c, c, qn2); }
        }
        __syncwarp();
    }
    for (int c = lane; c < n_cand; c += 32) {
        const int l = c / kc, p = c - l * kc;
        const size_t at = (size_t)qi * stride_q + (size_t)l * stride_l + p;
        const int idx = cand_i[at];
        T d = big<T>();
        if (idx >= 0) {
            if (REFINE) {
                d = (T)exact_l2(sq, db + (size_t)(idx - row_offset) * dp, dp);
                /* a full list (sorted ascending) dropped only templates whose key was >= its last entry */
                if (flag && p == kc - 1) b_key = fminf(b_key, (float)cand_d[at]);
            } else d = cand_d[at];
        }
        sd[c] = d;
        si[c] = idx >= 0 ? idx : INT_MAX;
    }
    __syncwarp();
    T d_kth = big<T>();
    bool have_k = true;
    for (int r = 0; r < k; ++r) {
        T bd = big<T>();
        int bi = INT_MAX, bc = -1;
        for (int c = lane; c < n_cand; c += 32) {
            const T d = sd[c];
            const int i = si[c];
            if (d < bd || (d == bd && i < bi)) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bi != INT_MAX);   /* an entry with id INT_MAX is an empty slot */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = w < 0 ? INT_MAX : __shfl_sync(0xffffffffu, bi, w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        if (bi == INT_MAX) have_k = false; else d_kth = bd;
        if (lane == 0) {
            out_idx[(size_t)qi * k + r] = bi == INT_MAX ? -1 : bi;
            out_dist[(size_t)qi * k + r] = bi == INT_MAX ? big<T>() : bd;
            if (bc >= 0) { sd[bc] = big<T>(); si[bc] = INT_MAX; }
        }
        __syncwarp();
    }
    if (REFINE && flag) {
        /* Certificate of the SIMT scan (same argument as knn_tc_select_kernel): its keys are |t'|^2 - 2 q'.t' evaluated
         * as one fp32 FMA chain of dp + 1 terms, error <= (dp + 2) 2^-23 (T^2 + 2 |q'| T); every template a chunk's list
         * dropped has a key >= that list's last entry, so its d^2 >= |q'|^2 + min over lists - eps */
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            qn2 += __shfl_xor_sync(0xffffffffu, qn2, off);
            b_key = fminf(b_key, __shfl_xor_sync(0xffffffffu, b_key, off));
        }
        bool proven = true;
        if (b_key < INFINITY) {
            const float qn = sqrtf(qn2), tmax = sqrtf(max_norm2);
            const float eps = (float)(dp + 2) * 1.1920929e-7f * (max_norm2 + 2.f * qn * tmax) + 1.9073486328125e-6f * qn2;
            const float dk = (float)d_kth;
            proven = have_k && dk * dk * 1.000002f < qn2 + b_key - eps;
        }
        if (!proven && lane == 0) {
            const int slot = atomicAdd(flag, 1);
            flag[4 + slot] = (int)qi;
        }
    }
}

__global__ void pad_rows_kernel(const float* __restrict__ src, int dim, float* __restrict__ dst, int dp, long long n) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n * dp) return;
    const long long r = i / dp;
    const int c = (int)(i - r * dp);
    dst[i] = c < dim ? src[r * dim + c] : 0.f;
}

__global__ void row_norm_kernel(const float* __restrict__ x, int dp, float* __restrict__ norm, long long n) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= n) return;
    float s = 0.f;
    for (int d = 0; d < dp; ++d) { const float v = x[r * dp + d]; s = fmaf(v, v, s); }
    norm[r] = s;
}

/* ------------------------------------------------- DB mean / centred statistics ---- */
/* column sums of the padded DB in double (fp32 partials flushed every 64 rows); one launch, dp <= 128 columns:
 * thread (row lane, column) with blockDim = (dp, 1024 / dp rows) */
__global__ void center_sum_kernel(const float* __restrict__ desc, int dp, long long t_local, double* __restrict__ sum) {
    const int col = threadIdx.x % dp, lane_r = threadIdx.x / dp, rows_per_block = blockDim.x / dp;
    double tot = 0.0;
    float acc = 0.f;
    int n = 0;
    for (long long r = (long long)blockIdx.x * rows_per_block + lane_r; r < t_local; r += (long long)gridDim.x * rows_per_block) {
        acc += desc[r * dp + col];
        if (++n == 64) { tot += acc; acc = 0.f; n = 0; }
    }
    tot += acc;
    atomicAdd(sum + col, tot);
}

/* centred squared norms + out[0] = max |t-c| over elements, out[1] = max |t-c|^2 over rows (bit patterns of floats >= 0) */
__global__ void center_stats_kernel(const float* __restrict__ desc, int dp, long long t_local, const float* __restrict__ ctr,
                                    float* __restrict__ norm_c, unsigned* __restrict__ out) {
    float m = 0.f, mn = 0.f;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < t_local; r += (long long)gridDim.x * blockDim.x) {
        float n = 0.f;
        for (int d = 0; d < dp; ++d) {
            const float v = desc[r * dp + d] - ctr[d];
            m = fmaxf(m, fabsf(v));
            n = fmaf(v, v, n);
        }
        norm_c[r] = n;
        mn = fmaxf(mn, n);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        mn = fmaxf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (isfinite(m)) atomicMax(out, __float_as_uint(m));
        if (isfinite(mn)) atomicMax(out + 1, __float_as_uint(mn));
    }
}

/* --------------------------------------------- exact re-scan of unproven queries ---- */
/* The matcher's last resort and its ground truth: for every flagged query the reference's own expression
 * sqrt(sum (q-t)^2) over ALL rows of the shard (no ranking shortcut), k smallest by (distance, row).
 * grid (slots, chunks): CTA (x, y) serves flagged queries x, x + gridDim.x, ... on row chunk y (up to 64 chunks, so that a
 * single flagged query - the common case - is spread over 64 SMs: ~10 us against 1 M rows).  A thread keeps a sorted
 * top-KT of its rows in registers (rows ascend per thread: a strict '<' keeps the lower row among equals), the CTA merges
 * them in shared memory and writes its chunk's k best; the last chunk to arrive (counter per query) merges the chunks. */
constexpr int kFbThreads = 256, kFbMaxChunks = 64, kFbMaxK = 16;

__device__ void fb_block_topk(float* s_d, int* s_i, int n, int k, float* out_d, int* out_i) {
    /* warp 0: k rounds of (distance, row) argmin over n shared-memory entries */
    const int lane = threadIdx.x & 31;
    if (threadIdx.x >= 32) return;
    for (int r = 0; r < k; ++r) {
        float bd = INFINITY;
        int bi = INT_MAX, bc = -1;
        for (int c = lane; c < n; c += 32) {
            const float d = s_d[c];
            const int i = s_i[c];
            if (i != INT_MAX && (bc < 0 || d < bd || (d == bd && i < bi))) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bc >= 0);     /* lane holding the smallest (key, id), -1: none */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = __shfl_sync(0xffffffffu, bi, w < 0 ? 0 : w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        if (lane == 0) {
            out_d[r] = bc >= 0 ? bd : INFINITY;
            out_i[r] = bc >= 0 ? bi : INT_MAX;
            if (bc >= 0) s_i[bc] = INT_MAX;
        }
        __syncwarp();
    }
}

template <int KT>
__global__ void __launch_bounds__(kFbThreads)
knn_exact_fallback_kernel(const int* __restrict__ flag, const float* __restrict__ q, int dp, const float* __restrict__ db,
                          long long t_local, long long row_offset, int k, int n_chunks, float* __restrict__ part_d,
                          int* __restrict__ part_i, unsigned* __restrict__ done, int* __restrict__ out_idx,
                          float* __restrict__ out_dist) {
    extern __shared__ __align__(16) float fb_smem[];
    float* sq = fb_smem;                                        /* [dp] */
    float* s_d = fb_smem + dp;                                  /* [max(kFbThreads * KT, n_chunks * k)] */
    int* s_i = reinterpret_cast<int*>(s_d + (kFbThreads * KT > kFbMaxChunks * kFbMaxK ? kFbThreads * KT : kFbMaxChunks * kFbMaxK));
    __shared__ float r_d[kFbMaxK];
    __shared__ int r_i[kFbMaxK];
    __shared__ bool s_last;
    const int count = flag[0];
    if (count == 0) return;
    const int chunk = blockIdx.y;
    const long long rows_per = (t_local + n_chunks - 1) / n_chunks;
    const long long r0 = chunk * rows_per, r1 = (r0 + rows_per < t_local) ? r0 + rows_per : t_local;
    for (int f = blockIdx.x; f < count; f += gridDim.x) {
        const int qi = flag[4 + f];
        __syncthreads();
        for (int d = threadIdx.x; d < dp; d += kFbThreads) sq[d] = q[(size_t)qi * dp + d];
        __syncthreads();
        float bd[KT];
        int bi[KT];
#pragma unroll
        for (int p = 0; p < KT; ++p) { bd[p] = INFINITY; bi[p] = INT_MAX; }
        /* four rows in flight per thread: the loop is bound by the latency of the row reads */
        for (long long r = r0 + threadIdx.x; r < r1; r += 4 * kFbThreads) {
            float dj[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const long long rr = r + (long long)j * kFbThreads;
                dj[j] = rr < r1 ? exact_l2(sq, db + (size_t)rr * dp, dp) : INFINITY;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float d = dj[j];
                if (d < bd[KT - 1]) {       /* into the last slot, then bubble towards the front (all indices compile-time) */
                    bd[KT - 1] = d;
                    bi[KT - 1] = (int)(r + (long long)j * kFbThreads);
#pragma unroll
                    for (int p = KT - 1; p > 0; --p) {
                        const bool sw = bd[p] < bd[p - 1];
                        const float td = bd[p - 1];
                        const int ti = bi[p - 1];
                        bd[p - 1] = sw ? bd[p] : td;
                        bi[p - 1] = sw ? bi[p] : ti;
                        bd[p] = sw ? td : bd[p];
                        bi[p] = sw ? ti : bi[p];
                    }
                }
            }
        }
#pragma unroll
        for (int p = 0; p < KT; ++p) { s_d[threadIdx.x * KT + p] = bd[p]; s_i[threadIdx.x * KT + p] = bi[p]; }
        __syncthreads();
        fb_block_topk(s_d, s_i, kFbThreads * KT, k, r_d, r_i);
        __syncthreads();
        if (threadIdx.x < k) {
            part_d[((size_t)f * n_chunks + chunk) * k + threadIdx.x] = r_d[threadIdx.x];
            part_i[((size_t)f * n_chunks + chunk) * k + threadIdx.x] = r_i[threadIdx.x];
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(done + f, 1u) == (unsigned)n_chunks - 1;
        __syncthreads();
        if (s_last) {
            __threadfence();
            for (int c = threadIdx.x; c < n_chunks * k; c += kFbThreads) {
                s_d[c] = __ldcg(part_d + (size_t)f * n_chunks * k + c);
                s_i[c] = __ldcg(part_i + (size_t)f * n_chunks * k + c);
            }
            __syncthreads();
            fb_block_topk(s_d, s_i, n_chunks * k, k, r_d, r_i);
            __syncthreads();
            if (threadIdx.x < k) {
                const bool ok = r_i[threadIdx.x] != INT_MAX;
                out_idx[(size_t)qi * k + threadIdx.x] = ok ? (int)(r_i[threadIdx.x] + row_offset) : -1;
                out_dist[(size_t)qi * k + threadIdx.x] = ok ? r_d[threadIdx.x] : INFINITY;
            }
            if (threadIdx.x == 0) done[f] = 0;                  /* ready for the next call */
        }
    }
}

int pad_dim(int dim) {
    if (dim <= 4) return 4;
    if (dim <= 16) return 16;
    if (dim <= 64) return 64;
    if (dim <= 128) return 128;
    return -1;
}

struct ScanGeom {
    int rq, q_blocks, n_chunks, kc;
    long long chunk_len;
    size_t smem;
};

ScanGeom plan_scan(const pe_db* db, long long nq, int k, bool hamming) {
    ScanGeom g;
    g.kc = hamming ? k : k + 2;
    int rq_max;
    if (hamming) rq_max = db->words <= 4 ? 8 : 4;
    else rq_max = db->dim_pad <= 16 ? 8 : (db->dim_pad == 64 ? 2 : 1);
    if (g.kc > 10 && rq_max > 4) rq_max = 4;
    int rq = rq_max;
    /* small query sets: fewer queries per thread so that lanes are not left idle */
    while (rq > 1 && nq <= (long long)kScanThreads * (rq / 2)) rq /= 2;
    if (!hamming && db->dim_pad <= 16 && rq == 2) rq = 1; /* instantiated: 1, 4, 8 */
    if (hamming && rq == 2) rq = 1;
    g.rq = rq;
    g.q_blocks = (int)((nq + (long long)kScanThreads * rq - 1) / ((long long)kScanThreads * rq));
    if (g.q_blocks < 1) g.q_blocks = 1;
    /* about four CTAs per SM in flight over the whole grid; chunks are whole tiles */
    long long want = (4LL * db->num_sms + g.q_blocks - 1) / g.q_blocks;
    const int tile_rows = hamming ? kTile : l2_tile_rows(db->dim_pad);
    const long long tiles = (db->t_local + tile_rows - 1) / tile_rows;
    if (want > tiles) want = tiles;
    if (want < 1) want = 1;
    const long long tiles_per_chunk = (tiles + want - 1) / want;
    g.chunk_len = tiles_per_chunk * tile_rows;
    g.n_chunks = (int)((db->t_local + g.chunk_len - 1) / g.chunk_len);
    if (g.n_chunks < 1) g.n_chunks = 1;
    const size_t lists = (size_t)2 * g.kc * rq * kScanThreads * 4;
    if (hamming) g.smem = (size_t)2 * kTile * db->words * 4 + lists;
    else g.smem = (size_t)2 * tile_rows * (db->dim_pad + 1) * 4 + lists;
    return g;
}

template <int DP, int RQ>
int launch_l2_scan(pe_db* db, const ScanGeom& g, const float* dq, int nq, float* cd, int* ci) {
    PE_CUDA(cudaFuncSetAttribute(knn_l2_scan_kernel<DP, RQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    dim3 grid(g.q_blocks, g.n_chunks);
    knn_l2_scan_kernel<DP, RQ><<<grid, kScanThreads, g.smem, db->stream>>>(
        dq, nq, db->d_desc, db->d_norm_c, db->d_center, db->t_local, g.chunk_len, g.n_chunks, g.kc, db->row_offset, cd, ci);
    PE_LAUNCHED();
    return PE_OK;
}

template <int W, int RQ>
int launch_ham_scan(pe_db* db, const ScanGeom& g, const uint32_t* dq, int nq, int* cd, int* ci) {
    PE_CUDA(cudaFuncSetAttribute(knn_hamming_scan_kernel<W, RQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    dim3 grid(g.q_blocks, g.n_chunks);
    knn_hamming_scan_kernel<W, RQ><<<grid, kScanThreads, g.smem, db->stream>>>(
        dq, nq, db->d_codes, db->t_local, g.chunk_len, g.n_chunks, g.kc, db->row_offset, cd, ci);
    PE_LAUNCHED();
    return PE_OK;
}

}  // namespace

/* local top-k of the shard held by `db` for nq device-resident queries (dense nq x dim) */
int knn_l2_local(pe_db* db, const float* d_queries, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist) {
    PE_REQUIRE(!db->hamming, "pe_knn on a Hamming DB");
    PE_REQUIRE(k >= 1 && k <= 16, "k=%d outside 1..16", k);
    PE_REQUIRE(nq >= 0 && nq < (1LL << 31) / 16, "nq out of range");
    PE_CUDA(cudaSetDevice(db->device));
    if (nq == 0) return PE_OK;
    const int dp = db->dim_pad;
    const float* dq = d_queries;
    if (dp != db->dim) {
        PE_CHECK(db->q.reserve((size_t)nq * dp * sizeof(float)));
        const long long tot = nq * dp;
        pad_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, db->stream>>>(d_queries, db->dim, db->q.as<float>(), dp, nq);
        PE_LAUNCHED();
        dq = db->q.as<float>();
    }
    ScanGeom g = plan_scan(db, nq, k, false);
    const bool tensor = g_opt_knn_tensor.load() != 0 && knn_tc_usable(db, k);
    int tc_tiles_per_chunk = 0;
    if (tensor) knn_tc_plan(db, nq, k, &g.n_chunks, &g.kc, &tc_tiles_per_chunk);
    const size_t n_cand = (size_t)nq * g.n_chunks * g.kc;
    PE_CHECK(db->cand_d.reserve(n_cand * sizeof(float)));
    PE_CHECK(db->cand_i.reserve(n_cand * sizeof(int)));
    float* cd = db->cand_d.as<float>();
    int* ci = db->cand_i.as<int>();
    int st = PE_ERR_UNSUPPORTED;
    /* exactness certificate (pe_set_option("knn_certify")): the select pass lists the queries it cannot prove */
    int* d_flag = nullptr;
    if (g_opt_knn_certify.load() != 0 && db->t_local > 0 && db->d_center) {
        PE_CHECK(db->flag.reserve((size_t)(nq + 4) * sizeof(int)));
        d_flag = db->flag.as<int>();
        PE_CUDA(cudaMemsetAsync(d_flag, 0, 4 * sizeof(int), db->stream));
    } else if (db->flag.p) {
        PE_CUDA(cudaMemsetAsync(db->flag.p, 0, 4 * sizeof(int), db->stream));
    }
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    if (tensor) {
        PE_CHECK(knn_tc_scan(db, dq, nq, k, g.n_chunks, g.kc, tc_tiles_per_chunk, cd, ci));
    } else if (db->t_local > 0) {
#define L2_CASE(DPV, RQV) if (dp == DPV && g.rq == RQV) st = launch_l2_scan<DPV, RQV>(db, g, dq, (int)nq, cd, ci)
        L2_CASE(4, 1); L2_CASE(4, 4); L2_CASE(4, 8);
        L2_CASE(16, 1); L2_CASE(16, 4); L2_CASE(16, 8);
        L2_CASE(64, 1); L2_CASE(64, 2);
        L2_CASE(128, 1);
#undef L2_CASE
        if (st != PE_OK) {
            if (st == PE_ERR_UNSUPPORTED) set_error("no L2 scan kernel for dim_pad=%d rq=%d", dp, g.rq);
            return st;
        }
    } else {
        PE_CUDA(cudaMemsetAsync(ci, 0xff, n_cand * sizeof(int), db->stream)); /* all -1: empty shard */
    }
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    if (tensor) {
        /* the tensor scan emits 32-template blocks: exact pass over the best blocks */
        PE_CHECK(knn_tc_select(db, cd, ci, g.n_chunks, g.kc, dq, nq, k, d_out_idx, d_out_dist, d_flag));
    } else {
        const size_t smem = (size_t)(2 * g.n_chunks * g.kc + dp) * 4;
        PE_REQUIRE(smem <= 200 * 1024, "candidate set too large for the select kernel");
        PE_CUDA(cudaFuncSetAttribute(knn_select_kernel<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        knn_select_kernel<float, true><<<(unsigned)nq, 32, smem, db->stream>>>(
            cd, ci, (long long)g.n_chunks * g.kc, g.kc, g.n_chunks, g.kc, dq, db->d_desc, dp, db->row_offset, k,
            d_out_idx, d_out_dist, db->d_center, db->max_norm2_c, db->t_local > 0 ? d_flag : nullptr);
        PE_LAUNCHED();
    }
    if (d_flag) PE_CHECK(knn_exact_fallback(db, d_flag, dq, nq, k, d_out_idx, d_out_dist));
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    return PE_OK;
}

/* re-scan of the queries listed in d_flag with the direct expression; one early-exit launch when none is listed */
int knn_exact_fallback(pe_db* db, const int* d_flag, const float* d_q, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist) {
    /* row chunks per query: as many as a 64 MB partial-result buffer allows, 4..64 (a power of two) */
    int n_chunks = kFbMaxChunks;
    while (n_chunks > 4 && (size_t)nq * n_chunks * k * 8 > ((size_t)64 << 20)) n_chunks >>= 1;
    int slots = (4 * db->num_sms + n_chunks - 1) / n_chunks;
    if ((int64_t)slots > nq) slots = (int)nq;
    if (slots < 1) slots = 1;
    const size_t part = (size_t)nq * n_chunks * k;
    PE_CHECK(db->fb_d.reserve(part * sizeof(float)));
    PE_CHECK(db->fb_i.reserve(part * sizeof(int)));
    if (db->fb_done.cap < (size_t)nq * sizeof(unsigned)) PE_CHECK(db->fb_done.reserve((size_t)nq * sizeof(unsigned), true));
    const int kt = k <= 1 ? 1 : (k <= 4 ? 4 : (k <= 8 ? 8 : 16));
    const size_t lists = (size_t)(kFbThreads * kt > kFbMaxChunks * kFbMaxK ? kFbThreads * kt : kFbMaxChunks * kFbMaxK);
    const size_t smem = (size_t)(db->dim_pad + 2 * lists) * 4;
    dim3 grid((unsigned)slots, (unsigned)n_chunks);
#define FB_LAUNCH(KT) do { \
        PE_CUDA(cudaFuncSetAttribute(knn_exact_fallback_kernel<KT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        knn_exact_fallback_kernel<KT><<<grid, kFbThreads, smem, db->stream>>>(d_flag, d_q, db->dim_pad, db->d_desc, db->t_local, db->row_offset, k, \
            n_chunks, db->fb_d.as<float>(), db->fb_i.as<int>(), db->fb_done.as<unsigned>(), d_out_idx, d_out_dist); } while (0)
    if (kt == 1) FB_LAUNCH(1); else if (kt == 4) FB_LAUNCH(4); else if (kt == 8) FB_LAUNCH(8); else FB_LAUNCH(16);
#undef FB_LAUNCH
    PE_LAUNCHED();
    return PE_OK;
}

/* DB mean, centred norms and spread of an L2 DB (after db->d_desc is in place) */
int knn_center_stats(pe_db* db) {
    const int dp = db->dim_pad;
    PE_CUDA(cudaSetDevice(db->device));
    if (!db->d_center) PE_CUDA(cudaMalloc(&db->d_center, 128 * sizeof(float)));
    if (!db->d_norm_c) PE_CUDA(cudaMalloc(&db->d_norm_c, (size_t)(db->t_local > 0 ? db->t_local : 1) * sizeof(float)));
    for (int d = 0; d < 128; ++d) db->center[d] = 0.f;
    db->max_abs_c = db->max_norm2_c = 0.f;
    if (db->t_local > 0) {
        void* d_tmp = nullptr;
        PE_CUDA(cudaMalloc(&d_tmp, 128 * sizeof(double) + 2 * sizeof(unsigned)));
        double* d_sum = reinterpret_cast<double*>(d_tmp);
        unsigned* d_max = reinterpret_cast<unsigned*>(d_sum + 128);
        PE_CUDA(cudaMemsetAsync(d_tmp, 0, 128 * sizeof(double) + 2 * sizeof(unsigned), db->stream));
        center_sum_kernel<<<2 * db->num_sms, 1024, 0, db->stream>>>(db->d_desc, dp, db->t_local, d_sum);
        PE_LAUNCHED();
        double h_sum[128];
        PE_CUDA(cudaMemcpyAsync(h_sum, d_sum, sizeof(h_sum), cudaMemcpyDeviceToHost, db->stream));
        PE_CUDA(cudaStreamSynchronize(db->stream));
        for (int d = 0; d < dp; ++d) {
            const double m = h_sum[d] / (double)db->t_local;
            db->center[d] = std::isfinite(m) ? (float)m : 0.f;
        }
        PE_CUDA(cudaMemcpyAsync(db->d_center, db->center, sizeof(db->center), cudaMemcpyHostToDevice, db->stream));
        center_stats_kernel<<<2 * db->num_sms, 256, 0, db->stream>>>(db->d_desc, dp, db->t_local, db->d_center, db->d_norm_c, d_max);
        PE_LAUNCHED();
        unsigned h[2];
        PE_CUDA(cudaMemcpyAsync(h, d_max, sizeof(h), cudaMemcpyDeviceToHost, db->stream));
        PE_CUDA(cudaStreamSynchronize(db->stream));
        cudaFree(d_tmp);
        memcpy(&db->max_abs_c, &h[0], 4);
        memcpy(&db->max_norm2_c, &h[1], 4);
    } else {
        PE_CUDA(cudaMemcpyAsync(db->d_center, db->center, sizeof(db->center), cudaMemcpyHostToDevice, db->stream));
        PE_CUDA(cudaStreamSynchronize(db->stream));
    }
    return PE_OK;
}

int knn_hamming_local(pe_db* db, const uint32_t* d_qcodes, int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist) {
    PE_REQUIRE(db->hamming, "pe_knn_hamming on an L2 DB");
    PE_REQUIRE(k >= 1 && k <= 16, "k=%d outside 1..16", k);
    PE_REQUIRE(nq >= 0 && nq < (1LL << 31) / 16, "nq out of range");
    PE_CUDA(cudaSetDevice(db->device));
    if (nq == 0) return PE_OK;
    ScanGeom g = plan_scan(db, nq, k, true);
    const bool tensor = g_opt_knn_tensor.load() != 0 && knn_tc_usable(db, k);
    int tc_tiles_per_chunk = 0;
    if (tensor) knn_tc_plan(db, nq, k, &g.n_chunks, &g.kc, &tc_tiles_per_chunk);
    const size_t n_cand = (size_t)nq * g.n_chunks * g.kc;
    PE_CHECK(db->cand_d.reserve(n_cand * sizeof(int)));
    PE_CHECK(db->cand_i.reserve(n_cand * sizeof(int)));
    int* cd = db->cand_d.as<int>();
    int* ci = db->cand_i.as<int>();
    int st = PE_ERR_UNSUPPORTED;
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    if (tensor) {
        /* -dot of the +-1 expansions on the tensor pipe (exact integers in fp32), 32-code candidate blocks */
        PE_CHECK(knn_tc_scan(db, d_qcodes, nq, k, g.n_chunks, g.kc, tc_tiles_per_chunk, db->cand_d.as<float>(), ci));
        PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
        PE_CHECK(knn_tc_select_hamming(db, db->cand_d.as<float>(), ci, g.n_chunks, g.kc, d_qcodes, nq, k, d_out_idx, d_out_dist));
        PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
        return PE_OK;
    }
    if (db->t_local > 0) {
#define HAM_CASE(WV, RQV) if (db->words == WV && g.rq == RQV) st = launch_ham_scan<WV, RQV>(db, g, d_qcodes, (int)nq, cd, ci)
        HAM_CASE(1, 1); HAM_CASE(1, 4); HAM_CASE(1, 8);
        HAM_CASE(2, 1); HAM_CASE(2, 4); HAM_CASE(2, 8);
        HAM_CASE(4, 1); HAM_CASE(4, 4); HAM_CASE(4, 8);
        HAM_CASE(8, 1); HAM_CASE(8, 4);
#undef HAM_CASE
        if (st != PE_OK) {
            if (st == PE_ERR_UNSUPPORTED) set_error("no Hamming scan kernel for words=%d rq=%d", db->words, g.rq);
            return st;
        }
    } else {
        PE_CUDA(cudaMemsetAsync(ci, 0xff, n_cand * sizeof(int), db->stream));
    }
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    const size_t smem = (size_t)(2 * g.n_chunks * g.kc) * 4;
    PE_REQUIRE(smem <= 200 * 1024, "candidate set too large for the select kernel");
    PE_CUDA(cudaFuncSetAttribute(knn_select_kernel<int, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    knn_select_kernel<int, false><<<(unsigned)nq, 32, smem, db->stream>>>(
        cd, ci, (long long)g.n_chunks * g.kc, g.kc, g.n_chunks, g.kc, nullptr, nullptr, 0, 0, k, d_out_idx, d_out_dist);
    PE_LAUNCHED();
    PE_CHECK(prof_mark(db->profile, &db->prof_events, db->stream));
    return PE_OK;
}

/* merge `n_lists` per-shard lists; list l of query q starts at l*stride_l + q*k (global indices, exact distances) */
int knn_merge_f32(pe_db* db, const float* d_cd, const int* d_ci, long long stride_l, int n_lists, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist) {
    if (nq == 0) return PE_OK;
    const size_t smem = (size_t)2 * n_lists * k * 4;
    knn_select_kernel<float, false><<<(unsigned)nq, 32, smem, db->stream>>>(
        d_cd, d_ci, k, stride_l, n_lists, k, nullptr, nullptr, 0, 0, k, d_out_idx, d_out_dist);
    PE_LAUNCHED();
    return PE_OK;
}
int knn_merge_i32(pe_db* db, const int* d_cd, const int* d_ci, long long stride_l, int n_lists, int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist) {
    if (nq == 0) return PE_OK;
    const size_t smem = (size_t)2 * n_lists * k * 4;
    knn_select_kernel<int, false><<<(unsigned)nq, 32, smem, db->stream>>>(
        d_cd, d_ci, k, stride_l, n_lists, k, nullptr, nullptr, 0, 0, k, d_out_idx, d_out_dist);
    PE_LAUNCHED();
    return PE_OK;
}

int db_set_descriptors(pe_db* db, const float* src, bool src_on_device) {
    const int dp = pad_dim(db->dim);
    if (dp < 0) { set_error("descriptor dim %d > 128 unsupported", db->dim); return PE_ERR_UNSUPPORTED; }
    db->dim_pad = dp;
    const size_t rows = db->t_local > 0 ? (size_t)db->t_local : 1;
    PE_CUDA(cudaMalloc(&db->d_desc, rows * dp * sizeof(float)));
    PE_CUDA(cudaMalloc(&db->d_norm, rows * sizeof(float)));
    if (db->t_local == 0) return PE_OK;
    const float* dsrc = src;
    float* tmp = nullptr;
    if (!src_on_device) {
        if (dp == db->dim) {
            PE_CUDA(cudaMemcpyAsync(db->d_desc, src, rows * dp * sizeof(float), cudaMemcpyHostToDevice, db->stream));
            dsrc = nullptr;
        } else {
            PE_CUDA(cudaMalloc(&tmp, rows * db->dim * sizeof(float)));
            PE_CUDA(cudaMemcpyAsync(tmp, src, rows * db->dim * sizeof(float), cudaMemcpyHostToDevice, db->stream));
            dsrc = tmp;
        }
    }
    if (dsrc) {
        const long long tot = (long long)rows * dp;
        pad_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, db->stream>>>(dsrc, db->dim, db->d_desc, dp, db->t_local);
        PE_LAUNCHED();
    }
    row_norm_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, db->stream>>>(db->d_desc, dp, db->d_norm, db->t_local);
    PE_LAUNCHED();
    PE_CUDA(cudaStreamSynchronize(db->stream));
    if (tmp) cudaFree(tmp);
    PE_CHECK(knn_center_stats(db));
    return knn_tc_prepare(db);
}

}  // namespace pe
