/*
 * knn_tc.cu - tcgen05 (5th-gen tensor core) scan of the brute-force L2 matcher for descriptors of
 * up to 16 dimensions: the hot loop of `matcher->knnMatch(DBtest, matches, knn)`
 * (reference src/networkevaluator.cpp:197,267; src/networksolver.cpp:472).
 *
 * The scan ranks templates by  R(q,t) = |t|^2 - 2 q.t  (same order as sqrt(sum (q-t)^2) for a fixed q)
 * and emits, per (query, DB chunk), the k+2 best candidates; knn_select_kernel (knn.cu) then recomputes
 * the reference's exact fp32 distance for every candidate and keeps the k smallest by (distance, index).
 *
 * Arithmetic: fp32 operands are split into two fp16 terms (x ~= hi + lo, 22 significant bits) after a
 * power-of-two scaling that keeps them in fp16's normal range; R is accumulated in fp32 in TMEM by
 *     D  = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi + A_n.B_n          (4 x tcgen05.mma, M=128 N=128 K=16)
 * with A = -2 s_q q (queries), B = s_t t (templates) and the last product adding s_q s_t |t|^2 from a
 * three-term fp16 split of the squared norm.
 *
 * Kernel (persistent, one CTA per SM, 352 threads):
 *   warp 0      one lane streams 10 KB template tiles (128 rows: hi | lo | norm) with cp.async.bulk into an
 *               8-stage ring (mbarrier complete_tx)
 *   warps 1-2   one lane each issues the MMAs of one query tile (two tiles = 256 queries share every template
 *               tile), four 128-column accumulators in TMEM (double buffered), tcgen05.commit -> mbarriers
 *   warps 3-10  epilogue (four warps per query tile, one accumulator lane = one query per thread):
 *               tcgen05.ld 32 columns at a time (double buffered), running minimum, and only when a value beats
 *               the query's admission threshold the sorted (k+2)-list in registers is updated.  The
 *               threshold is shared between the CTAs that scan different chunks for the same query
 *               through one float per query in global memory (atomicMin), so updates stay rare.
 */
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "common.h"
#include "umma.cuh"

namespace pe {

using namespace umma;

namespace {

constexpr int kTcRows = 128;                 /* templates per tile / queries per query tile */
constexpr int kTcTileBytes = 10240;          /* hi 4096 | lo 4096 | norm 2048 */
constexpr int kTcStages = 8;
constexpr int kTcThreads = 352;        /* producer warp, 2 MMA warps, 2 x 4 epilogue warps */
constexpr int kAOpBytes = 10240;             /* per query tile: hi 4096 | lo 4096 | a_n 2048 */
constexpr int kSmemA = 0;                                  /* 2 query tiles */
constexpr int kSmemZero = kSmemA + 2 * kAOpBytes;          /* 2048 zero bytes: second K chunk of A_n */
constexpr int kSmemB = kSmemZero + 2048;
constexpr int kSmemScratch = kSmemB + kTcStages * kTcTileBytes;   /* [32][256] floats: rare-path staging  */
constexpr int kSmemList = kSmemScratch + 32 * 256 * 4;            /* [7][256] floats + [7][256] ints     */
constexpr int kSmemBar = kSmemList + 2 * 7 * 256 * 4;
constexpr int kSmemTotal = kSmemBar + 256;

/* monotone float <-> int map so that atomicMin on ints orders floats (negative values included) */
__device__ __forceinline__ int f2ord(float f) {
    const int b = __float_as_int(f);
    return b ^ ((b >> 31) & 0x7FFFFFFF);
}
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o ^ ((o >> 31) & 0x7FFFFFFF)); }

__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
    return static_cast<uint32_t>(__half_as_ushort(a)) | (static_cast<uint32_t>(__half_as_ushort(b)) << 16);
}

/* ------------------------------------------------------------------ DB pack ---- */
/* The L2 distance is translation invariant, so the scan works on descriptors centred on the DB mean c:
 * |t-c| and |q-c| are what the cancellation in |t|^2 - 2 q.t is measured against, not the distance of the
 * cluster from the origin. */
__global__ void tc_mean_kernel(const float* __restrict__ desc, int ld, long long t_local, double* __restrict__ sum16) {
    __shared__ double red[16][8];
    float acc[16];
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[d] = 0.f;
    double tot[16];
#pragma unroll
    for (int d = 0; d < 16; ++d) tot[d] = 0.0;
    int n = 0;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < t_local; r += (long long)gridDim.x * blockDim.x) {
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[d] += d < ld ? desc[r * ld + d] : 0.f;
        if (++n == 64) {   /* flush the fp32 partials into doubles before they lose bits */
#pragma unroll
            for (int d = 0; d < 16; ++d) { tot[d] += acc[d]; acc[d] = 0.f; }
            n = 0;
        }
    }
#pragma unroll
    for (int d = 0; d < 16; ++d) {
        double v = tot[d] + acc[d];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[d][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 16) {
        double v = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += red[threadIdx.x][w];
        atomicAdd(sum16 + threadIdx.x, v);
    }
}

struct TcCenter { float c[16]; };

/* out[0] = max |t-c| over all elements, out[1] = max |t-c|^2 over rows (as uint bit patterns of non-negative floats) */
__global__ void tc_stats_kernel(const float* __restrict__ desc, int ld, long long t_local, const TcCenter ctr, unsigned* __restrict__ out) {
    float m = 0.f, mn = 0.f;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < t_local; r += (long long)gridDim.x * blockDim.x) {
        float n = 0.f;
#pragma unroll
        for (int d = 0; d < 16; ++d) {
            const float v = d < ld ? desc[r * ld + d] - ctr.c[d] : 0.f;
            m = fmaxf(m, fabsf(v));
            n = fmaf(v, v, n);
        }
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

/* desc: [t][ld] fp32 (ld = 4 or 16, zero padded).  out tile j (128 rows): hi[2][128][8] lo[2][128][8] n[128][8] halves */
__global__ void tc_pack_db_kernel(const float* __restrict__ desc, int ld, long long t_local, const TcCenter ctr,
                                  float s_t, float s_n, uint4* __restrict__ out) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long tiles = (t_local + kTcRows - 1) / kTcRows;
    if (r >= tiles * kTcRows) return;
    const long long tile = r / kTcRows;
    const int i = (int)(r - tile * kTcRows);
    uint4* base = out + tile * (kTcTileBytes / 16);
    uint32_t hi[8], lo[8];
    if (r < t_local) {
        float v[16], nrm = 0.f;
#pragma unroll
        for (int d = 0; d < 16; ++d) {
            v[d] = d < ld ? desc[r * ld + d] - ctr.c[d] : 0.f;
            nrm = fmaf(v[d], v[d], nrm);
        }
#pragma unroll
        for (int d = 0; d < 16; d += 2) {
            __half h0, l0, h1, l1;
            split_f16(v[d] * s_t, h0, l0);
            split_f16(v[d + 1] * s_t, h1, l1);
            hi[d / 2] = pack_h2(h0, h1);
            lo[d / 2] = pack_h2(l0, l1);
        }
        const float n = nrm * s_n;
        const __half n1 = __float2half_rn(n);
        const float r1 = n - __half2float(n1);
        const __half n2 = __float2half_rn(r1);
        const __half n3 = __float2half_rn(r1 - __half2float(n2));
        base[512 + i] = make_uint4(pack_h2(n1, n2), pack_h2(n3, __ushort_as_half(0)), 0u, 0u);
    } else {
        /* padding rows: zero vector, largest norm term; the epilogue also rejects them by index */
#pragma unroll
        for (int j = 0; j < 8; ++j) hi[j] = lo[j] = 0u;
        const __half big = __ushort_as_half(0x7BFF); /* 65504 */
        base[512 + i] = make_uint4(pack_h2(big, big), pack_h2(big, __ushort_as_half(0)), 0u, 0u);
    }
    base[0 + i] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    base[128 + i] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    base[256 + i] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    base[384 + i] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
}

/* ------------------------------------------------------------------- scan ---- */
struct TcArgs {
    const float* q;            /* [nq][q_ld], q_ld = 4 or 16 (zero padded) */
    int nq, q_ld;
    const uint4* db;           /* packed tiles */
    long long t_local, row_offset;
    int tiles_total, tiles_per_chunk, n_chunks, n_qpairs;
    int log2_st, log2_sn;      /* s_t = 2^log2_st, s_n = s_t * 2^-c = 2^log2_sn  (norm term scale) */
    int* tau;                  /* [nq] shared admission thresholds (ordered-int floats) */
    TcCenter ctr;              /* DB mean: queries are centred the same way as the packed templates */
    float* cand_d;
    int* cand_i;
    int* err;
    unsigned long long* dbg;   /* PE_TC_DEBUG: per-role cycle counters (null in production) */
};
#define TC_DBG_T0(var) long long var = 0; if (a.dbg) var = clock64()
#define TC_DBG_ADD(slot, var) do { if (a.dbg) { const long long n_ = clock64(); dbg_acc[slot] += n_ - var; var = n_; } } while (0)

/* sorted insert into the calling thread's k+2 list (entries strided by 256 threads in shared memory); strict '>'
 * keeps the earlier (lower-index) entry first among equals.  Returns the new k+2-th best.  Deliberately not
 * inlined: it runs in about one warp-tile in six and must not bloat the hot loop's instruction footprint. */
__device__ __noinline__ float tc_list_insert(float* ld, int* li, int kc, float v, int idx) {
    int p = kc - 1;
    while (p > 0 && ld[(p - 1) * 256] > v) {
        ld[p * 256] = ld[(p - 1) * 256];
        li[p * 256] = li[(p - 1) * 256];
        --p;
    }
    ld[p * 256] = v;
    li[p * 256] = idx;
    return ld[(kc - 1) * 256];
}

template <int KC>
__global__ void __launch_bounds__(kTcThreads, 1) knn_tc_scan_kernel(const TcArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemBar);
    uint64_t* full = bars;                     /* [kTcStages] tile landed */
    uint64_t* empty = bars + kTcStages;        /* [kTcStages] tile consumed by both query tiles' MMAs */
    uint64_t* tfull = bars + 2 * kTcStages;    /* [4] accumulator ready */
    uint64_t* tempty = tfull + 4;              /* [4] accumulator drained */
    uint64_t* a_ready = tempty + 4;            /* query operands of the next item written */
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_ready + 1);
    volatile int* s_abort = reinterpret_cast<volatile int*>(tmem_slot + 1);
    float* s_red = reinterpret_cast<float*>(tmem_slot + 2);   /* [8] query abs-max per epilogue warp */
    float* s_scratch = reinterpret_cast<float*>(smem + kSmemScratch);   /* [32][256] rare-path staging */
    float* s_ld = reinterpret_cast<float*>(smem + kSmemList);           /* [KC][256] candidate values */
    int* s_li = reinterpret_cast<int*>(s_ld + KC * 256);                /* [KC][256] candidate rows   */

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 2048 / 16; i += kTcThreads) reinterpret_cast<uint4*>(smem + kSmemZero)[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 2); }
        for (int b = 0; b < 4; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 4); }
        mbar_init(a_ready, 256);
        *s_abort = 0;
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const int n_items = a.n_chunks * a.n_qpairs;

    if (warp == 0) {
        /* ================= producer ================= */
        if (lane == 0) {
            uint32_t g = 0;
            for (int item = blockIdx.x; item < n_items && !*s_abort; item += gridDim.x) {
                const int chunk = item / a.n_qpairs;
                const int t0 = chunk * a.tiles_per_chunk;
                const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
                for (int t = t0; t < t1; ++t, ++g) {
                    const int s = g % kTcStages;
                    if (!mbar_wait(&empty[s], ((g / kTcStages) & 1) ^ 1)) { atomicExch(a.err, 1); *s_abort = 1; break; }
                    mbar_expect_tx(&full[s], kTcTileBytes);
                    bulk_g2s(smem + kSmemB + s * kTcTileBytes, a.db + (size_t)t * (kTcTileBytes / 16), kTcTileBytes, &full[s]);
                }
            }
        }
    } else if (warp <= 2) {
        /* ================= MMA issuers: warp 1 feeds query tile 0, warp 2 query tile 1 ================= */
        if (lane == 0) {
            const int qt = warp - 1;
            const uint32_t idesc = instr_desc(128, 128, FMT_F16, FMT_F16);
            const uint32_t sB = smem_u32(smem + kSmemB), sZ = smem_u32(smem + kSmemZero);
            const uint32_t a_base = smem_u32(smem + kSmemA) + qt * kAOpBytes;
            const uint64_t a_hi = smem_desc(a_base, 2048, 128), a_lo = smem_desc(a_base + 4096, 2048, 128);
            const uint64_t a_n = smem_desc(a_base + 8192, sZ - (a_base + 8192), 128);
            const uint64_t b_hi0 = smem_desc(sB, 2048, 128), b_lo0 = smem_desc(sB + 4096, 2048, 128);
            const uint64_t b_n0 = smem_desc(sB + 8192, 0, 128);   /* chunk 1 aliases chunk 0: A_n's chunk 1 is zero */
            uint32_t g = 0, it = 0;
            long long dbg_acc[4] = {0, 0, 0, 0};
            TC_DBG_T0(tm);
            for (int item = blockIdx.x; item < n_items && !*s_abort; item += gridDim.x, ++it) {
                const int chunk = item / a.n_qpairs;
                const int t0 = chunk * a.tiles_per_chunk;
                const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
                if (!mbar_wait(a_ready, it & 1)) { atomicExch(a.err, 2); *s_abort = 1; break; }
                fence_after_sync();
                for (int t = t0; t < t1; ++t, ++g) {
                    const int s = g % kTcStages;
                    const int b = (g & 1) * 2 + qt;
                    TC_DBG_ADD(2, tm);
                    if (!mbar_wait(&tempty[b], ((g >> 1) & 1) ^ 1)) { atomicExch(a.err, 4); *s_abort = 1; break; }
                    TC_DBG_ADD(1, tm);
                    if (!mbar_wait(&full[s], (g / kTcStages) & 1)) { atomicExch(a.err, 3); *s_abort = 1; break; }
                    fence_after_sync();
                    TC_DBG_ADD(0, tm);
                    const uint32_t off = static_cast<uint32_t>(s) * (kTcTileBytes >> 4);
                    const uint32_t d = tmem + b * 128;
                    mma_f16(d, a_hi, b_hi0 + off, idesc, 0);
                    mma_f16(d, a_hi, b_lo0 + off, idesc, 1);
                    mma_f16(d, a_lo, b_hi0 + off, idesc, 1);
                    mma_f16(d, a_n, b_n0 + off, idesc, 1);
                    mma_commit(&tfull[b]);
                    mma_commit(&empty[s]);
                }
            }
            if (a.dbg && qt == 0) {
                TC_DBG_ADD(2, tm);
                atomicAdd(a.dbg + 5, (unsigned long long)dbg_acc[0]);
                atomicAdd(a.dbg + 6, (unsigned long long)dbg_acc[1]);
                atomicAdd(a.dbg + 7, (unsigned long long)dbg_acc[2]);
                atomicAdd(a.dbg + 8, (unsigned long long)g);
            }
        }
    } else {
        /* ================= epilogue: warps 3..6 drain query tile 0, warps 7..10 query tile 1 ================= */
        const int quarter = warp & 3;                 /* TMEM lanes this warp may read: 32 * (warp % 4) */
        const int qt = (warp - 3) >> 2;               /* query tile of the pair */
        const int row = quarter * 32 + lane;          /* query row inside the query tile */
        const int et = qt * 128 + row;                /* 0..255: this thread's column in the shared-memory lists */
        float* my_ld = s_ld + et;
        int* my_li = s_li + et;
        float* my_scratch = s_scratch + et;
        uint32_t g = 0;
        long long dbg_acc[5] = {0, 0, 0, 0, 0};
        TC_DBG_T0(te);
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int chunk = item / a.n_qpairs, qpair = item - chunk * a.n_qpairs;
            const int t0 = chunk * a.tiles_per_chunk;
            const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
            /* ---- query operands: centre, scale, split, store in the K-major core-matrix layout */
            const int qidx = (qpair * 2 + qt) * kTcRows + row;
            const bool q_ok = qidx < a.nq;
            float qv[16];
            {
                float m = 0.f;
#pragma unroll
                for (int d4 = 0; d4 < 4; ++d4) {
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (q_ok && 4 * d4 < a.q_ld) {
                        v = __ldg(reinterpret_cast<const float4*>(a.q + (size_t)qidx * a.q_ld) + d4);
                        v.x -= a.ctr.c[4 * d4 + 0]; v.y -= a.ctr.c[4 * d4 + 1]; v.z -= a.ctr.c[4 * d4 + 2]; v.w -= a.ctr.c[4 * d4 + 3];
                    }
                    qv[4 * d4 + 0] = v.x; qv[4 * d4 + 1] = v.y; qv[4 * d4 + 2] = v.z; qv[4 * d4 + 3] = v.w;
                    m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                if (lane == 0) s_red[qt * 4 + quarter] = m;
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                const float m = fmaxf(fmaxf(s_red[qt * 4 + 0], s_red[qt * 4 + 1]), fmaxf(s_red[qt * 4 + 2], s_red[qt * 4 + 3]));
                /* s_q = 2^eq with 2 * max|q| * s_q in [2^10, 2^11); a_n = s_q * s_t / s_n must stay a finite fp16 power of two */
                int eq = (m > 0.f && isfinite(m)) ? 9 - ilogbf(m) : 0;
                int ea = eq + a.log2_st - a.log2_sn;
                if (ea > 15) { eq -= ea - 15; ea = 15; }
                const float s_q = ldexpf(1.f, eq);
                const __half an = ea < -24 ? __ushort_as_half(0) : __float2half_rn(ldexpf(1.f, ea));
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int d = 0; d < 16; d += 2) {
                    __half h0, l0, h1, l1;
                    split_f16(-2.f * s_q * qv[d], h0, l0);
                    split_f16(-2.f * s_q * qv[d + 1], h1, l1);
                    hi[d / 2] = pack_h2(h0, h1);
                    lo[d / 2] = pack_h2(l0, l1);
                }
                uint4* op = reinterpret_cast<uint4*>(smem + kSmemA + qt * kAOpBytes);
                op[0 + row] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                op[128 + row] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                op[256 + row] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                op[384 + row] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                op[512 + row] = make_uint4(pack_h2(an, an), pack_h2(an, __ushort_as_half(0)), 0u, 0u);
#pragma unroll
                for (int p = 0; p < KC; ++p) { my_ld[p * 256] = INFINITY; my_li[p * 256] = -1; }
            }
            fence_async_smem();
            mbar_arrive(a_ready);
            asm volatile("bar.sync 1, 256;" ::: "memory");   /* s_red may be rewritten by the next item */

            float thr = INFINITY;                              /* k+2-th best of this chunk so far */
            /* the CTAs that scan other chunks for this query publish their k+2-th best in tau: nothing above it
             * can reach the final top-k.  Admission is v <= tau, i.e. v < nextup(tau).  tau is re-read every fourth
             * tile and used from the following tile on (a stale tau is only more permissive). */
            int tau_raw = q_ok ? __ldcg(a.tau + qidx) : 0x7F7F7F7F;
            float tau_up = INFINITY;
            for (int t = t0; t < t1; ++t, ++g) {
                const int col0 = t * kTcRows;                 /* local template index of column 0 */
                const int b = (g & 1) * 2 + qt;
                if ((g & 3) == 1) {
                    const float tg = ord2f(tau_raw);
                    tau_up = __int_as_float(tg >= 0.f ? __float_as_int(tg) + 1 : __float_as_int(tg) - 1);
                    if (tg == 0.f) tau_up = FLT_MIN;
                } else if ((g & 3) == 0 && q_ok) {
                    tau_raw = __ldcg(a.tau + qidx);
                }
                float th = fminf(thr, tau_up);
                TC_DBG_ADD(3, te);
                if (!mbar_wait(&tfull[b], (g >> 1) & 1)) { atomicExch(a.err, 5); *s_abort = 1; break; }
                fence_after_sync();
                TC_DBG_ADD(0, te);
                const uint32_t tbase = tmem + b * 128 + (static_cast<uint32_t>(quarter * 32) << 16);
                uint32_t v0[32], v1[32], v2[32], v3[32];
                auto min32 = [](const uint32_t (&v)[32]) {
                    float m8[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        m8[j] = fminf(fminf(__uint_as_float(v[j]), __uint_as_float(v[j + 8])), fminf(__uint_as_float(v[j + 16]), __uint_as_float(v[j + 24])));
                    return fminf(fminf(fminf(m8[0], m8[1]), fminf(m8[2], m8[3])), fminf(fminf(m8[4], m8[5]), fminf(m8[6], m8[7])));
                };
                /* hot path: the whole 128-column accumulator row in one round trip (four loads in flight) */
                tmem_ld32(tbase, v0);
                tmem_ld32(tbase + 32, v1);
                tmem_ld32(tbase + 64, v2);
                tmem_ld32(tbase + 96, v3);
                tmem_ld_wait();
                const float m0 = min32(v0), m1 = min32(v1), m2 = min32(v2), m3 = min32(v3);
                const unsigned hit = (m0 < th ? 1u : 0u) | (m1 < th ? 2u : 0u) | (m2 < th ? 4u : 0u) | (m3 < th ? 8u : 0u);
                const unsigned any = __reduce_or_sync(0xffffffffu, hit);
                TC_DBG_ADD(1, te);
                if (any) {
                    /* rare path, kept small (instruction cache): re-read a flagged 32-column block, park it in shared
                     * memory, build the mask of admissible columns and insert those (usually one) */
                    dbg_acc[4] += 1;
                    bool changed = false;
#pragma unroll 1
                    for (int blk = 0; blk < 4; ++blk) {
                        if (!((any >> blk) & 1u)) continue;
                        uint32_t w[32];
                        tmem_ld32(tbase + blk * 32, w);
                        tmem_ld_wait();
                        unsigned mask = 0;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            my_scratch[j * 256] = __uint_as_float(w[j]);
                            mask |= (__uint_as_float(w[j]) < th ? 1u : 0u) << j;
                        }
                        while (mask) {
                            const int j = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const float x = my_scratch[j * 256];
                            const int li = col0 + blk * 32 + j;
                            if (x < th && li < a.t_local) {
                                thr = tc_list_insert(my_ld, my_li, KC, x, li);
                                th = fminf(thr, tau_up);
                                changed = true;
                            }
                        }
                    }
                    if (changed && thr < INFINITY && q_ok) atomicMin(a.tau + qidx, f2ord(thr));
                    __syncwarp();
                    TC_DBG_ADD(2, te);
                }
                fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[b]);
            }
            /* ---- this item's candidates: [query][chunk][KC], global template indices */
            if (q_ok) {
                const size_t o = ((size_t)qidx * a.n_chunks + chunk) * KC;
#pragma unroll
                for (int p = 0; p < KC; ++p) {
                    const int li = my_li[p * 256];
                    a.cand_d[o + p] = my_ld[p * 256];
                    a.cand_i[o + p] = li >= 0 ? (int)(li + a.row_offset) : -1;
                }
            }
        }
        if (a.dbg && lane == 0) {
            TC_DBG_ADD(3, te);
            for (int j = 0; j < 5; ++j) atomicAdd(a.dbg + j, (unsigned long long)dbg_acc[j]);
            atomicAdd(a.dbg + 9, (unsigned long long)g);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

}  // namespace

/* packs the (already padded, normed) fp32 DB of `db` into tensor tiles; called once per DB */
int knn_tc_prepare(pe_db* db) {
    db->tc_ready = false;
    if (db->hamming || (db->dim_pad != 16 && db->dim_pad != 4) || db->t_local <= 0) return PE_OK;
    PE_CUDA(cudaSetDevice(db->device));
    const int ld = db->dim_pad;
    void* d_tmp = nullptr;
    PE_CUDA(cudaMalloc(&d_tmp, 16 * sizeof(double) + 2 * sizeof(unsigned)));
    double* d_sum = reinterpret_cast<double*>(d_tmp);
    unsigned* d_max = reinterpret_cast<unsigned*>(d_sum + 16);
    PE_CUDA(cudaMemsetAsync(d_tmp, 0, 16 * sizeof(double) + 2 * sizeof(unsigned), db->stream));
    tc_mean_kernel<<<296, 256, 0, db->stream>>>(db->d_desc, ld, db->t_local, d_sum);
    PE_LAUNCHED();
    double h_sum[16];
    PE_CUDA(cudaMemcpyAsync(h_sum, d_sum, sizeof(h_sum), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    TcCenter ctr;
    for (int d = 0; d < 16; ++d) {
        const double m = h_sum[d] / (double)db->t_local;
        ctr.c[d] = std::isfinite(m) ? (float)m : 0.f;
        db->tc_center[d] = ctr.c[d];
    }
    tc_stats_kernel<<<296, 256, 0, db->stream>>>(db->d_desc, ld, db->t_local, ctr, d_max);
    PE_LAUNCHED();
    unsigned h[2];
    PE_CUDA(cudaMemcpyAsync(h, d_max, sizeof(h), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    cudaFree(d_tmp);
    float mt, mn;
    memcpy(&mt, &h[0], 4);
    memcpy(&mn, &h[1], 4);
    /* s_t: max |s_t (t-c)| in [2^10, 2^11);  s_n: max s_n |t-c|^2 in [2^14, 2^15) */
    db->tc_log2_st = mt > 0.f ? 10 - ilogbf(mt) : 0;
    db->tc_log2_sn = mn > 0.f ? 14 - ilogbf(mn) : 0;
    const long long tiles = (db->t_local + kTcRows - 1) / kTcRows;
    PE_CUDA(cudaMalloc(&db->d_tc, (size_t)tiles * kTcTileBytes));
    tc_pack_db_kernel<<<(unsigned)((tiles * kTcRows + 255) / 256), 256, 0, db->stream>>>(
        db->d_desc, ld, db->t_local, ctr, ldexpf(1.f, db->tc_log2_st), ldexpf(1.f, db->tc_log2_sn),
        reinterpret_cast<uint4*>(db->d_tc));
    PE_LAUNCHED();
    /* pinned, device-visible flag: the kernels write it only when a pipeline barrier times out */
    PE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&db->d_tc_err), sizeof(int), cudaHostAllocMapped));
    *db->d_tc_err = 0;
    PE_CUDA(cudaStreamSynchronize(db->stream));
    db->tc_ready = true;
    return PE_OK;
}

bool knn_tc_usable(const pe_db* db, int k) { return db->tc_ready && k + 2 <= 7; }

/* plans the chunking; returns n_chunks and kc through the arguments */
void knn_tc_plan(const pe_db* db, int64_t nq, int k, int* n_chunks, int* kc, int* tiles_per_chunk) {
    const int n_qpairs = (int)((nq + 2 * kTcRows - 1) / (2 * kTcRows));
    const long long tiles = (db->t_local + kTcRows - 1) / kTcRows;
    long long c = (4LL * db->num_sms + n_qpairs - 1) / n_qpairs;
    const long long cmax = tiles / 32 > 0 ? tiles / 32 : 1;
    if (c > cmax) c = cmax;
    const long long tpc = (tiles + c - 1) / c;
    *tiles_per_chunk = (int)tpc;
    *n_chunks = (int)((tiles + tpc - 1) / tpc);
    *kc = k <= 1 ? 3 : 7;
}

int knn_tc_scan(pe_db* db, const float* d_q, int64_t nq, int k, int n_chunks, int kc, int tiles_per_chunk,
                float* cand_d, int* cand_i) {
    PE_CHECK(db->tau.reserve((size_t)nq * sizeof(int)));
    PE_CUDA(cudaMemsetAsync(db->tau.p, 0x7f, (size_t)nq * sizeof(int), db->stream));
    TcArgs a;
    a.q = d_q;
    a.nq = (int)nq;
    a.q_ld = db->dim_pad;
    a.db = reinterpret_cast<const uint4*>(db->d_tc);
    a.t_local = db->t_local;
    a.row_offset = db->row_offset;
    a.tiles_total = (int)((db->t_local + kTcRows - 1) / kTcRows);
    a.tiles_per_chunk = tiles_per_chunk;
    a.n_chunks = n_chunks;
    a.n_qpairs = (int)((nq + 2 * kTcRows - 1) / (2 * kTcRows));
    a.log2_st = db->tc_log2_st;
    a.log2_sn = db->tc_log2_sn;
    a.tau = db->tau.as<int>();
    for (int d = 0; d < 16; ++d) a.ctr.c[d] = db->tc_center[d];
    a.cand_d = cand_d;
    a.cand_i = cand_i;
    a.err = db->d_tc_err;
    a.dbg = nullptr;
    static const bool dbg_on = getenv("PE_TC_DEBUG") != nullptr;
    unsigned long long* d_dbg = nullptr;
    if (dbg_on) {
        PE_CUDA(cudaMalloc(&d_dbg, 16 * sizeof(unsigned long long)));
        PE_CUDA(cudaMemsetAsync(d_dbg, 0, 16 * sizeof(unsigned long long), db->stream));
        a.dbg = d_dbg;
    }
    const int items = a.n_chunks * a.n_qpairs;
    const int grid = items < db->num_sms ? items : db->num_sms;
    if (kc == 3) {
        PE_CUDA(cudaFuncSetAttribute(knn_tc_scan_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
        knn_tc_scan_kernel<3><<<grid, kTcThreads, kSmemTotal, db->stream>>>(a);
    } else {
        PE_CUDA(cudaFuncSetAttribute(knn_tc_scan_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
        knn_tc_scan_kernel<7><<<grid, kTcThreads, kSmemTotal, db->stream>>>(a);
    }
    PE_LAUNCHED();
    if (dbg_on) {
        unsigned long long h[16];
        PE_CUDA(cudaMemcpyAsync(h, d_dbg, sizeof(h), cudaMemcpyDeviceToHost, db->stream));
        PE_CUDA(cudaStreamSynchronize(db->stream));
        cudaFree(d_dbg);
        const double ew = 8.0 * grid, tiles_e = h[9] > 0 ? (double)h[9] : 1.0, tiles_m = h[8] > 0 ? (double)h[8] : 1.0;
        fprintf(stderr, "[tc dbg] grid %d items %d  epilogue per warp-tile: wait_tfull %.0f hot %.0f slow %.0f other %.0f cycles, slow entries %.3f/tile | "
                        "mma per tile: wait_full %.0f wait_tempty %.0f issue %.0f (epi warps %.0f)\n", grid, items,
                h[0] / tiles_e, h[1] / tiles_e, h[2] / tiles_e, h[3] / tiles_e, h[4] / tiles_e, h[5] / tiles_m, h[6] / tiles_m, h[7] / tiles_m, ew);
    }
    return PE_OK;
}

/* after a synchronisation: nonzero means a pipeline barrier timed out inside the scan kernel */
int knn_tc_check(pe_db* db) {
    if (!db->d_tc_err) return PE_OK;
    const int e = *reinterpret_cast<volatile int*>(db->d_tc_err);
    if (e != 0) {
        set_error("tensor kNN scan: pipeline barrier %d timed out", e);
        *db->d_tc_err = 0;
        return PE_ERR_CUDA;
    }
    return PE_OK;
}

}  // namespace pe
