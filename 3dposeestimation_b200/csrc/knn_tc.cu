/*
 * knn_tc.cu - tcgen05 (5th-gen tensor core) scan of the brute-force L2 matcher for descriptors of
 * up to 128 dimensions and k <= 16: the hot loop of `matcher->knnMatch(DBtest, matches, knn)`
 * (reference src/networkevaluator.cpp:197,267; src/networksolver.cpp:472) - the 16-d nets, the 64-d
 * manifold_wang (network/manifold_wang.prototxt:104) and the 128 float outputs of bin_ip2
 * (network/manifold_rgbdnor_16_bin.prototxt:142-154) alike.
 *
 * The scan ranks templates by  R(q,t) = |t|^2 - 2 q.t  (same order as sqrt(sum (q-t)^2) for a fixed q)
 * and emits, per (query, DB chunk), the k+2 best 32-template blocks (keyed by their smallest R);
 * knn_tc_select_kernel then recomputes the reference's exact fp32 distance for every template of the best
 * blocks and keeps the k smallest by (distance, index).  The scan therefore never has to locate a column.
 *
 * Arithmetic: fp32 operands are split into two fp16 terms (x ~= hi + lo, 22 significant bits) after a
 * power-of-two scaling that keeps them in fp16's normal range; R is accumulated in fp32 in TMEM by
 *     D  = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi + A_n.B_n          (4 x tcgen05.mma, M=128 N=128 K=16)
 * with A = -2 s_q q (queries), B = s_t t (templates) and the last product adding s_q s_t |t|^2 from a
 * three-term fp16 split of the squared norm.  Descriptors wider than 16 dimensions run dim/16 K chunks
 * (3 MMAs each) into the same accumulator; a template tile then travels through the shared-memory ring
 * as dim/16 stages of 8 KB (hi | lo of one chunk; the first also carries the 2 KB norm term).
 *
 * Certificate: knn_tc_select_kernel proves per query that no template outside the re-ranked blocks can
 * beat the reported k-th neighbour (rigorous error bound of the split arithmetic, see the kernel) and
 * flags the queries it cannot prove; those are re-scanned with the direct fp32 (q-t)^2 expression
 * (knn.cu knn_exact_fallback), so the result never depends on the precision of the ranking pass.
 *
 * Kernel (persistent, one CTA per SM, 608 threads):
 *   warp 0      one lane streams 10 KB template tiles (128 rows: hi | lo | norm) with cp.async.bulk into an
 *               8-stage ring (mbarrier complete_tx)
 *   warps 1-2   one lane each issues the MMAs of one query tile (two tiles = 256 queries share every template
 *               tile), four 128-column accumulators in TMEM (double buffered), tcgen05.commit -> mbarriers
 *   warps 3-18  epilogue (per query tile two sets of four warps that drain alternate tiles; one accumulator lane =
 *               one query per thread):
 *               tcgen05.ld of the accumulator row, minimum of each 32-column block, and only when a block minimum
 *               beats the query's admission threshold is the sorted (k+2)-list in shared memory updated.  The
 *               threshold is shared between the CTAs that scan different chunks for the same query
 *               through one float per query in global memory (atomicMin), so updates stay rare.
 */
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "common.h"
#include "knn_exact.cuh"
#include "umma.cuh"

namespace pe {

using namespace umma;

namespace {

constexpr int kTcRows = 128;                 /* templates per tile / queries per query tile */
constexpr int kTcChunkBytes = 8192;          /* one 16-dim K chunk of a tile: hi 4096 | lo 4096 */
constexpr int kTcStageBytes = 10240;         /* ring stage: a chunk + the 2048-byte norm term (first chunk of a tile only) */
constexpr int kTcBlock = 32;                 /* templates per candidate block */
constexpr int kTcMaxSel = 19;                /* candidate blocks re-ranked exactly per query: k + 3, k <= 16 */
constexpr int kTcMaxStages = 8;
constexpr int kTcThreads = 608;               /* producer warp, 2 MMA warps, 2 x 2 x 4 epilogue warps */
constexpr int kTcListCols = 512;              /* one list column per epilogue thread */
constexpr int kHamBytes = 16384;             /* Hamming: per tile up to 4 K-steps of [2 chunks][128 rows][16 x e4m3] */
constexpr int kTcSmemMax = 232448;           /* 227 KB of dynamic shared memory per CTA on sm_100 */

/* shared-memory plan of one launch (bytes); MODE 0 = L2 (split fp16), MODE 1 = Hamming (+-1 as e4m3) */
struct TcPlan {
    int nch;           /* L2: K chunks of 16 dims per tile (dim_pad / 16, at least 1); Hamming: 1 */
    int stages;        /* ring stages (one K chunk each) */
    int stage_bytes;
    int a_bytes;       /* one query tile: nch x (hi | lo) + the a_n term */
    int off_zero;      /* L2: 2048 zero bytes, second K chunk of A_n */
    int off_b, off_list, off_bar, total;
};

/* false when even two ring stages do not fit (e.g. D = 128 with k + 2 = 18 list entries still fits: 2 stages) */
bool tc_make_plan(bool hamming, int dim_pad, int kc, TcPlan* p) {
    p->nch = hamming ? 1 : (dim_pad > 16 ? dim_pad / 16 : 1);
    p->a_bytes = hamming ? kHamBytes : p->nch * kTcChunkBytes + 2048;
    p->stage_bytes = hamming ? kHamBytes : kTcStageBytes;
    p->off_zero = 2 * p->a_bytes;
    p->off_b = p->off_zero + (hamming ? 0 : 2048);
    const int list_bytes = kc > 3 ? 2 * kc * kTcListCols * 4 : 0;
    const int avail = kTcSmemMax - 512 - p->off_b - list_bytes;
    p->stages = avail / p->stage_bytes;
    if (p->stages > kTcMaxStages) p->stages = kTcMaxStages;
    if (p->stages < 2) return false;
    p->off_list = p->off_b + p->stages * p->stage_bytes;
    p->off_bar = p->off_list + list_bytes;
    p->total = p->off_bar + 512;
    return true;
}

int tc_kc_for(int k) { return k <= 1 ? 3 : (k <= 5 ? 7 : 18); }

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
/* desc: [t][ld] fp32 (ld = 4, 16, 64 or 128, zero padded), ctr: the DB mean (knn.cu knn_center_stats).
 * out tile j (128 rows): [chunk 0: hi[2][128][8] | lo[2][128][8] | n[128][8]] [chunk c >= 1: hi | lo] halves, chunk c =
 * dims 16c .. 16c + 15 of the centred, scaled rows */
__global__ void tc_pack_db_kernel(const float* __restrict__ desc, int ld, int nch, long long t_local, const float* __restrict__ ctr,
                                  float s_t, float s_n, uint4* __restrict__ out) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long tiles = (t_local + kTcRows - 1) / kTcRows;
    if (r >= tiles * kTcRows) return;
    const long long tile = r / kTcRows;
    const int i = (int)(r - tile * kTcRows);
    const size_t tile16 = (size_t)(nch * kTcChunkBytes + 2048) / 16;
    uint4* base = out + tile * tile16;
    float nrm = 0.f;
    for (int c = 0; c < nch; ++c) {
        uint32_t hi[8], lo[8];
        if (r < t_local) {
            float v[16];
#pragma unroll
            for (int d = 0; d < 16; ++d) {
                const int dd = 16 * c + d;
                v[d] = dd < ld ? desc[r * ld + dd] - ctr[dd] : 0.f;
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
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) hi[j] = lo[j] = 0u;
        }
        uint4* cb = base + (c == 0 ? 0 : (kTcStageBytes + (size_t)(c - 1) * kTcChunkBytes) / 16);
        cb[0 + i] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        cb[128 + i] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        cb[256 + i] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        cb[384 + i] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
    }
    if (r < t_local) {
        const float n = nrm * s_n;
        const __half n1 = __float2half_rn(n);
        const float r1 = n - __half2float(n1);
        const __half n2 = __float2half_rn(r1);
        const __half n3 = __float2half_rn(r1 - __half2float(n2));
        base[512 + i] = make_uint4(pack_h2(n1, n2), pack_h2(n3, __ushort_as_half(0)), 0u, 0u);
    } else {
        /* padding rows: zero vector, largest norm term; the epilogue also rejects them by index */
        const __half big = __ushort_as_half(0x7BFF); /* 65504 */
        base[512 + i] = make_uint4(pack_h2(big, big), pack_h2(big, __ushort_as_half(0)), 0u, 0u);
    }
}

/* 16 code bits -> 16 e4m3 bytes: byte i = base with its sign flipped when bit i is set (base 0xB8.. = -1: bit -> +-1;
 * base 0x38.. = +1: bit -> -(+-1)) */
__device__ __forceinline__ uint4 tc_expand16(uint32_t bits, uint32_t base) {
    uint32_t w[4];
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const uint32_t nib = (bits >> (4 * n)) & 0xFu;
        w[n] = base ^ (((nib * 0x00204081u) & 0x01010101u) << 7);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

/* codes: [t][words] u32.  out tile j (128 rows): [2 * words chunks][128 rows][16 bytes].  Rows past t_local repeat the
 * last real row, so the minimum of a partly filled block is that of its real rows; blocks with no real row are
 * skipped by the scan (n_blocks). */
__global__ void tc_pack_codes_kernel(const uint32_t* __restrict__ codes, int words, long long t_local, uint4* __restrict__ out) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long tiles = (t_local + kTcRows - 1) / kTcRows;
    if (r >= tiles * kTcRows) return;
    const long long tile = r / kTcRows;
    const int i = (int)(r - tile * kTcRows);
    const long long src = r < t_local ? r : t_local - 1;
    uint4* base = out + tile * ((size_t)words * 4096 / 16);
    for (int w = 0; w < words; ++w) {
        const uint32_t code = codes[src * words + w];
        base[(2 * w) * 128 + i] = tc_expand16(code & 0xFFFFu, 0xB8B8B8B8u);
        base[(2 * w + 1) * 128 + i] = tc_expand16(code >> 16, 0xB8B8B8B8u);
    }
}

/* ------------------------------------------------------------------- scan ---- */
struct TcArgs {
    const float* q;            /* L2: [nq][q_ld], q_ld = 4 or 16 (zero padded) */
    const uint32_t* qcodes;    /* Hamming: [nq][words] */
    int nq, q_ld, words;
    int n_blocks;              /* 32-template blocks that hold at least one real row */
    const uint4* db;           /* packed tiles */
    long long t_local, row_offset;
    int tiles_total, tiles_per_chunk, n_chunks, n_qpairs;
    int log2_st, log2_sn;      /* s_t = 2^log2_st, s_n = s_t * 2^-c = 2^log2_sn  (norm term scale) */
    int* tau;                  /* [nq] shared admission thresholds (ordered-int floats) */
    const float* ctr;          /* [128] DB mean: queries are centred the same way as the packed templates */
    float* key_scale;          /* [nq] 1 / (s_q s_t): what a key must be multiplied with to become |t'|^2 - 2 q'.t' */
    TcPlan plan;
    float* cand_d;
    int* cand_i;
    int* err;
    unsigned long long* dbg;   /* PE_TC_DEBUG: per-role cycle counters (null in production) */
};
#define TC_DBG_T0(var) long long var = 0; if (DBG) var = clock64()
#define TC_DBG_ADD(slot, var) do { if (DBG) { const long long n_ = clock64(); dbg_acc[slot] += n_ - var; var = n_; } } while (0)

/* insert into the calling thread's k+2 list (entries strided by kTcListCols threads in shared memory): the new entry
 * replaces the worst one - largest key, among equal keys the largest block id, so the earlier (lower-index) block
 * stays as a strict '>' in a sorted insert would have it (KC = 7 for k <= 5, 18 for k <= 16).  The list is NOT kept sorted (the select kernel ranks all
 * candidates anyway): all entries are read at once, one is written, i.e. two shared-memory round trips where the
 * shifting insert made up to a dozen dependent ones (~1,800 cycles per entry under the tensor pipe's operand traffic).
 * The four 32-template blocks of a drained tile go through ONE load / store of the list.  Returns the new k+2-th best =
 * the largest key left.  Deliberately not inlined: it runs in about one warp-tile in six and must not bloat the hot
 * loop's instruction footprint. */
template <int KC>
__device__ __noinline__ float tc_list_insert4(float* ld, int* li, float tau_up, float m0, float m1, float m2, float m3,
                                             int blk0, int n_blocks) {
    float key[KC];
    int id[KC];
#pragma unroll
    for (int p = 0; p < KC; ++p) {
        key[p] = ld[p * kTcListCols];
        id[p] = li[p * kTcListCols];
    }
    float thr = key[0];
#pragma unroll
    for (int p = 1; p < KC; ++p) thr = fmaxf(thr, key[p]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float mj = j == 0 ? m0 : (j == 1 ? m1 : (j == 2 ? m2 : m3));
        if (mj < fminf(thr, tau_up) && blk0 + j < n_blocks) {
            float wk = key[0];
            int wi = id[0], wp = 0;
#pragma unroll
            for (int p = 1; p < KC; ++p) {
                const bool worse = key[p] > wk || (key[p] == wk && id[p] > wi);
                wk = worse ? key[p] : wk;
                wi = worse ? id[p] : wi;
                wp = worse ? p : wp;
            }
            thr = mj;
#pragma unroll
            for (int p = 0; p < KC; ++p) {
                key[p] = p == wp ? mj : key[p];
                id[p] = p == wp ? blk0 + j : id[p];
                thr = fmaxf(thr, key[p]);
            }
        }
    }
#pragma unroll
    for (int p = 0; p < KC; ++p) {
        ld[p * kTcListCols] = key[p];
        li[p * kTcListCols] = id[p];
    }
    return thr;
}

__device__ __forceinline__ float fmin3(float x, float y, float z) { return fminf(fminf(x, y), z); }   /* one FMNMX3 */

/* minimum of 32 accumulator columns in 16 three-input instructions */
__device__ __forceinline__ float tc_min32(const uint32_t (&v)[32]) {
    float t[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) t[i] = fmin3(__uint_as_float(v[3 * i]), __uint_as_float(v[3 * i + 1]), __uint_as_float(v[3 * i + 2]));
    const float u0 = fmin3(t[0], t[1], t[2]), u1 = fmin3(t[3], t[4], t[5]), u2 = fmin3(t[6], t[7], t[8]);
    const float u3 = fmin3(t[9], __uint_as_float(v[30]), __uint_as_float(v[31]));
    return fminf(fmin3(u0, u1, u2), u3);
}

/* FAST: one K chunk per tile and a full ring of 8 stages (descriptors of up to 16 dimensions, all Hamming widths with
 * k <= 5): chunk count and ring size are compile-time constants and both issuers run.  Otherwise (wider descriptors,
 * shorter rings) ONE issuer walks the tiles in order: a parity wait on a ring barrier is only sound while the waiter is at
 * most one phase away from the barrier, which two issuers that work on different tiles cannot guarantee once a tile spans
 * several stages - and with dim / 16 x 3 MMAs per chunk the tensor pipe, not the issue bookkeeping, is the bound there. */
template <int KC, bool DBG, int MODE, bool FAST>
__global__ void __launch_bounds__(kTcThreads, 1) knn_tc_scan_kernel(const TcArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const TcPlan& L = a.plan;
    const int n_stages = FAST ? kTcMaxStages : L.stages, nch = FAST ? 1 : L.nch;
    const bool single = !FAST;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
    uint64_t* full = bars;                     /* [kTcMaxStages] chunk landed */
    uint64_t* empty = bars + kTcMaxStages;     /* [kTcMaxStages] chunk consumed by the MMAs of both query tiles */
    uint64_t* tfull = bars + 2 * kTcMaxStages; /* [4] accumulator ready */
    uint64_t* tempty = tfull + 4;              /* [4] accumulator drained */
    uint64_t* a_ready = tempty + 4;            /* query operands of the next item written */
    uint64_t* item_done = a_ready + 1;         /* every MMA of the item has completed (one commit per issuer) */
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(item_done + 1);
    volatile int* s_abort = reinterpret_cast<volatile int*>(tmem_slot + 1);
    float* s_red = reinterpret_cast<float*>(tmem_slot + 2);   /* [8] query abs-max per operand warp */
    float* s_ld = reinterpret_cast<float*>(smem + L.off_list);          /* [KC][512] candidate keys (KC > 3) */
    int* s_li = reinterpret_cast<int*>(s_ld + KC * kTcListCols);        /* [KC][512] candidate blocks */

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (MODE == 0)
        for (int i = tid; i < 2048 / 16; i += kTcThreads) reinterpret_cast<uint4*>(smem + L.off_zero)[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) {
        for (int s = 0; s < kTcMaxStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 4; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 4); }
        mbar_init(a_ready, 256);
        mbar_init(item_done, single ? 1 : 2);
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
        /* ================= producer: one ring stage per K chunk of a template tile ================= */
        if (lane == 0) {
            uint32_t g = 0;                                   /* stages issued so far */
            const uint32_t ham_bytes = (uint32_t)a.words * 4096u;
            const size_t tile16 = MODE ? ham_bytes / 16 : (size_t)(nch * kTcChunkBytes + 2048) / 16;
            bool ok = true;
            for (int item = blockIdx.x; item < n_items && ok && !*s_abort; item += gridDim.x) {
                const int chunk = item / a.n_qpairs;
                const int t0 = chunk * a.tiles_per_chunk;
                const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
                for (int t = t0; t < t1 && ok; ++t) {
                    for (int c = 0; c < nch; ++c, ++g) {
                        const uint32_t s = g % (uint32_t)n_stages;
                        if (!mbar_wait(&empty[s], ((g / (uint32_t)n_stages) & 1u) ^ 1u)) { atomicExch(a.err, 1); *s_abort = 1; ok = false; break; }
                        const uint32_t bytes = MODE ? ham_bytes : (c == 0 ? (uint32_t)kTcStageBytes : (uint32_t)kTcChunkBytes);
                        const size_t off16 = MODE || c == 0 ? 0 : (kTcStageBytes + (size_t)(c - 1) * kTcChunkBytes) / 16;
                        mbar_expect_tx(&full[s], bytes);
                        bulk_g2s(smem + L.off_b + s * L.stage_bytes, a.db + (size_t)t * tile16 + off16, bytes, &full[s]);
                    }
                }
            }
        }
    } else if (warp <= 2) {
        /* ================= MMA issuers: warp 1 issues the even tiles of the CTA's tile sequence, warp 2 the odd ones,
         * each for both query tiles, so the bookkeeping of one issuer runs under the MMAs of the other.  The whole warp
         * runs the loop (warp-uniform operands), one elected lane issues.  A tile of a D-dim DB is D/16 K chunks; chunk
         * c adds hi.hi + hi.lo + lo.hi of dims 16c..16c+15 (and, with c = 0, the norm term) to BOTH accumulators of the
         * set, then releases its ring stage; the accumulators are committed behind the last chunk. ============ */
        if (!(single && warp == 2)) {
            const int set = warp - 1;
            const uint32_t idesc = instr_desc(128, 128, FMT_F16, FMT_F16);
            const uint32_t sB = smem_u32(smem + L.off_b), sZ = smem_u32(smem + L.off_zero);
            const uint32_t a0 = smem_u32(smem), a1 = a0 + L.a_bytes;
            const uint32_t an_off = (uint32_t)nch * kTcChunkBytes;     /* a_n behind the chunks of a query tile */
            /* L2: hi | lo of a chunk, norm term.  Hamming: K step j of a tile is the two 16-byte chunks at j * 4096 */
            const uint64_t a0_hi = smem_desc(a0, 2048, 128), a0_lo = smem_desc(a0 + 4096, 2048, 128);
            const uint64_t a0_n = smem_desc(a0 + an_off, MODE ? 2048 : sZ - (a0 + an_off), 128);
            const uint64_t a1_hi = smem_desc(a1, 2048, 128), a1_lo = smem_desc(a1 + 4096, 2048, 128);
            const uint64_t a1_n = smem_desc(a1 + an_off, MODE ? 2048 : sZ - (a1 + an_off), 128);
            const uint64_t b_hi0 = smem_desc(sB, 2048, 128), b_lo0 = smem_desc(sB + 4096, 2048, 128);
            const uint64_t b_n0 = smem_desc(sB + 8192, MODE ? 2048 : 0, 128);   /* L2: chunk 1 aliases chunk 0, A_n's chunk 1 is zero */
            uint32_t g = 0, it = 0;                          /* g: tiles of the CTA's sequence so far */
            long long dbg_acc[4] = {0, 0, 0, 0};
            TC_DBG_T0(tm);
            bool ok = true;
            for (int item = blockIdx.x; item < n_items && ok && !*s_abort; item += gridDim.x, ++it) {
                const int chunk = item / a.n_qpairs;
                const int t0 = chunk * a.tiles_per_chunk;
                const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
                if (!mbar_wait(a_ready, it & 1)) { if (lane == 0) atomicExch(a.err, 2); ok = false; break; }
                fence_after_sync();
                const uint32_t g_end = g + (uint32_t)(t1 - t0);
#pragma unroll 1
                for (uint32_t gi = single ? g : g + ((g ^ (uint32_t)set) & 1u); gi < g_end && ok; gi += single ? 1 : 2) {
                    const uint32_t ph = (gi >> 1) & 1u;
                    const uint32_t aset = single ? (gi & 1u) : (uint32_t)set;       /* accumulator pair = tile parity */
                    const uint32_t d0 = tmem + (aset * 2) * 128, d1 = d0 + 128;
                    uint64_t* const tf0 = &tfull[aset * 2], * const te0 = &tempty[aset * 2];
#pragma unroll 1
                    for (int c = 0; c < nch; ++c) {
                        const uint32_t gc = gi * (uint32_t)nch + (uint32_t)c;
                        const uint32_t s = gc % (uint32_t)n_stages;
                        const bool first = c == 0, last = c == nch - 1;
                        TC_DBG_ADD(2, tm);
                        if (!mbar_wait(&full[s], (gc / (uint32_t)n_stages) & 1u)) { if (lane == 0) atomicExch(a.err, 3); ok = false; break; }
                        TC_DBG_ADD(0, tm);
                        /* the two accumulators of the set are handed back half a tile apart: issue for each as soon as it is
                         * free instead of waiting for both */
                        if (first && !mbar_wait(te0, ph ^ 1u)) { if (lane == 0) atomicExch(a.err, 4); ok = false; break; }
                        fence_after_sync();
                        TC_DBG_ADD(1, tm);
                        const uint32_t off = s * ((uint32_t)L.stage_bytes >> 4);
                        const uint32_t ao = (uint32_t)c * (kTcChunkBytes >> 4);
                        if (elect_one()) {
                            if (MODE == 0) {
                                mma_f16(d0, a0_hi + ao, b_hi0 + off, idesc, first ? 0u : 1u);
                                mma_f16(d0, a0_hi + ao, b_lo0 + off, idesc, 1);
                                mma_f16(d0, a0_lo + ao, b_hi0 + off, idesc, 1);
                                if (first) mma_f16(d0, a0_n, b_n0 + off, idesc, 1);
                            } else {
                                /* -dot of the +-1 expansions, 32 bits per MMA: R = 2 * hamming - bits */
                                for (int j = 0; j < a.words; ++j) mma_f8(d0, a0_hi + j * 256, b_hi0 + off + j * 256, idesc, j > 0);
                            }
                            if (last) mma_commit(tf0);
                        }
                        __syncwarp();
                        TC_DBG_ADD(2, tm);
                        if (first && !mbar_wait(te0 + 1, ph ^ 1u)) { if (lane == 0) atomicExch(a.err, 4); ok = false; break; }
                        fence_after_sync();
                        TC_DBG_ADD(1, tm);
                        if (elect_one()) {
                            if (MODE == 0) {
                                mma_f16(d1, a1_hi + ao, b_hi0 + off, idesc, first ? 0u : 1u);
                                mma_f16(d1, a1_hi + ao, b_lo0 + off, idesc, 1);
                                mma_f16(d1, a1_lo + ao, b_hi0 + off, idesc, 1);
                                if (first) mma_f16(d1, a1_n, b_n0 + off, idesc, 1);
                            } else {
                                for (int j = 0; j < a.words; ++j) mma_f8(d1, a1_hi + j * 256, b_hi0 + off + j * 256, idesc, j > 0);
                            }
                            if (last) mma_commit(tf0 + 1);
                            mma_commit(&empty[s]);
                        }
                        __syncwarp();
                    }
                }
                if (!ok) break;
                /* the operand warps may rewrite the query tiles once both issuers' MMAs of this item are done */
                if (elect_one()) mma_commit(item_done);
                __syncwarp();
                g = g_end;
            }
            if (!ok) *s_abort = 1;
            if (DBG && lane == 0) {
                TC_DBG_ADD(2, tm);
                atomicAdd(a.dbg + 5, (unsigned long long)dbg_acc[0]);
                atomicAdd(a.dbg + 6, (unsigned long long)dbg_acc[1]);
                atomicAdd(a.dbg + 7, (unsigned long long)dbg_acc[2]);
                if (set == 0) atomicAdd(a.dbg + 8, (unsigned long long)g);
            }
        }
    } else {
        /* ================= epilogue: 16 warps = 2 query tiles x 2 tile sets x 4 lane quarters.  Set 0 drains the
         * even tiles of the CTA's tile sequence, set 1 the odd ones (each into its own list), so a warp has two tile
         * periods for one accumulator and four warps per scheduler hide each other's TMEM and barrier latency. ===== */
        const int ew = warp - 3;
        const int quarter = warp & 3;                 /* TMEM lanes this warp may read: 32 * (warp % 4) */
        const int qt = (ew >> 2) & 1;                 /* query tile of the pair */
        const int set = ew >> 3;                      /* parity of the tiles this warp drains */
        const int row = quarter * 32 + lane;          /* query row inside the query tile */
        const int et = ew * 32 + lane;                /* 0..511: this thread's column in the shared-memory lists */
        float* my_ld = s_ld + et;
        int* my_li = s_li + et;
        uint64_t* const my_tfull = &tfull[set * 2 + qt];
        uint64_t* const my_tempty = &tempty[set * 2 + qt];
        const uint32_t tbase = tmem + (set * 2 + qt) * 128 + (static_cast<uint32_t>(quarter * 32) << 16);
        uint32_t g = 0, n_item = 0;
        long long dbg_acc[5] = {0, 0, 0, 0, 0};
        TC_DBG_T0(te);
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int chunk = item / a.n_qpairs, qpair = item - chunk * a.n_qpairs;
            const int t0 = chunk * a.tiles_per_chunk;
            const int t1 = min(t0 + a.tiles_per_chunk, a.tiles_total);
            const int qidx = (qpair * 2 + qt) * kTcRows + row;
            const bool q_ok = qidx < a.nq;
            if (set == 0 && MODE == 0) {
                /* ---- query operands (256 threads): centre, scale, split, store in the K-major core-matrix layout.  Two
                 * passes over the query row (it stays in L1): the tile's abs-max first, then one 16-dim chunk at a time */
                const float4* q4 = reinterpret_cast<const float4*>(a.q + (size_t)(q_ok ? qidx : 0) * a.q_ld);
                const float4* c4 = reinterpret_cast<const float4*>(a.ctr);
                const int n4 = a.q_ld >> 2;
                float m = 0.f;
                if (q_ok)
                    for (int d4 = 0; d4 < n4; ++d4) {
                        const float4 v = __ldg(q4 + d4), c = __ldg(c4 + d4);
                        m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x - c.x), fabsf(v.y - c.y)), fmaxf(fabsf(v.z - c.z), fabsf(v.w - c.w))));
                    }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                if (lane == 0) s_red[qt * 4 + quarter] = m;
                asm volatile("bar.sync 1, 256;" ::: "memory");
                m = fmaxf(fmaxf(s_red[qt * 4 + 0], s_red[qt * 4 + 1]), fmaxf(s_red[qt * 4 + 2], s_red[qt * 4 + 3]));
                /* s_q = 2^eq with 2 * max|q| * s_q in [2^10, 2^11); a_n = s_q * s_t / s_n must stay a finite fp16 power of two */
                int eq = (m > 0.f && isfinite(m)) ? 9 - ilogbf(m) : 0;
                int ea = eq + a.log2_st - a.log2_sn;
                if (ea > 15) { eq -= ea - 15; ea = 15; }
                const float s_q = ldexpf(1.f, eq);
                const __half an = ea < -24 ? __ushort_as_half(0) : __float2half_rn(ldexpf(1.f, ea));
                if (q_ok && chunk == 0) a.key_scale[qidx] = ldexpf(1.f, -(eq + a.log2_st));
                uint4* op = reinterpret_cast<uint4*>(smem + qt * L.a_bytes);
#pragma unroll 1
                for (int c = 0; c < nch; ++c) {
                    float qv[16];
#pragma unroll
                    for (int d4 = 0; d4 < 4; ++d4) {
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (q_ok && 4 * c + d4 < n4) {
                            const float4 cc = __ldg(c4 + 4 * c + d4);
                            v = __ldg(q4 + 4 * c + d4);
                            v.x -= cc.x; v.y -= cc.y; v.z -= cc.z; v.w -= cc.w;
                        }
                        qv[4 * d4 + 0] = v.x; qv[4 * d4 + 1] = v.y; qv[4 * d4 + 2] = v.z; qv[4 * d4 + 3] = v.w;
                    }
                    uint32_t hi[8], lo[8];
#pragma unroll
                    for (int d = 0; d < 16; d += 2) {
                        __half h0, l0, h1, l1;
                        split_f16(-2.f * s_q * qv[d], h0, l0);
                        split_f16(-2.f * s_q * qv[d + 1], h1, l1);
                        hi[d / 2] = pack_h2(h0, h1);
                        lo[d / 2] = pack_h2(l0, l1);
                    }
                    uint4* oc = op + c * (kTcChunkBytes / 16);
                    oc[0 + row] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    oc[128 + row] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                    oc[256 + row] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    oc[384 + row] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                }
                op[nch * (kTcChunkBytes / 16) + row] = make_uint4(pack_h2(an, an), pack_h2(an, __ushort_as_half(0)), 0u, 0u);
                fence_async_smem();
                mbar_arrive(a_ready);
                asm volatile("bar.sync 1, 256;" ::: "memory");   /* s_red may be rewritten by the next item */
            }
            if (set == 0 && MODE == 1) {
                /* ---- query operands, Hamming: bit b of the code -> -(+-1) as e4m3 (0x38 = +1, 0xB8 = -1), 16 bits per
                 * 16-byte core-matrix row, K chunk c = bits 16c .. 16c+15 */
                uint4* op = reinterpret_cast<uint4*>(smem + qt * L.a_bytes);
                for (int w = 0; w < a.words; ++w) {
                    const uint32_t code = q_ok ? __ldg(a.qcodes + (size_t)qidx * a.words + w) : 0u;
                    op[(2 * w) * 128 + row] = tc_expand16(code & 0xFFFFu, 0x38383838u);
                    op[(2 * w + 1) * 128 + row] = tc_expand16(code >> 16, 0x38383838u);
                }
                fence_async_smem();
                mbar_arrive(a_ready);
            }
            /* the k+2 best blocks of this thread's tiles: in registers for k = 1, in shared memory otherwise */
            float rk[KC <= 3 ? KC : 1];
            int ri[KC <= 3 ? KC : 1];
            if (KC <= 3) {
#pragma unroll
                for (int p = 0; p < (KC <= 3 ? KC : 1); ++p) { rk[p] = INFINITY; ri[p] = -1; }
            } else {
#pragma unroll
                for (int p = 0; p < KC; ++p) { my_ld[p * kTcListCols] = INFINITY; my_li[p * kTcListCols] = -1; }
            }

            float thr = INFINITY;                              /* k+2-th best of this list so far */
            /* the warps that scan other tiles for this query publish their k+2-th best in tau: nothing above it can
             * reach the final top-k.  Admission is v <= tau, i.e. v < nextup(tau).  tau is re-read every other
             * drained tile and used from the following one on (a stale tau is only more permissive). */
            int* const tau_p = a.tau + (q_ok ? qidx : 0);
            int tau_raw = q_ok ? __ldcg(tau_p) : 0x7F7F7F7F;
            float tau_up = INFINITY;
            const uint32_t g_end = g + (uint32_t)(t1 - t0);
            const uint32_t g_first = g + ((g ^ (uint32_t)set) & 1u);               /* first tile of this set */
            const int n_mine = g_first < g_end ? (int)((g_end - g_first + 1) >> 1) : 0;
            int blk0 = (t0 + (int)(g_first - g)) * (kTcRows / kTcBlock);            /* local block id of columns 0..31 */
            uint32_t ph = (g_first >> 1) & 1u;
#pragma unroll 1
            for (int i = 0; i < n_mine; ++i, blk0 += 2 * (kTcRows / kTcBlock), ph ^= 1u) {
                if (i & 1) {
                    const float tg = ord2f(tau_raw);
                    tau_up = __int_as_float(tg >= 0.f ? __float_as_int(tg) + 1 : __float_as_int(tg) - 1);
                    if (tg == 0.f) tau_up = FLT_MIN;
                } else if (q_ok) {
                    tau_raw = __ldcg(tau_p);
                }
                float th = fminf(thr, tau_up);
                TC_DBG_ADD(3, te);
                if (!mbar_wait(my_tfull, ph)) { atomicExch(a.err, 5); *s_abort = 1; break; }
                fence_after_sync();
                TC_DBG_ADD(0, te);
                uint32_t v0[32], v1[32];
                tmem_ld32(tbase, v0);
                tmem_ld32(tbase + 32, v1);
                tmem_ld_wait();
                const float m0 = tc_min32(v0), m1 = tc_min32(v1);
                tmem_ld32(tbase + 64, v0);
                tmem_ld32(tbase + 96, v1);
                tmem_ld_wait();
                fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(my_tempty);     /* the accumulator goes back before anything else happens */
                const float m2 = tc_min32(v0), m3 = tc_min32(v1);
                const bool hit = fminf(fminf(m0, m1), fminf(m2, m3)) < th;
                TC_DBG_ADD(1, te);
                if (__any_sync(0xffffffffu, hit)) {
                    /* rare path: a 32-template block enters the list with its minimum as the key; which template of
                     * the block it was is found by the exact pass over the block (knn_tc_select_kernel) */
                    if (DBG) dbg_acc[4] += 1;
                    if (hit) {
                        if (KC <= 3) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const float mj = j == 0 ? m0 : (j == 1 ? m1 : (j == 2 ? m2 : m3));
                                if (mj < th && blk0 + j < a.n_blocks) {
                                    /* sorted insert, strict '>' keeps the earlier block first among equals */
#pragma unroll
                                    for (int p = (KC <= 3 ? KC : 1) - 1; p > 0; --p) {
                                        const bool shift = rk[p - 1] > mj;
                                        const bool here = !shift && rk[p] > mj;
                                        ri[p] = shift ? ri[p - 1] : (here ? blk0 + j : ri[p]);
                                        rk[p] = shift ? rk[p - 1] : (here ? mj : rk[p]);
                                    }
                                    if (rk[0] > mj) { rk[0] = mj; ri[0] = blk0 + j; }
                                    thr = rk[(KC <= 3 ? KC : 1) - 1];
                                    th = fminf(thr, tau_up);
                                }
                            }
                        } else {
                            thr = tc_list_insert4<KC>(my_ld, my_li, tau_up, m0, m1, m2, m3, blk0, a.n_blocks);
                            th = fminf(thr, tau_up);
                        }
                        if (thr < INFINITY && q_ok) atomicMin(tau_p, f2ord(thr));
                    }
                    __syncwarp();
                    TC_DBG_ADD(2, te);
                }
            }
            /* the item's MMAs must have completed before the operand warps (set 0) overwrite the query tiles.  A parity
             * wait on the other set's accumulator barrier would not do: that issuer may be more than one phase behind. */
            if (set == 0 && !mbar_wait(item_done, n_item & 1u)) { atomicExch(a.err, 6); *s_abort = 1; }
            ++n_item;
            g = g_end;
            /* ---- this item's candidates: [query][chunk][set][KC] (key, local 32-template block id or -1) */
            if (q_ok) {
                const size_t o = (((size_t)qidx * a.n_chunks + chunk) * 2 + set) * KC;
                if (KC <= 3) {
#pragma unroll
                    for (int p = 0; p < (KC <= 3 ? KC : 1); ++p) { a.cand_d[o + p] = rk[p]; a.cand_i[o + p] = ri[p]; }
                } else {
#pragma unroll
                    for (int p = 0; p < KC; ++p) {
                        a.cand_d[o + p] = my_ld[p * kTcListCols];
                        a.cand_i[o + p] = my_li[p * kTcListCols];
                    }
                }
            }
        }
        if (DBG && lane == 0) {
            TC_DBG_ADD(3, te);
            for (int j = 0; j < 5; ++j) atomicAdd(a.dbg + j, (unsigned long long)dbg_acc[j]);
            atomicAdd(a.dbg + 9, (unsigned long long)g);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

/* ------------------------------------------------------------------ select ---- */
/* One warp per query.  The scan left n_lists x kc (key, block) candidates; the n_sel best blocks by
 * (key, block id) are evaluated exactly -- the reference's fp32 sqrt(sum (q-t)^2) for each of their 32 templates --
 * and the k smallest by (distance, index) are the answer, as BFMatcher::knnMatch orders them
 * (reference src/networkevaluator.cpp:197; ties -> lower index).
 *
 * Certificate (flag != null).  Every block outside the re-ranked ones has a scan key >= B:
 *   - a candidate that was not re-ranked has a key >= the (n_sel + 1)-th smallest candidate key;
 *   - a block that is in no list was rejected or evicted while some list of this query held kc keys <= its own (own
 *     list's threshold, or the shared threshold tau = another list's kc-th best), and list entries are only ever
 *     replaced by smaller keys: its key is >= K*, the kc-th smallest candidate key.
 * A scan key is s_q s_t (|t'|^2 - 2 q'.t') (primes: centred on the DB mean) up to the error of the arithmetic: each
 * operand is a two-term fp16 split (residual <= 2^-22 relative), the lo.lo product is dropped, the tensor core
 * accumulates the <= 8 x 3 + 1 products of a tile with truncation (<= 2^-23 of the running magnitude each), the
 * three-term norm split is exact to 2^-33:  |key / (s_q s_t) - R| <= 2^-17 |q'| T + 2^-19 T^2  with T = max |t'| over
 * the DB (a factor >= 4 above the worst case of the terms just listed).  So every template that was not re-ranked has
 * d^2 >= |q'|^2 + B / (s_q s_t) - eps, and the reported k-th neighbour is proven whenever its exact d_k^2 lies below
 * that.  Queries that cannot be proven (near-ties at the list boundary, or a DB whose spread T^2 is > ~1e5 x the
 * neighbour distances, e.g. tight far-apart clusters) are appended to flag[] and re-scanned with the direct expression. */
constexpr int kTcSelWarps = 4;               /* queries (warps) per CTA of the L2 select kernel */
__global__ void __launch_bounds__(32 * kTcSelWarps)
knn_tc_select_kernel(const float* __restrict__ cand_d, const int* __restrict__ cand_i, int n_cand, int kc,
                     const float* __restrict__ q, const float* __restrict__ db, int dp, long long t_local,
                     long long row_offset, int k, int n_sel, long long nq, const float* __restrict__ key_scale,
                     const float* __restrict__ ctr, float max_norm2, float eps_rel, int* __restrict__ flag,
                     int* __restrict__ out_idx, float* __restrict__ out_dist) {
    extern __shared__ __align__(16) float sel_smem_all[];
    /* one warp per query, kTcSelWarps independent queries per CTA (32-thread CTAs cap an SM at 32 resident warps, and
     * this kernel is a chain of dependent reductions and random row reads: it lives on resident warps) */
    float* sel_smem = sel_smem_all + (size_t)(threadIdx.x >> 5) * (dp + 2 * n_cand + 2 * n_sel * kTcBlock);
    float* sq = sel_smem;                                   /* [dp] */
    float* sd = sel_smem + dp;                              /* [n_cand] keys */
    int* si = reinterpret_cast<int*>(sd + n_cand);          /* [n_cand] block ids */
    float* ed = reinterpret_cast<float*>(si + n_cand);      /* [n_sel * 32] exact distances */
    int* ei = reinterpret_cast<int*>(ed + n_sel * kTcBlock);
    const int lane = threadIdx.x & 31;
    const long long qi = (long long)blockIdx.x * kTcSelWarps + (threadIdx.x >> 5);
    if (qi >= nq) return;
    float qn2 = 0.f;                                        /* |q - c|^2 */
    for (int d = lane; d < dp; d += 32) {
        const float v = q[(size_t)qi * dp + d];
        sq[d] = v;
        if (flag) { const float c = v - ctr[d]; qn2 = fmaf(c, c, qn2); }
    }
    for (int c = lane; c < n_cand; c += 32) {
        const int blk = cand_i[(size_t)qi * n_cand + c];
        sd[c] = blk >= 0 ? cand_d[(size_t)qi * n_cand + c] : INFINITY;
        si[c] = blk >= 0 ? blk : INT_MAX;
    }
    for (int c = lane; c < n_sel * kTcBlock; c += 32) { ed[c] = INFINITY; ei[c] = INT_MAX; }
    __syncwarp();
    /* extraction in ascending (key, id) order: rounds 0 .. n_sel - 1 are re-ranked, the certificate also wants the
     * kc-th smallest key (K*) and the first key that is NOT re-ranked */
    float key_kstar = INFINITY, key_next = INFINITY;
    int my_blk = 0, n_rerank = 0;                           /* lane r keeps the block of round r (n_sel <= 18) */
    const int rounds = flag ? (n_sel + 1 > kc ? n_sel + 1 : kc) : n_sel;
    for (int r = 0; r < rounds; ++r) {
        float bd = INFINITY;
        int bi = INT_MAX, bc = -1;
        for (int c = lane; c < n_cand; c += 32) {
            const float d = sd[c];
            const int i = si[c];
            if (i != INT_MAX && (d < bd || (d == bd && i < bi) || bc < 0)) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bc >= 0);     /* lane holding the smallest (key, id), -1: none */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = __shfl_sync(0xffffffffu, bi, w < 0 ? 0 : w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        if (bc < 0) break;                                  /* no candidate left (warp-uniform) */
        if (r == kc - 1) key_kstar = bd;
        if (r == n_sel) key_next = bd;
        if (r < n_sel) {
            /* the block's 32 rows are re-ranked after the extraction rounds; their lines are requested now, so the
             * n_sel gathers (random rows of the DB) overlap instead of queueing behind one another */
            if (lane == r) my_blk = bi;
            const long long row = (long long)bi * kTcBlock + lane;
            if (row < t_local) {
                const char* p = reinterpret_cast<const char*>(db + (size_t)row * dp);
                for (int b = 0; b < dp * 4; b += 128) asm volatile("prefetch.L2 [%0];" ::"l"(p + b));
            }
            n_rerank = r + 1;
        }
        __syncwarp();
        if (lane == 0) si[bc] = INT_MAX;
        __syncwarp();
        if (r >= n_sel && r >= kc - 1) break;
    }
    for (int r = 0; r < n_rerank; ++r) {
        const int blk = __shfl_sync(0xffffffffu, my_blk, r);
        const long long row = (long long)blk * kTcBlock + lane;
        if (row < t_local) {
            ed[r * kTcBlock + lane] = exact_l2(sq, db + (size_t)row * dp, dp);
            ei[r * kTcBlock + lane] = (int)row;
        }
    }
    __syncwarp();
    float d_kth = INFINITY;
    bool have_k = true;
    for (int r = 0; r < k; ++r) {
        float bd = INFINITY;
        int bi = INT_MAX, bc = -1;
        for (int c = lane; c < n_sel * kTcBlock; c += 32) {
            const float d = ed[c];
            const int i = ei[c];
            if (i != INT_MAX && (bc < 0 || d < bd || (d == bd && i < bi))) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bc >= 0);     /* lane holding the smallest (key, id), -1: none */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = __shfl_sync(0xffffffffu, bi, w < 0 ? 0 : w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        if (bc >= 0) d_kth = bd; else have_k = false;
        if (lane == 0) {
            out_idx[(size_t)qi * k + r] = bc >= 0 ? (int)(bi + row_offset) : -1;
            out_dist[(size_t)qi * k + r] = bc >= 0 ? bd : INFINITY;
            if (bc >= 0) ei[bc] = INT_MAX;
        }
        __syncwarp();
    }
    if (flag) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) qn2 += __shfl_xor_sync(0xffffffffu, qn2, off);
        /* B: lower bound of the scan keys of everything that was not re-ranked (see the comment above) */
        const float b_key = n_sel >= kc ? key_kstar : key_next;
        bool proven = true;
        if (b_key < INFINITY) {
            const float qn = sqrtf(qn2), tmax = sqrtf(max_norm2);
            const float eps = eps_rel * qn * tmax + 1.9073486328125e-6f * (max_norm2 + qn2);    /* eps_rel (2^-17), 2^-19 */
            const float lb = qn2 + b_key * key_scale[qi] - eps;          /* d^2 of any template that was not re-ranked */
            const float dk2 = d_kth * d_kth;
            proven = have_k && dk2 * 1.000002f < lb;
        }
        if (!proven && lane == 0) {
            const int slot = atomicAdd(flag, 1);
            flag[4 + slot] = (int)qi;
        }
    }
}

/* Hamming twin of knn_tc_select_kernel: exact XOR + POPC distance over the 32 codes of each of the best blocks.  With
 * integer keys the block order (key, id) is exact, and a block holding one of the k nearest (ties -> lower row) is
 * preceded by at most k - 1 others, so the first n_sel >= k blocks contain the answer (no certificate needed). */
__global__ void __launch_bounds__(32)
knn_tc_select_hamming_kernel(const float* __restrict__ cand_d, const int* __restrict__ cand_i, int n_cand,
                             const uint32_t* __restrict__ q, const uint32_t* __restrict__ codes, int words, long long t_local,
                             long long row_offset, int k, int n_sel, int* __restrict__ out_idx, int* __restrict__ out_dist) {
    extern __shared__ __align__(16) float sel_smem[];
    float* sd = sel_smem;                                   /* [n_cand] keys */
    int* si = reinterpret_cast<int*>(sd + n_cand);          /* [n_cand] block ids */
    int* ed = si + n_cand;                                  /* [n_sel * 32] exact distances */
    int* ei = ed + n_sel * kTcBlock;
    const int lane = threadIdx.x;
    const long long qi = blockIdx.x;
    uint32_t qc[8];
    for (int w = 0; w < words; ++w) qc[w] = q[(size_t)qi * words + w];
    for (int c = lane; c < n_cand; c += 32) {
        const int blk = cand_i[(size_t)qi * n_cand + c];
        sd[c] = blk >= 0 ? cand_d[(size_t)qi * n_cand + c] : INFINITY;
        si[c] = blk >= 0 ? blk : INT_MAX;
    }
    __syncwarp();
    for (int r = 0; r < n_sel; ++r) {
        float bd = INFINITY;
        int bi = INT_MAX, bc = -1;
        for (int c = lane; c < n_cand; c += 32) {
            const float d = sd[c];
            const int i = si[c];
            if (i != INT_MAX && (bc < 0 || d < bd || (d == bd && i < bi))) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bc >= 0);     /* lane holding the smallest (key, id), -1: none */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = __shfl_sync(0xffffffffu, bi, w < 0 ? 0 : w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        int d = INT_MAX, idx = INT_MAX;
        if (bc >= 0) {
            const long long row = (long long)bi * kTcBlock + lane;
            if (row < t_local) {
                d = 0;
                for (int w = 0; w < words; ++w) d += __popc(qc[w] ^ __ldg(codes + (size_t)row * words + w));
                idx = (int)row;
            }
        }
        ed[r * kTcBlock + lane] = d;
        ei[r * kTcBlock + lane] = idx;
        __syncwarp();
        if (lane == 0 && bc >= 0) si[bc] = INT_MAX;
        __syncwarp();
    }
    for (int r = 0; r < k; ++r) {
        int bd = INT_MAX, bi = INT_MAX, bc = -1;
        for (int c = lane; c < n_sel * kTcBlock; c += 32) {
            const int d = ed[c], i = ei[c];
            if (i != INT_MAX && (bc < 0 || d < bd || (d == bd && i < bi))) { bd = d; bi = i; bc = c; }
        }
        {
            const int w = warp_argmin(bd, bi, bc >= 0);     /* lane holding the smallest (key, id), -1: none */
            bd = __shfl_sync(0xffffffffu, bd, w < 0 ? 0 : w);
            bi = __shfl_sync(0xffffffffu, bi, w < 0 ? 0 : w);
            bc = w < 0 ? -1 : __shfl_sync(0xffffffffu, bc, w);
        }
        if (lane == 0) {
            out_idx[(size_t)qi * k + r] = bc >= 0 ? (int)(bi + row_offset) : -1;
            out_dist[(size_t)qi * k + r] = bc >= 0 ? bd : INT_MAX;
            if (bc >= 0) ei[bc] = INT_MAX;
        }
        __syncwarp();
    }
}

}  // namespace

/* packs the (already padded, centred-stats-known) fp32 DB of `db` into tensor tiles; called once per DB */
int knn_tc_prepare(pe_db* db) {
    db->tc_ready = false;
    if (db->t_local <= 0) return PE_OK;
    if (db->hamming) {
        /* +-1 (e4m3) tiles of the packed codes: 32, 64 or 128 bits -> 1, 2 or 4 K steps of 32 */
        if (db->words != 1 && db->words != 2 && db->words != 4) return PE_OK;
        PE_CUDA(cudaSetDevice(db->device));
        const long long tiles = (db->t_local + kTcRows - 1) / kTcRows;
        PE_CUDA(cudaMalloc(&db->d_tc, (size_t)tiles * db->words * 4096));
        tc_pack_codes_kernel<<<(unsigned)((tiles * kTcRows + 255) / 256), 256, 0, db->stream>>>(
            db->d_codes, db->words, db->t_local, reinterpret_cast<uint4*>(db->d_tc));
        PE_LAUNCHED();
        PE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&db->d_tc_err), sizeof(int), cudaHostAllocMapped));
        *db->d_tc_err = 0;
        PE_CUDA(cudaStreamSynchronize(db->stream));
        db->tc_ready = true;
        return PE_OK;
    }
    const int ld = db->dim_pad;
    if (ld != 4 && ld != 16 && ld != 64 && ld != 128) return PE_OK;
    if (!db->d_center) return PE_OK;                        /* knn_center_stats did not run */
    PE_CUDA(cudaSetDevice(db->device));
    const float mt = db->max_abs_c, mn = db->max_norm2_c;
    /* s_t: max |s_t (t-c)| in [2^10, 2^11);  s_n: max s_n |t-c|^2 in [2^14, 2^15) */
    db->tc_log2_st = mt > 0.f ? 10 - ilogbf(mt) : 0;
    db->tc_log2_sn = mn > 0.f ? 14 - ilogbf(mn) : 0;
    const int nch = ld > 16 ? ld / 16 : 1;
    const long long tiles = (db->t_local + kTcRows - 1) / kTcRows;
    PE_CUDA(cudaMalloc(&db->d_tc, (size_t)tiles * ((size_t)nch * kTcChunkBytes + 2048)));
    tc_pack_db_kernel<<<(unsigned)((tiles * kTcRows + 255) / 256), 256, 0, db->stream>>>(
        db->d_desc, ld, nch, db->t_local, db->d_center, ldexpf(1.f, db->tc_log2_st), ldexpf(1.f, db->tc_log2_sn),
        reinterpret_cast<uint4*>(db->d_tc));
    PE_LAUNCHED();
    /* pinned, device-visible flag: the kernels write it only when a pipeline barrier times out */
    PE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&db->d_tc_err), sizeof(int), cudaHostAllocMapped));
    *db->d_tc_err = 0;
    PE_CUDA(cudaStreamSynchronize(db->stream));
    db->tc_ready = true;
    return PE_OK;
}

/* the tensor scan serves every (dim, k <= 16) whose shared-memory plan leaves at least two ring stages */
bool knn_tc_usable(const pe_db* db, int k) {
    if (!db->tc_ready || k < 1 || k > 16) return false;
    TcPlan p;
    return tc_make_plan(db->hamming, db->dim_pad, tc_kc_for(k), &p);
}

/* plans the chunking; returns the number of candidate lists per query (two per chunk: even and odd tiles) and kc */
void knn_tc_plan(const pe_db* db, int64_t nq, int k, int* n_lists, int* kc, int* tiles_per_chunk) {
    const int n_qpairs = (int)((nq + 2 * kTcRows - 1) / (2 * kTcRows));
    const long long tiles = (db->t_local + kTcRows - 1) / kTcRows;
    /* items = chunks x query pairs are dealt round-robin to num_sms persistent CTAs: pick the chunk count whose makespan
     * ceil(items / CTAs) x (tiles per chunk + a fixed per-item cost of ~8 tile times) is smallest, among 2..8 items per CTA
     * (the first guess, 4 items per CTA rounded up, left a fifth round with a few items at 2 GPUs: 5 rounds for 4.1) */
    const long long cmax = tiles / 32 > 0 ? tiles / 32 : 1;
    long long c = 1, best = -1;
    for (long long cc = 1; cc <= cmax; ++cc) {
        const long long items = cc * n_qpairs;
        if (items > 8LL * db->num_sms && cc > 1) break;
        const long long rounds = (items + db->num_sms - 1) / db->num_sms;
        const long long cost = rounds * ((tiles + cc - 1) / cc + 8);
        if (best < 0 || cost < best) { best = cost; c = cc; }
    }
    const long long tpc = (tiles + c - 1) / c;
    *tiles_per_chunk = (int)tpc;
    *n_lists = 2 * (int)((tiles + tpc - 1) / tpc);
    *kc = tc_kc_for(k);
}

int knn_tc_scan(pe_db* db, const void* d_q, int64_t nq, int k, int n_lists, int kc, int tiles_per_chunk,
                float* cand_d, int* cand_i) {
    PE_CHECK(db->tau.reserve((size_t)nq * sizeof(int)));
    PE_CHECK(db->key_scale.reserve((size_t)nq * sizeof(float)));
    PE_CUDA(cudaMemsetAsync(db->tau.p, 0x7f, (size_t)nq * sizeof(int), db->stream));
    TcArgs a;
    PE_REQUIRE(tc_make_plan(db->hamming, db->dim_pad, kc, &a.plan), "tensor scan: no shared-memory plan for dim %d, k + 2 = %d", db->dim_pad, kc);
    a.q = db->hamming ? nullptr : static_cast<const float*>(d_q);
    a.qcodes = db->hamming ? static_cast<const uint32_t*>(d_q) : nullptr;
    a.nq = (int)nq;
    a.q_ld = db->dim_pad;
    a.words = db->words;
    a.n_blocks = (int)((db->t_local + kTcBlock - 1) / kTcBlock);
    a.db = reinterpret_cast<const uint4*>(db->d_tc);
    a.t_local = db->t_local;
    a.row_offset = db->row_offset;
    a.tiles_total = (int)((db->t_local + kTcRows - 1) / kTcRows);
    a.tiles_per_chunk = tiles_per_chunk;
    a.n_chunks = n_lists / 2;
    a.n_qpairs = (int)((nq + 2 * kTcRows - 1) / (2 * kTcRows));
    a.log2_st = db->tc_log2_st;
    a.log2_sn = db->tc_log2_sn;
    a.tau = db->tau.as<int>();
    a.ctr = db->d_center;
    a.key_scale = db->key_scale.as<float>();
    a.cand_d = cand_d;
    a.cand_i = cand_i;
    a.err = db->d_tc_err;
    a.dbg = nullptr;
    static const bool dbg_on = getenv("PE_TC_DEBUG") != nullptr;
    unsigned long long* d_dbg = nullptr;
    if (dbg_on && !db->hamming) {
        PE_CUDA(cudaMalloc(&d_dbg, 16 * sizeof(unsigned long long)));
        PE_CUDA(cudaMemsetAsync(d_dbg, 0, 16 * sizeof(unsigned long long), db->stream));
        a.dbg = d_dbg;
    }
    const int items = a.n_chunks * a.n_qpairs;
    const int grid = items < db->num_sms ? items : db->num_sms;
    const int smem_bytes = a.plan.total;
    const bool fast = a.plan.nch == 1 && a.plan.stages == kTcMaxStages;
#define TC_LAUNCH2(KCV, DBGV, MODEV, FASTV) do { \
        PE_CUDA(cudaFuncSetAttribute(knn_tc_scan_kernel<KCV, DBGV, MODEV, FASTV>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemMax)); \
        knn_tc_scan_kernel<KCV, DBGV, MODEV, FASTV><<<grid, kTcThreads, smem_bytes, db->stream>>>(a); } while (0)
#define TC_LAUNCH(KCV, DBGV, MODEV) do { if (fast) TC_LAUNCH2(KCV, DBGV, MODEV, true); else TC_LAUNCH2(KCV, DBGV, MODEV, false); } while (0)
    if (db->hamming) {
        if (kc == 3) TC_LAUNCH(3, false, 1); else if (kc == 7) TC_LAUNCH(7, false, 1); else TC_LAUNCH(18, false, 1);
    } else if (kc == 3) { if (dbg_on) TC_LAUNCH(3, true, 0); else TC_LAUNCH(3, false, 0); }
    else if (kc == 7) { if (dbg_on) TC_LAUNCH(7, true, 0); else TC_LAUNCH(7, false, 0); }
    else TC_LAUNCH(18, false, 0);
#undef TC_LAUNCH2
#undef TC_LAUNCH
    PE_LAUNCHED();
    if (dbg_on && !db->hamming && a.dbg) {
        unsigned long long h[16];
        PE_CUDA(cudaMemcpyAsync(h, d_dbg, sizeof(h), cudaMemcpyDeviceToHost, db->stream));
        PE_CUDA(cudaStreamSynchronize(db->stream));
        cudaFree(d_dbg);
        const double ew = 16.0 * grid, tiles_e = h[9] > 0 ? (double)h[9] : 1.0, tiles_m = h[8] > 0 ? (double)h[8] : 1.0;
        fprintf(stderr, "[tc dbg] grid %d items %d  epilogue per warp-tile: wait_tfull %.0f hot %.0f slow %.0f other %.0f cycles, slow entries %.3f/tile | "
                        "mma per tile: wait_full %.0f wait_tempty %.0f issue %.0f (epi warps %.0f)\n", grid, items,
                h[0] / tiles_e, h[1] / tiles_e, h[2] / tiles_e, h[3] / tiles_e, h[4] / tiles_e, h[5] / tiles_m, h[6] / tiles_m, h[7] / tiles_m, ew);
    }
    return PE_OK;
}

/* exact re-ranking of the candidate blocks the scan produced: the local top-k of this shard.  d_flag (may be null):
 * [0] counter + [4..] ids of the queries whose result the certificate could not prove */
int knn_tc_select(pe_db* db, const float* cand_d, const int* cand_i, int n_lists, int kc, const float* d_q,
                  int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist, int* d_flag) {
    const int n_cand = n_lists * kc;
    /* blocks re-ranked exactly: a block holding one of the k nearest is preceded only by blocks whose key lies within the
     * split's ~2^-22 of it, so a small margin beyond k suffices - and since round 2 the certificate proves every result, so
     * the margin only trades select time against the flag rate, never correctness.  tools/knn_nsel_experiment.py (32,768
     * queries vs a 125 k-row shard, the 8-GPU shape): k + 3 blocks 0.092 ms, k + 2 0.075, k + 1 0.062, all with 0 queries
     * flagged; k + 0 flags 13 (k = 1) / 113 (k = 5).  k + 2 keeps one block of slack over what the data needed. */
    static const int extra = getenv("PE_KNN_NSEL_EXTRA") ? atoi(getenv("PE_KNN_NSEL_EXTRA")) : 2;     /* experiment knob */
    const int n_sel = k + extra < kTcMaxSel ? (k + extra < 1 ? 1 : k + extra) : kTcMaxSel;
    const size_t smem = (size_t)kTcSelWarps * (db->dim_pad + 2 * n_cand + 2 * n_sel * kTcBlock) * 4;
    PE_REQUIRE(smem <= 200 * 1024, "candidate set too large for the select kernel");
    PE_CUDA(cudaFuncSetAttribute(knn_tc_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    /* relative error bound of a scan key in units of |q'| T (see the kernel); PE_KNN_EPS_LOG2 overrides it for experiments */
    static const float eps_rel = std::ldexp(1.f, getenv("PE_KNN_EPS_LOG2") ? atoi(getenv("PE_KNN_EPS_LOG2")) : -17);
    knn_tc_select_kernel<<<(unsigned)((nq + kTcSelWarps - 1) / kTcSelWarps), 32 * kTcSelWarps, smem, db->stream>>>(
        cand_d, cand_i, n_cand, kc, d_q, db->d_desc, db->dim_pad, db->t_local, db->row_offset, k, n_sel, (long long)nq,
        db->key_scale.as<float>(), db->d_center, db->max_norm2_c, eps_rel, d_flag, d_out_idx, d_out_dist);
    PE_LAUNCHED();
    return PE_OK;
}

int knn_tc_select_hamming(pe_db* db, const float* cand_d, const int* cand_i, int n_lists, int kc, const uint32_t* d_q,
                          int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist) {
    const int n_cand = n_lists * kc;
    const int n_sel = k > 8 ? k : 8;
    const size_t smem = (size_t)(2 * n_cand + 2 * n_sel * kTcBlock) * 4;
    PE_REQUIRE(smem <= 200 * 1024, "candidate set too large for the select kernel");
    PE_CUDA(cudaFuncSetAttribute(knn_tc_select_hamming_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    knn_tc_select_hamming_kernel<<<(unsigned)nq, 32, smem, db->stream>>>(cand_d, cand_i, n_cand, d_q, db->d_codes, db->words,
                                                                        db->t_local, db->row_offset, k, n_sel, d_out_idx, d_out_dist);
    PE_LAUNCHED();
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
