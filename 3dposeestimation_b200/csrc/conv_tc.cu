/*
 * conv_tc.cu - conv1 + bias + max-pool + activation of the manifold nets on the tcgen05 tensor path.
 *
 * Replaces the first Convolution / Pooling / activation group that `CNN.Forward()` runs inside
 * networkEvaluator::computeDescriptors (reference src/networkevaluator.cpp:33-48; layers conv1, pool1,
 * relu1 of network/manifold*.prototxt, e.g. manifold_rgbdnor_16.prototxt:6-38): 16 filters of 8x8xC,
 * stride 1, no padding, 2x2/2 MAX pooling with Caffe's ceil sizing (57 -> 29), then ReLU / ELU / PReLU.
 *
 * Formulation (implicit GEMM without im2col, see umma.cuh):
 *   - the patch lives in shared memory as an fp16 HWC8 image (one pixel = 8 channel slots = 16 bytes = one
 *     core-matrix row), split in two terms x ~= hi + lo.  A K-major no-swizzle descriptor with LBO = one pixel
 *     and SBO = eight pixels turns 256 consecutive pixels into the 256 rows of the B operand and (kx, kx+1) x 8
 *     channels into one K step, so the tensor core reads the sliding windows straight out of the image;
 *   - image rows are stored in the order (y mod 8, y div 8), so the four base rows r, r + 8, r + 16, r + 24 of a
 *     tile are adjacent: one B window = 4 base rows x 64 pixels = 256 accumulator columns;
 *   - the A operand is a Toeplitz copy of the weights: accumulator lane (j, f) is output row r + j, filter f, and a
 *     K step over input row r + ky' uses weight row ky' - j (zero outside 0..7).  The Toeplitz operand is itself an
 *     overlapping window over a short array of 256-byte weight slabs (SBO = 128, start = (14 - ky') slabs).  With
 *     16 filters alone M would be 16; eight row offsets fill the 128 lanes at the price of 15 K rows per 8 useful;
 *   - fp32 accuracy: D = Z_hi.X_hi + Z_lo.X_hi + Z_hi.X_lo in fp32 (TMEM), operands scaled by powers of two.  The two
 *     cross terms are ~2^-11 of the product; with F8 (default, pe_set_option("forward_fp8")) they come from ONE
 *     kind::f8f6f4 MMA over e4m3 copies ([x | lo_x] per pixel against [lo_w | w] per filter row, K = 32), which costs
 *     the same 128 cycles as an fp16 MMA and leaves descriptors within ~3e-5 of the row max (fp16 x3: ~1e-6).
 *   Per patch: 2 tiles (conv rows 0..31, 32..63) x 15 input rows x 4 pixel pairs x (2 | 3) products = 240 | 360
 *   MMAs of 128x256, 128 cycles each: the tensor pipe is the bound (ncu: 95 % active).  Which operand slides matters:
 *   with the image as the 128-row A operand and N = 128 (720 MMAs) an MMA took 78 cycles instead of 64, because
 *   shared memory had to deliver 8 KB per 64 cycles and three of four windows are misaligned; at N = 256 the
 *   operand traffic per cycle is lower and the misalignment disappears behind the math.
 *
 * Kernel (persistent, one CTA per SM, 544 threads): warp 0 issues the MMAs, warps 1-4 / 5-8 drain the accumulator
 * of tile 0 / 1 (TMEM loads one 16-column group ahead, pool x in registers, pool y by one shuffle, bias, activation,
 * fp16 hi / lo split, 32-bit stores after a pixel trade between neighbouring filters), warps 9-16 stream the next
 * patch in 16-row blocks (fp32 HWC -> fp16 hi + fp16 lo | e4m3 pair, L1-allocating loads, bank-conflict-free stores)
 * as soon as a block's last reader tile has completed.
 */
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include <cuda_fp8.h>

#include "common.h"
#include "umma.cuh"

namespace pe {

using namespace umma;

namespace {

constexpr int kRowBytes = 1024;                    /* 64 pixels x 16 bytes */
constexpr int kImgRows = 72;                       /* 64 image rows + 8 zero rows (base row 56 reads up to row 70) */
constexpr int kImgBytes = kImgRows * kRowBytes;    /* per term */
constexpr int kZSlabs = 127;                       /* 8 x (7 zero + 8 weight rows) + 7 zero slabs of 256 bytes */
constexpr int kZBytes = kZSlabs * 256;             /* per term */
constexpr int kZStride = 15 * 256;                 /* from the array of kx to the array of kx + 1 */
constexpr int kOffImgHi = 0, kOffImgLo = kImgBytes;
constexpr int kOffZHi = 2 * kImgBytes, kOffZLo = kOffZHi + kZBytes;
constexpr int kOffMisc = kOffZLo + kZBytes;        /* bias[16] slope[16] barriers */
constexpr int kC1TcSmem = kOffMisc + 512;
constexpr int kC1TcThreads = 544;
constexpr int kLog2Sx = 8;                         /* image scale 2^8: |x| <= 255 stays finite in fp16 */
/* fp8 cross terms (F8 kernels): per pixel 16 bytes [x8: e4m3(x s_x 2^-4) x 8 ch | lo8: e4m3((x s_x - hi) 2^6) x 8 ch]
 * against weight slabs [lo_w8: e4m3((w s_w - hi_w) 2^4) | w8: e4m3(w s_w 2^-6)], so one kind::f8f6f4 MMA (K = 32 =
 * two pixels x 16 slots) adds hi_x.lo_w + lo_x.w at the scale s_x s_w of the fp16 hi.hi product.  With |w s_w| < 2^11
 * and |x| <= 28: |x8| <= 448, |lo8| <= 128, |lo_w8| <= 8, |w8| < 32 - all inside e4m3's finite range */
constexpr int kF8XDown = 4, kF8LoUp = 6;
constexpr float kF8MaxInput = 28.f;

constexpr int kP1H = 30, kP1W = 32, kP1Plane = kP1H * kP1W, kP1Patch = 16 * kP1Plane;   /* as forward.cu */

__device__ __forceinline__ int row_pos(int y) { return (y & 7) * 9 + (y >> 3); }

/* pool1 as conv2_tc reads it: 29 x 29 x 16 activations, fp16 hi / lo, two blocks of 8 channels, rows in
 * (y mod 8, y div 8) order (40 row slots), 32 pixels per row, scaled by 2^4 */
constexpr int kP1Rows = 40;
constexpr int kP1PlaneBytes = kP1Rows * 512;             /* one term of one channel block */
constexpr int kP1PackedBytes = 4 * kP1PlaneBytes;        /* hi: cblk0, cblk1; lo: cblk0, cblk1 = 81,920 B per patch */
constexpr float kP1Scale = 16.f;
__host__ __device__ __forceinline__ int p1_row_pos(int y) { return (y & 7) * 5 + (y >> 3); }

__device__ __forceinline__ float act_apply(float v, int act, float slope) {
    const float pos = fmaxf(v, 0.f), neg = fminf(v, 0.f);
    if (act == PE_ACT_RELU) return pos;
    if (act == PE_ACT_ELU) return pos + (expf(neg) - 1.0f);
    if (act == PE_ACT_PRELU) return pos + slope * neg;
    return v;
}

__device__ __forceinline__ uint32_t pack2(__half a, __half b) {
    return static_cast<uint32_t>(__half_as_ushort(a)) | (static_cast<uint32_t>(__half_as_ushort(b)) << 16);
}

struct C1Args {
    const float* patches;      /* [n][64][64][CP] fp32 */
    long long n;
    const uint4* ztoeplitz;    /* hi term then lo term, kZBytes each */
    const float* bias;         /* [16] */
    const float* slope;        /* [16] or null */
    int act;
    float inv_scale;           /* 1 / (s_x s_w) */
    float* pool1;              /* [n][16][30][32] fp32 (SIMT conv2) */
    unsigned char* p1_packed;  /* or, for conv2_tc: [n][hi,lo][2 channel blocks][40 rows][32 px][8 ch] fp16 */
    int* err;                  /* pinned flag: 1.. barrier time-out, 100 input magnitude above 255 */
    int dbg;                   /* PE_C1_DBG experiments (results are wrong when bits 1..256 are set): 1 no stores, 2 no shuffles, 4 loaders
                                  do not store after the first patch, 8 no MMAs, 16 no global loads after the first patch;
                                  512 = round-1 issue order of the split products (results stay correct) */
};

template <int CP, bool F8, bool PACKED>
__global__ void __launch_bounds__(kC1TcThreads, 1) conv1_tc_kernel(const C1Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* s_bias = reinterpret_cast<float*>(smem + kOffMisc);
    float* s_slope = s_bias + 16;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffMisc + 128);
    uint64_t* img_full = bars;          /* [4] 16-row block converted (256 loader threads) */
    uint64_t* img_empty = bars + 4;     /* [4] last reader tile of the block has completed */
    uint64_t* acc_full = bars + 8;      /* [4] tile accumulated */
    uint64_t* acc_empty = bars + 12;    /* [4] tile drained (4 epilogue warps) */
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
    volatile int* s_abort = reinterpret_cast<volatile int*>(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    /* one-time: zero the image buffers (pad rows and the unused channel slot stay zero), load the Toeplitz weights */
    for (int i = tid; i < 2 * kImgBytes / 16; i += kC1TcThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    /* second weight buffer: the fp16 lo term, or (F8) the e4m3 [lo_w | w] slabs that follow it in ztoeplitz */
    for (int i = tid; i < kZBytes / 16; i += kC1TcThreads) {
        reinterpret_cast<uint4*>(smem + kOffZHi)[i] = a.ztoeplitz[i];
        reinterpret_cast<uint4*>(smem + kOffZLo)[i] = a.ztoeplitz[(F8 ? 2 : 1) * (kZBytes / 16) + i];
    }
    if (tid < 16) {
        s_bias[tid] = a.bias[tid];
        s_slope[tid] = a.slope ? a.slope[tid] : 0.f;
    }
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) {
            mbar_init(&img_full[i], 256);
            mbar_init(&img_empty[i], 1);
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 4);
        }
        *s_abort = 0;
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        /* ======================= MMA issuer: the whole warp runs the loops (warp-uniform operands), one elected lane
         * issues.  A = Toeplitz weight window (128 rows = 8 output-row offsets x 16 filters), B = 256 consecutive
         * pixels of the image = base rows r, r + 8, r + 16, r + 24 (adjacent in the (y mod 8, y div 8) order) ===== */
        const uint32_t idesc256 = instr_desc(128, 256, FMT_F16, FMT_F16);
        /* tile 1, input rows ky' >= 8: base row 56 (columns 192..255) only feeds conv row 56 = row offset 0, whose weight
         * rows end at ky' = 7, so these steps run on 192 columns (-5.8 % of the patch's tensor-pipe time) */
        const uint32_t idesc192 = instr_desc(128, 192, FMT_F16, FMT_F16);
        const uint32_t s0 = smem_u32(smem);
        const uint64_t i_hi = smem_desc(s0 + kOffImgHi, 16, 128), i_lo = smem_desc(s0 + kOffImgLo, 16, 128);
        const uint64_t z_hi = smem_desc(s0 + kOffZHi, kZStride, 128), z_lo = smem_desc(s0 + kOffZLo, kZStride, 128);
        uint32_t it = 0;
        bool ok = true;
        for (long long patch = blockIdx.x; patch < a.n && ok && !*s_abort; patch += gridDim.x, ++it) {
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                /* tile 0 reads image rows 0..38 (16-row blocks 0, 1 and, from ky' = 8 on, 2), tile 1 rows 32..70 */
                int e = 0;
                if (!mbar_wait(&acc_empty[t], (it & 1) ^ 1)) e = 1;
                else if (!mbar_wait(&img_full[2 * t], it & 1)) e = 2;
                else if (!mbar_wait(&img_full[2 * t + 1], it & 1)) e = 3;
                if (e) { if (lane == 0) atomicExch(a.err, e); ok = false; break; }
                fence_after_sync();
                const uint32_t d = tmem + t * 256;
                /* Issue order.  The tensor core adds every MMA into the fp32 accumulator with truncation (round toward zero):
                 * a bias of ~2^-24 of the accumulator's magnitude per MMA that adds up along the chain (the effect Ootomo &
                 * Yokota 2022 describe for split-precision GEMMs).  The cross terms are 2^-11 of the result, so they are
                 * issued FIRST, while the accumulator is small, and only the 60 hi.hi MMAs truncate at full magnitude
                 * (measured on B200, tools/c1_order.sh: descriptor error 6.6e-6 -> 4.6e-6 of the row max; with the same
                 * change in conv2 and two accumulators in the InnerProducts 1.3e-5 -> 4.6e-6).  Several input rows go into
                 * one elected block: below ~6 MMAs per block the issue loop, not the tensor pipe, sets the pace (4 per
                 * block: conv1 0.76 ms; 12-16 per block: 0.715 ms; the interleaved order, PE_C1_DBG=512: 0.70 ms). */
                auto rows = [&](int ky, uint32_t& i_row, uint32_t& z_row, uint32_t& idesc) {
                    i_row = static_cast<uint32_t>(row_pos(32 * t + ky)) * (kRowBytes >> 4);
                    z_row = static_cast<uint32_t>(14 - ky) * (256 >> 4);
                    idesc = (t == 1 && ky >= 8) ? idesc192 : idesc256;
                };
                if (!(a.dbg & 512)) {
                    /* cross terms: two input rows per elected block (16 MMAs), then hi.hi: three rows per block (12 MMAs) */
#pragma unroll 1
                    for (int ky0 = 0; ky0 < 15; ky0 += 2) {
                        if (t == 0 && ky0 == 8) {
                            if (!mbar_wait(&img_full[2], it & 1)) { if (lane == 0) atomicExch(a.err, 6); ok = false; break; }
                            fence_after_sync();
                        }
                        if (!(a.dbg & 8) && elect_one()) {
#pragma unroll
                            for (int dk = 0; dk < 2; ++dk) {
                                const int ky = ky0 + dk;
                                if (ky < 15) {
                                    uint32_t i_row, z_row, idesc;
                                    rows(ky, i_row, z_row, idesc);
#pragma unroll
                                    for (int p = 0; p < 4; ++p) {
                                        const uint32_t io = i_row + p * 2, zo = z_row + p * 2 * (kZStride >> 4);
                                        if (F8) mma_f8(d, z_lo + zo, i_lo + io, idesc, (ky | p) != 0);     /* hi_x.lo_w + lo_x.w in e4m3, K = 32 */
                                        else { mma_f16(d, z_lo + zo, i_hi + io, idesc, (ky | p) != 0); mma_f16(d, z_hi + zo, i_lo + io, idesc, 1); }
                                    }
                                }
                            }
                        }
                        __syncwarp();
                    }
                    if (ok) {
#pragma unroll 1
                        for (int ky0 = 0; ky0 < 15; ky0 += 3) {
                            if (!(a.dbg & 8) && elect_one()) {
#pragma unroll 1
                                for (int ky = ky0; ky < ky0 + 3; ++ky) {
                                    uint32_t i_row, z_row, idesc;
                                    rows(ky, i_row, z_row, idesc);
#pragma unroll
                                    for (int p = 0; p < 4; ++p) mma_f16(d, z_hi + z_row + p * 2 * (kZStride >> 4), i_hi + i_row + p * 2, idesc, 1);
                                }
                            }
                            __syncwarp();
                        }
                    }
                } else {
                    /* PE_C1_DBG=512: the three products of a K step back to back (round-1 order), for the A/B measurement */
#pragma unroll 1
                    for (int ky = 0; ky < 15; ++ky) {
                        if (t == 0 && ky == 8) {
                            if (!mbar_wait(&img_full[2], it & 1)) { if (lane == 0) atomicExch(a.err, 6); ok = false; break; }
                            fence_after_sync();
                        }
                        uint32_t i_row, z_row, idesc;
                        rows(ky, i_row, z_row, idesc);
                        if (!(a.dbg & 8) && elect_one()) {
#pragma unroll
                            for (int p = 0; p < 4; ++p) {
                                const uint32_t io = i_row + p * 2, zo = z_row + p * 2 * (kZStride >> 4);
                                mma_f16(d, z_hi + zo, i_hi + io, idesc, (ky | p) != 0);
                                if (F8) { mma_f8(d, z_lo + zo, i_lo + io, idesc, 1); }
                                else { mma_f16(d, z_lo + zo, i_hi + io, idesc, 1); mma_f16(d, z_hi + zo, i_lo + io, idesc, 1); }
                            }
                        }
                        __syncwarp();
                    }
                }
                if (!ok) break;
                if (elect_one()) {
                    mma_commit(&img_empty[2 * t]);
                    mma_commit(&img_empty[2 * t + 1]);
                    mma_commit(&acc_full[t]);
                }
                __syncwarp();
            }
        }
        if (!ok) *s_abort = 1;
    } else if (warp <= 8) {
        /* ======================= epilogue: warps 1-4 drain tile 0 (conv rows 0..31), warps 5-8 tile 1 (32..63).
         * Accumulator lane = (row offset j, filter f), column = (base row b, x).  Lanes j = 2q and 2q + 1 of a pool
         * window sit 16 apart in the same warp: pool x inside the thread, y by one shuffle, then the even-j lane keeps
         * the even pooled columns and the odd-j lane the odd ones, so one store instruction fills whole 32-byte
         * sectors of the packed image ======================= */
        const int t = (warp - 1) >> 2;
        const int quarter = warp & 3;                 /* the TMEM lane quarter this warp may read */
        const int m = quarter * 32 + lane;            /* accumulator lane */
        const int f = m & 15, jpar = (m >> 4) & 1;
        const float bias = s_bias[f];
        const float slope = a.act == PE_ACT_PRELU ? s_slope[f] : 0.f;
        const int fodd = f & 1;
        const bool is_elu = a.act == PE_ACT_ELU;
        const bool is_lin = a.act != PE_ACT_RELU && a.act != PE_ACT_ELU && a.act != PE_ACT_PRELU;
        int n_b = 0;                                  /* base rows of this tile with a valid conv row for this warp */
        for (int b = 0; b < 4; ++b) n_b += (32 * t + 8 * b + 2 * quarter) <= 56;
        float amax = 0.f;
        uint32_t sink = 0;
        uint32_t it = 0;
        for (long long patch = blockIdx.x; patch < a.n && !*s_abort; patch += gridDim.x, ++it) {
            float* out = a.pool1 + patch * (long long)kP1Patch + f * kP1Plane;
            unsigned char* outp = a.p1_packed + ((a.dbg & 128) ? (long long)blockIdx.x : patch) * (long long)kP1PackedBytes + (f >> 3) * kP1PlaneBytes + (f & 6) * 2;
            if (!mbar_wait(&acc_full[t], it & 1)) { atomicExch(a.err, 4); *s_abort = 1; break; }
            fence_after_sync();
            const uint32_t tbase = tmem + t * 256 + (static_cast<uint32_t>(quarter * 32) << 16);
            /* one 16-column group (base row b, quarter g) of this thread's accumulator lane: pool x in the thread, send the
             * partner lane its half, pool y, bias, activation, store.  Under the tensor pipe's operand traffic every
             * shuffle / store spends ~150 cycles in the memory-instruction queue, so the four shuffles and the eight
             * stores of a group are issued back to back (row pointer + immediate offsets, no per-store address) */
            auto drain = [&](const uint32_t (&v)[16], int g, bool own_ok, bool oth_ok, unsigned char* rowp, float* rowf) {
                float mine[4], snd[4], oth[4], keep[4];
#pragma unroll
                for (int i2 = 0; i2 < 4; ++i2) {
                    const int x0 = 16 * g + 4 * i2;                 /* pooled columns x0 / 2 and x0 / 2 + 1 */
                    const float m0 = x0 + 1 <= 56 ? fmaxf(__uint_as_float(v[4 * i2]), __uint_as_float(v[4 * i2 + 1])) : __uint_as_float(v[4 * i2]);
                    const float m1 = x0 + 3 <= 56 ? fmaxf(__uint_as_float(v[4 * i2 + 2]), __uint_as_float(v[4 * i2 + 3])) : __uint_as_float(v[4 * i2 + 2]);
                    mine[i2] = jpar ? m1 : m0;
                    snd[i2] = jpar ? m0 : m1;
                }
#pragma unroll
                for (int i2 = 0; i2 < 4; ++i2) oth[i2] = (a.dbg & 2) ? snd[i2] : __shfl_xor_sync(0xffffffffu, snd[i2], 16);
#pragma unroll
                for (int i2 = 0; i2 < 4; ++i2) {
                    float c = own_ok ? mine[i2] : -FLT_MAX;
                    if (oth_ok) c = fmaxf(c, oth[i2]);
                    /* branch-free activation: ReLU / PReLU = pos + slope neg (slope 0 for ReLU), ELU = pos + (e^neg - 1) */
                    const float y = fmaf(c, a.inv_scale, bias);
                    const float pos = fmaxf(y, 0.f), neg = fminf(y, 0.f);
                    const float tail = is_elu ? expf(neg) - 1.0f : slope * neg;
                    keep[i2] = is_lin ? y : pos + tail;
                }
                if (a.dbg & 1) return;
                if (PACKED) {
                    /* conv2_tc's A operand: [term][channel block][row (y%8, y/8)][32 px][8 channels] fp16 */
                    uint32_t hl[4];
#pragma unroll
                    for (int i2 = 0; i2 < 4; ++i2) {
                        __half hi, lo;
                        amax = fmaxf(amax, fabsf(keep[i2]));
                        split_f16(keep[i2] * kP1Scale, hi, lo);
                        hl[i2] = pack2(hi, lo);
                    }
#pragma unroll
                    if (a.dbg & 32) { sink ^= hl[0] ^ hl[1] ^ hl[2] ^ hl[3]; return; }
                    /* a 2-byte store per lane costs ~15 cycles of the load/store unit per warp instruction.  Neighbouring
                     * filters (lanes f, f ^ 1) therefore trade two of their four pixels: the even-f lane ends up with both
                     * channels of pixels i2 = 0, 2, the odd-f lane with those of i2 = 1, 3, and every store is 32 bits
                     * (a warp instruction writes four whole 32-byte sectors) */
                    const uint32_t r0 = __shfl_xor_sync(0xffffffffu, fodd ? hl[0] : hl[1], 1);
                    const uint32_t r1 = __shfl_xor_sync(0xffffffffu, fodd ? hl[2] : hl[3], 1);
                    const uint32_t ea = fodd ? r0 : hl[0], oa = fodd ? hl[1] : r0;      /* even / odd channel, first pixel */
                    const uint32_t eb = fodd ? r1 : hl[2], ob = fodd ? hl[3] : r1;      /* second pixel (4 further right) */
                    unsigned char* q = rowp + g * 128;
                    if (a.dbg & 64) {
                        __stcs(reinterpret_cast<uint32_t*>(q), __byte_perm(ea, oa, 0x5410));
                        __stcs(reinterpret_cast<uint32_t*>(q + 2 * kP1PlaneBytes), __byte_perm(ea, oa, 0x7632));
                        if (g < 3 || (!fodd && jpar == 0)) {
                            __stcs(reinterpret_cast<uint32_t*>(q + 64), __byte_perm(eb, ob, 0x5410));
                            __stcs(reinterpret_cast<uint32_t*>(q + 64 + 2 * kP1PlaneBytes), __byte_perm(eb, ob, 0x7632));
                        }
                        return;
                    }
                    *reinterpret_cast<uint32_t*>(q) = __byte_perm(ea, oa, 0x5410);                           /* hi terms */
                    *reinterpret_cast<uint32_t*>(q + 2 * kP1PlaneBytes) = __byte_perm(ea, oa, 0x7632);       /* lo terms */
                    if (g < 3 || (!fodd && jpar == 0)) {
                        *reinterpret_cast<uint32_t*>(q + 64) = __byte_perm(eb, ob, 0x5410);
                        *reinterpret_cast<uint32_t*>(q + 64 + 2 * kP1PlaneBytes) = __byte_perm(eb, ob, 0x7632);
                    }
                } else {
#pragma unroll
                    for (int i2 = 0; i2 < 4; ++i2) {
                        const int pxe = 8 * g + 2 * i2;
                        if (pxe < 28 || (pxe == 28 && jpar == 0)) rowf[pxe] = keep[i2];
                    }
                }
            };
            /* the TMEM loads run one group ahead of the arithmetic (two 16-register buffers) */
            uint32_t v0[16], v1[16];
            tmem_ld16(tbase, v0);
#pragma unroll 1
            for (int b = 0; b < n_b; ++b) {
                const int ye = 32 * t + 8 * b + 2 * quarter;          /* conv rows ye (even-j lanes) and ye + 1 */
                const int py = ye >> 1;
                const bool row_o_ok = ye + 1 <= 56;
                const bool own_ok = jpar ? row_o_ok : true, oth_ok = jpar ? true : row_o_ok;
                unsigned char* rowp = outp + p1_row_pos(py) * 512 + jpar * 16 + fodd * 32;   /* pixel jpar (+ 2 for odd f) */
                float* rowf = out + py * kP1W + jpar;
                tmem_ld_wait();
                tmem_ld16(tbase + b * 64 + 16, v1);
                drain(v0, 0, own_ok, oth_ok, rowp, rowf);
                tmem_ld_wait();
                tmem_ld16(tbase + b * 64 + 32, v0);
                drain(v1, 1, own_ok, oth_ok, rowp, rowf);
                tmem_ld_wait();
                tmem_ld16(tbase + b * 64 + 48, v1);
                drain(v0, 2, own_ok, oth_ok, rowp, rowf);
                tmem_ld_wait();
                if (b + 1 < n_b) {
                    tmem_ld16(tbase + (b + 1) * 64, v0);
                } else {              /* every column this warp needs is in registers: hand the accumulator back */
                    fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[t]);
                }
                drain(v1, 3, own_ok, oth_ok, rowp, rowf);
            }
        }
        if (sink == 0x12345678u) atomicExch(a.err, 7);
        if (!(amax <= 4000.f)) atomicExch(a.err, 101);   /* packed activations must stay inside the fp16 split range */
    } else {
        /* ======================= loaders: fp32 HWC -> fp16 hi / lo HWC8, 16 rows at a time ======================= */
        const int lt = tid - 288;                      /* 0..255: row (lt >> 4) of the block, pixels 4 * (lt & 15) .. +3 */
        const int yr = lt >> 4, x0 = (lt & 15) * 4;
        const int rot = (lt >> 1) & 3;                 /* see the store loop */
        const bool rot1 = rot & 1, rot2 = rot & 2;
        const float sx = static_cast<float>(1 << kLog2Sx);
        float vmax = 0.f, vnan = 0.f;                  /* fmaxf drops NaN operands: x * 0 accumulates NaN / Inf inputs into vnan */
        uint32_t it = 0;
        for (long long patch = blockIdx.x; patch < a.n && !*s_abort; patch += gridDim.x, ++it) {
            const float* src = a.patches + patch * (long long)(4096 * CP);
#pragma unroll 1
            for (int rb = 0; rb < 4; ++rb) {
                const int y = rb * 16 + yr;
                float val[4 * CP];
                const float4* s4 = reinterpret_cast<const float4*>(src + (size_t)(y * 64 + x0) * CP);
#pragma unroll
                for (int i = 0; i < CP; ++i) {
                    float4 q = make_float4(0.5f, 0.25f, 0.125f, 0.75f);
                    /* a thread's seven 16-byte loads walk 112 contiguous bytes, so every 32-byte sector is touched by two
                     * consecutive loads: they must allocate in L1 (with L1::no_allocate each sector came from L2 twice
                     * and the doubled request stream delayed the epilogue's stores: 0.68 instead of 0.52 ms) */
                    if (a.dbg & 256)
                    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "l"(s4 + i));
                    else if (!((a.dbg & 16) && it > 0))
                    asm volatile("ld.global.L1::evict_first.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "l"(s4 + i));
                    val[4 * i + 0] = q.x; val[4 * i + 1] = q.y; val[4 * i + 2] = q.z; val[4 * i + 3] = q.w;
                }
                if (!mbar_wait_relaxed(&img_empty[rb], (it & 1) ^ 1)) { atomicExch(a.err, 5); *s_abort = 1; break; }
                unsigned char* dst = smem + row_pos(y) * kRowBytes + x0 * 16;
                if ((a.dbg & 4) && it > 0) { mbar_arrive(&img_full[rb]); continue; }
                uint32_t H[4][4], L[4][4];
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    uint32_t (&hi)[4] = H[p];
                    uint32_t (&lo)[4] = L[p];
#pragma unroll
                    for (int c2 = 0; c2 < 4; ++c2) {
                        __half h[2], l[2];
                        float xs[2], ls[2];
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int ch = 2 * c2 + e;
                            const float xv = ch < CP ? val[p * CP + (ch < CP ? ch : 0)] : 0.f;
                            vmax = fmaxf(vmax, fabsf(xv));
                            vnan = fmaf(xv, 0.f, vnan);
                            split_f16(xv * sx, h[e], l[e]);
                            xs[e] = xv * (sx / static_cast<float>(1 << kF8XDown));
                            ls[e] = (xv * sx - __half2float(h[e])) * static_cast<float>(1 << kF8LoUp);
                        }
                        hi[c2] = pack2(h[0], h[1]);
                        if (F8) {
                            /* bytes 0..7 of the pixel: x8 of channels 0..7, bytes 8..15: lo8 */
                            const uint32_t x2 = __nv_cvt_float2_to_fp8x2(make_float2(xs[0], xs[1]), __NV_SATFINITE, __NV_E4M3);
                            const uint32_t l2 = __nv_cvt_float2_to_fp8x2(make_float2(ls[0], ls[1]), __NV_SATFINITE, __NV_E4M3);
                            const int w = c2 >> 1, sh = (c2 & 1) * 16;
                            if ((c2 & 1) == 0) { lo[w] = x2; lo[2 + w] = l2; }
                            else { lo[w] |= x2 << sh; lo[2 + w] |= l2 << sh; }
                        } else {
                            lo[c2] = pack2(l[0], l[1]);
                        }
                    }
                }
                /* a quarter warp writes 8 x 16 bytes at a 64-byte stride: storing pixel s at step s would hit only two
                 * 16-byte bank groups (4-way conflict, 4x the shared-memory cycles, and this kernel runs the shared-memory
                 * pipe close to its limit with the tensor operands).  Lane l therefore stores pixel (s + (l >> 1)) & 3
                 * at step s - a rotation of its four pixels by two select stages - which spreads the eight lanes over
                 * all banks */
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    uint32_t t0 = H[0][w], t1 = H[1][w], t2 = H[2][w], t3 = H[3][w];
                    H[0][w] = rot1 ? t1 : t0; H[1][w] = rot1 ? t2 : t1; H[2][w] = rot1 ? t3 : t2; H[3][w] = rot1 ? t0 : t3;
                    t0 = H[0][w]; t1 = H[1][w]; t2 = H[2][w]; t3 = H[3][w];
                    H[0][w] = rot2 ? t2 : t0; H[1][w] = rot2 ? t3 : t1; H[2][w] = rot2 ? t0 : t2; H[3][w] = rot2 ? t1 : t3;
                    t0 = L[0][w]; t1 = L[1][w]; t2 = L[2][w]; t3 = L[3][w];
                    L[0][w] = rot1 ? t1 : t0; L[1][w] = rot1 ? t2 : t1; L[2][w] = rot1 ? t3 : t2; L[3][w] = rot1 ? t0 : t3;
                    t0 = L[0][w]; t1 = L[1][w]; t2 = L[2][w]; t3 = L[3][w];
                    L[0][w] = rot2 ? t2 : t0; L[1][w] = rot2 ? t3 : t1; L[2][w] = rot2 ? t0 : t2; L[3][w] = rot2 ? t1 : t3;
                }
#pragma unroll
                for (int st = 0; st < 4; ++st) {
                    const int po = ((st + rot) & 3) * 16;
                    *reinterpret_cast<uint4*>(dst + kOffImgHi + po) = make_uint4(H[st][0], H[st][1], H[st][2], H[st][3]);
                    *reinterpret_cast<uint4*>(dst + kOffImgLo + po) = make_uint4(L[st][0], L[st][1], L[st][2], L[st][3]);
                }
                fence_async_smem();
                mbar_arrive(&img_full[rb]);
            }
        }
        if (!(vmax <= (F8 ? kF8MaxInput : 255.f)) || vnan != 0.f) atomicExch(a.err, (F8 && vnan == 0.f) ? 102 : 100);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

/* ------------------------------------------------------------------ conv2 ---- */
/* conv2 (7 filters 5x5x16) + bias + pool (25 -> 13, ceil) + act on the same scheme: the A operand is the packed
 * pool1 image (one 80 KB cp.async.bulk per patch, double buffered), a K step is (ky', kx) x 16 channels (the two
 * 8-channel blocks are the two K chunks, LBO = one plane), N = 8 output rows x 8 filter slots = 64 through a
 * Toeplitz window over 128-byte weight slabs, one 128-lane tile = base rows 0, 8, 16, 24 x 32 pixels.
 * 12 x 5 x 3 = 180 MMAs (128x64x16) per patch. */
constexpr int kZ2Slabs = 127, kZ2Bytes = kZ2Slabs * 256, kZ2Stride = 12 * 256;   /* a slab = [hi 8 x 8][lo 8 x 8] halves */
constexpr int kC2OffImg = 0;                                   /* 2 x kP1PackedBytes */
constexpr int kC2OffZ = 2 * kP1PackedBytes;
constexpr int kC2OffMisc = kC2OffZ + kZ2Bytes;
constexpr int kC2TcSmem = kC2OffMisc + 256;
constexpr int kC2TcThreads = 192;
constexpr int kPool2Ld = 1184;

struct C2Args {
    const unsigned char* p1_packed;
    long long n;
    const uint4* ztoeplitz;    /* kZ2Bytes: hi and lo slabs interleaved */
    const float* bias;         /* [8] */
    const float* slope;        /* [8] or null */
    int act;
    float inv_scale;
    float* pool2;              /* [n][1184] */
    int* err;
};

__global__ void __launch_bounds__(kC2TcThreads, 1) conv2_tc_kernel(const C2Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* s_bias = reinterpret_cast<float*>(smem + kC2OffMisc);
    float* s_slope = s_bias + 8;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kC2OffMisc + 64);
    uint64_t* img_full = bars;          /* [2] patch landed (tx bytes) */
    uint64_t* img_empty = bars + 2;     /* [2] its MMAs have completed */
    uint64_t* acc_full = bars + 4;      /* [2] */
    uint64_t* acc_empty = bars + 6;     /* [2] four epilogue warps */
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    volatile int* s_abort = reinterpret_cast<volatile int*>(tmem_slot + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < kZ2Bytes / 16; i += kC2TcThreads) reinterpret_cast<uint4*>(smem + kC2OffZ)[i] = a.ztoeplitz[i];
    if (tid < 8) {
        s_bias[tid] = a.bias[tid];
        s_slope[tid] = a.slope ? a.slope[tid] : 0.f;
    }
    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&img_full[i], 1);
            mbar_init(&img_empty[i], 1);
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 4);
        }
        *s_abort = 0;
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc<256>(tmem_slot);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (long long patch = blockIdx.x; patch < a.n && !*s_abort; patch += gridDim.x, ++it) {
                const int s = it & 1;
                if (!mbar_wait(&img_empty[s], ((it >> 1) & 1) ^ 1)) { atomicExch(a.err, 21); *s_abort = 1; break; }
                mbar_expect_tx(&img_full[s], kP1PackedBytes);
                bulk_g2s(smem + kC2OffImg + s * kP1PackedBytes, a.p1_packed + patch * (long long)kP1PackedBytes, kP1PackedBytes, &img_full[s]);
            }
        }
    } else if (warp == 1) {
        /* MMA issuer: whole warp, one elected lane issues (see conv1) */
        const uint32_t idesc = instr_desc(128, 128, FMT_F16, FMT_F16);
        const uint32_t s0 = smem_u32(smem);
        const uint64_t b_cat = smem_desc(s0 + kC2OffZ, kZ2Stride, 128);
        uint32_t it = 0;
        bool ok = true;
        for (long long patch = blockIdx.x; patch < a.n && ok && !*s_abort; patch += gridDim.x, ++it) {
            const int s = it & 1;
            const uint32_t ph = (it >> 1) & 1;
            if (!mbar_wait(&acc_empty[s], ph ^ 1)) { if (lane == 0) atomicExch(a.err, 22); ok = false; break; }
            if (!mbar_wait(&img_full[s], ph)) { if (lane == 0) atomicExch(a.err, 23); ok = false; break; }
            fence_after_sync();
            const uint32_t img = s0 + kC2OffImg + s * kP1PackedBytes;
            const uint64_t a_hi = smem_desc(img, kP1PlaneBytes, 128), a_lo = smem_desc(img + 2 * kP1PlaneBytes, kP1PlaneBytes, 128);
            const uint32_t d = tmem + s * 128;
            /* lo image terms first: they are 2^-11 of the result, so the accumulator is still small while they are added and
             * only the 60 hi MMAs truncate (round toward zero) at full magnitude - see the issue order of conv1 */
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
#pragma unroll 1
            for (int ky = 0; ky < 12; ++ky) {
                const uint32_t a_row = static_cast<uint32_t>(p1_row_pos(ky)) * (512 >> 4);
                const uint32_t b_row = static_cast<uint32_t>(11 - ky) * (256 >> 4);
                const uint64_t a_op = pass == 0 ? a_lo : a_hi;
                if (elect_one()) {
#pragma unroll
                    for (int kx = 0; kx < 5; ++kx) {
                        const uint32_t ao = a_row + kx, bo = b_row + kx * 2 * (kZ2Stride >> 4);
                        mma_f16(d, a_op + ao, b_cat + bo, idesc, (pass | ky | kx) != 0);
                    }
                }
                __syncwarp();
            }
            }
            if (elect_one()) {
                mma_commit(&img_empty[s]);
                mma_commit(&acc_full[s]);
            }
            __syncwarp();
        }
        if (!ok) *s_abort = 1;
    } else {
        const int quarter = warp & 3;
        const int m = quarter * 32 + lane;
        const int g = m >> 5, x = m & 31;            /* base conv row 8g, pixel x */
        const int px = x >> 1;
        const bool own_ok = x <= 24, other_ok = (x ^ 1) <= 24;
        const int fsel = (x & 1) * 4;                 /* even lanes store filters 0..3, odd lanes 4..6 */
        uint32_t it = 0;
        for (long long patch = blockIdx.x; patch < a.n && !*s_abort; patch += gridDim.x, ++it) {
            const int s = it & 1;
            if (!mbar_wait(&acc_full[s], (it >> 1) & 1)) { atomicExch(a.err, 24); *s_abort = 1; break; }
            fence_after_sync();
            const uint32_t tbase = tmem + s * 128 + (static_cast<uint32_t>(quarter * 32) << 16);
            /* columns (row offset j, term, filter slot): the hi and lo weight slabs alternate, their sums are the result */
            uint32_t v[2][32];
            {
                uint32_t w[2][32];
                tmem_ld32(tbase, w[0]);
                tmem_ld32(tbase + 32, w[1]);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        v[0][j * 8 + q] = __float_as_uint(__uint_as_float(w[j >> 1][(j & 1) * 16 + q]) + __uint_as_float(w[j >> 1][(j & 1) * 16 + 8 + q]));
                tmem_ld32(tbase + 64, w[0]);
                tmem_ld32(tbase + 96, w[1]);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        v[1][j * 8 + q] = __float_as_uint(__uint_as_float(w[j >> 1][(j & 1) * 16 + q]) + __uint_as_float(w[j >> 1][(j & 1) * 16 + 8 + q]));
            }
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[s]);
            float* out = a.pool2 + patch * (long long)kPool2Ld;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int ya = 8 * g + 2 * jj;         /* conv rows ya, ya + 1 -> pooled row ya / 2 */
                const int py = ya >> 1;
                const bool row_b = ya + 1 <= 24;
#pragma unroll
                for (int f = 0; f < 7; ++f) {
                    float c = __uint_as_float(v[jj >> 1][(jj & 1) * 16 + f]);
                    if (row_b) c = fmaxf(c, __uint_as_float(v[jj >> 1][(jj & 1) * 16 + 8 + f]));
                    if (!own_ok) c = -FLT_MAX;
                    const float o = __shfl_xor_sync(0xffffffffu, c, 1);
                    if (other_ok) c = fmaxf(c, o);
                    if ((f & 4) == fsel && ya <= 24 && px <= 12)
                        out[f * 169 + py * 13 + px] = act_apply(fmaf(c, a.inv_scale, s_bias[f]), a.act, s_slope[f]);
                }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem);
}

}  // namespace

/* Toeplitz copy of the conv1 weights for the tensor kernel (see the header comment): per term (hi, lo) and per kx
 * an array of 256-byte slabs [16 filters][8 patch channels]; slab 7 + (7 - ky) of array kx holds weight row ky.
 * The Slice/Concat/Silence channel select (network/manifold_rgb_16.prototxt:6-33) is folded in here: a patch
 * channel that the net drops gets zero weights, so the image is converted in patch channel order. */
int conv1_tc_prepare(pe_net* net) {
    net->c1_tc_ready = false;
    const int cp = net->def.channels;
    if (cp != 4 && cp != 7) return PE_OK;     /* loader instantiations; other counts use the SIMT kernel */
    const std::vector<float>& w = net->blobs[net->blob_w1];   /* [16][cin][8][8] */
    const int cin = net->cin;
    float wmax = 0.f;
    for (float v : w) wmax = std::fmax(wmax, std::fabs(v));
    if (!(wmax > 0.f) || !std::isfinite(wmax)) return PE_OK;
    const int log2_sw = 10 - std::ilogb(wmax);
    const float sw = std::ldexp(1.f, log2_sw);
    std::vector<int> slot_of(8, -1);
    for (size_t s = 0; s < net->def.sel.size(); ++s) slot_of[net->def.sel[s]] = (int)s;
    std::vector<uint16_t> z((size_t)3 * kZBytes / 2, 0);     /* fp16 hi, fp16 lo, e4m3 [lo_w | w] */
    unsigned char* z8 = reinterpret_cast<unsigned char*>(z.data()) + (size_t)2 * kZBytes;
    for (int kx = 0; kx < 8; ++kx)
        for (int ky = 0; ky < 8; ++ky)
            for (int f = 0; f < 16; ++f)
                for (int ch = 0; ch < cp; ++ch) {
                    const int sl = slot_of[ch];
                    if (sl < 0 || sl >= cin) continue;
                    const float v = w[(((size_t)f * cin + sl) * 8 + ky) * 8 + kx] * sw;
                    const __half h = __float2half_rn(v);
                    const __half l = __float2half_rn(v - __half2float(h));
                    const size_t slab = (size_t)kx * 15 + 7 + (7 - ky);
                    const size_t at = slab * 128 + (size_t)f * 8 + ch;       /* in halves */
                    uint16_t hb, lb;
                    memcpy(&hb, &h, 2);
                    memcpy(&lb, &l, 2);
                    z[at] = hb;
                    z[(size_t)kZBytes / 2 + at] = lb;
                    /* the same slab in e4m3: bytes 0..7 of a filter row = lo_w (meets x8), bytes 8..15 = w (meets lo8) */
                    const size_t at8 = slab * 256 + (size_t)f * 16 + ch;
                    z8[at8] = __nv_cvt_float_to_fp8(std::ldexp(v - __half2float(h), kF8XDown), __NV_SATFINITE, __NV_E4M3);
                    z8[at8 + 8] = __nv_cvt_float_to_fp8(std::ldexp(v, -kF8LoUp), __NV_SATFINITE, __NV_E4M3);
                }
    PE_CUDA(cudaSetDevice(net->device));
    if (!net->d_z1) PE_CUDA(cudaMalloc(&net->d_z1, 3 * kZBytes));
    PE_CUDA(cudaMemcpy(net->d_z1, z.data(), 3 * kZBytes, cudaMemcpyHostToDevice));
    if (!net->tc_err) {
        PE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&net->tc_err), sizeof(int), cudaHostAllocMapped));
        *net->tc_err = 0;
    }
    net->c1_inv_scale = std::ldexp(1.f, -(log2_sw + kLog2Sx));
    net->fp8_state = 0;                       /* new weights: the e4m3 cross terms are re-qualified on the next batch */
    net->c1_tc_ready = true;
    return PE_OK;
}

/* Toeplitz copy of the conv2 weights [7][16][5][5]: per term, per (kx, channel block) an array of 128-byte slabs
 * [8 filter slots][8 channels]; slab 7 + (4 - ky) of array 2 kx + cblk holds weight row ky */
int conv2_tc_prepare(pe_net* net) {
    net->c2_tc_ready = false;
    if (!net->c1_tc_ready) return PE_OK;
    const std::vector<float>& w = net->blobs[net->blob_w2];
    float wmax = 0.f;
    for (float v : w) wmax = std::fmax(wmax, std::fabs(v));
    if (!(wmax > 0.f) || !std::isfinite(wmax)) return PE_OK;
    const int log2_sw = 10 - std::ilogb(wmax);
    const float sw = std::ldexp(1.f, log2_sw);
    std::vector<uint16_t> z((size_t)kZ2Bytes / 2, 0);
    for (int kx = 0; kx < 5; ++kx)
        for (int ky = 0; ky < 5; ++ky)
            for (int f = 0; f < 7; ++f)
                for (int c = 0; c < 16; ++c) {
                    const float v = w[(((size_t)f * 16 + c) * 5 + ky) * 5 + kx] * sw;
                    const __half h = __float2half_rn(v);
                    const __half l = __float2half_rn(v - __half2float(h));
                    uint16_t hb, lb;
                    memcpy(&hb, &h, 2);
                    memcpy(&lb, &l, 2);
                    const size_t slab = (size_t)(kx * 2 + c / 8) * 12 + 7 + (4 - ky);
                    const size_t at = slab * 128 + (size_t)f * 8 + (c % 8);      /* in halves: hi rows, then the 8 lo rows */
                    z[at] = hb;
                    z[at + 64] = lb;
                }
    PE_CUDA(cudaSetDevice(net->device));
    if (!net->d_z2) PE_CUDA(cudaMalloc(&net->d_z2, kZ2Bytes));
    PE_CUDA(cudaMemcpy(net->d_z2, z.data(), kZ2Bytes, cudaMemcpyHostToDevice));
    net->c2_inv_scale = std::ldexp(1.f, -(log2_sw + 4));      /* activations carry kP1Scale = 2^4 */
    net->c2_tc_ready = true;
    return PE_OK;
}

size_t conv_tc_p1_packed_bytes() { return (size_t)kP1PackedBytes; }

int conv2_tc_launch(pe_net* net, long long n) {
    C2Args a;
    a.p1_packed = net->p1_packed.as<unsigned char>();
    a.n = n;
    a.ztoeplitz = reinterpret_cast<const uint4*>(net->d_z2);
    a.bias = net->d_b2;
    a.slope = net->d_s2;
    a.act = net->act2;
    a.inv_scale = net->c2_inv_scale;
    a.pool2 = net->pool2.as<float>();
    a.err = net->tc_err;
    const int grid = (int)(n < net->num_sms ? n : net->num_sms);
    PE_CUDA(cudaFuncSetAttribute(conv2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kC2TcSmem));
    conv2_tc_kernel<<<grid, kC2TcThreads, kC2TcSmem, net->stream>>>(a);
    PE_LAUNCHED();
    return PE_OK;
}

bool conv1_tc_usable(const pe_net* net, int patch_c, const void* d_patches) {
    return g_opt_forward_tensor.load() != 0 && net->c1_tc_ready && patch_c == net->def.channels &&
           (reinterpret_cast<uintptr_t>(d_patches) & 15) == 0;
}

int conv1_tc_launch(pe_net* net, const float* d_patches, int patch_c, long long n, bool packed_out) {
    C1Args a;
    a.patches = d_patches;
    a.n = n;
    a.ztoeplitz = reinterpret_cast<const uint4*>(net->d_z1);
    a.bias = net->d_b1;
    a.slope = net->d_s1;
    a.act = net->act1;
    a.inv_scale = net->c1_inv_scale;
    a.pool1 = net->pool1.as<float>();
    a.p1_packed = packed_out ? net->p1_packed.as<unsigned char>() : nullptr;
    a.err = net->tc_err;
    static const int dbg = getenv("PE_C1_DBG") ? atoi(getenv("PE_C1_DBG")) : 0;
    a.dbg = dbg;
    const int grid = (int)(n < net->num_sms ? n : net->num_sms);
    /* forward_fp8: 0 never, 2 always, 1 (default) where the calibration of net_forward_dev qualified these weights */
    const int opt8 = g_opt_forward_fp8.load();
    const bool f8 = opt8 >= 2 || (opt8 == 1 && net->fp8_state == 1);
    auto launch = [&](auto kernel) -> int {
        PE_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kC1TcSmem));
        kernel<<<grid, kC1TcThreads, kC1TcSmem, net->stream>>>(a);
        return PE_OK;
    };
    if (patch_c == 7 && packed_out) PE_CHECK(f8 ? launch(conv1_tc_kernel<7, true, true>) : launch(conv1_tc_kernel<7, false, true>));
    else if (patch_c == 7) PE_CHECK(f8 ? launch(conv1_tc_kernel<7, true, false>) : launch(conv1_tc_kernel<7, false, false>));
    else if (packed_out) PE_CHECK(f8 ? launch(conv1_tc_kernel<4, true, true>) : launch(conv1_tc_kernel<4, false, true>));
    else PE_CHECK(f8 ? launch(conv1_tc_kernel<4, true, false>) : launch(conv1_tc_kernel<4, false, false>));
    PE_LAUNCHED();
    return PE_OK;
}

/* after a synchronisation: reports (and clears) a failure recorded by the tensor forward kernels */
int forward_tc_check(pe_net* net) {
    if (!net->tc_err) return PE_OK;
    const int e = *reinterpret_cast<volatile int*>(net->tc_err);
    if (e == 0) return PE_OK;
    *net->tc_err = 0;
    if (e == 102 && g_opt_forward_fp8.load() == 1) {
        /* qualified e4m3 cross terms (opt-in) and a later batch beyond |x| = 28: the qualification is withdrawn for
         * these weights; host entry points replay the batch on the fp16 split, device entry points report it */
        net->fp8_state = 2;
        set_error("tensor forward: an input value exceeds 28 with the e4m3 cross terms active (forward_fp8 = 1); the net now "
                  "uses the all-fp16 split - run the batch again");
        return PE_RETRY_FP16;
    }
    if (e >= 100) {
        set_error("tensor forward: %s is not finite or exceeds the split range (patches are expected in [-1, 1]: "
                  "reference src/datasetmanager.cpp:327-343); use pe_set_option(\"%s\", 0) for such data",
                  e == 100 ? "an input value (|x| > 255)" : e == 102 ? "an input value (|x| > 28, fp16 + e4m3 split)" : "an activation (|x| > 4000)",
                  e == 102 ? "forward_fp8" : "forward_tensor");
        return PE_ERR_INVALID;
    }
    set_error("tensor forward: pipeline barrier %d timed out", e);
    return PE_ERR_CUDA;
}

}  // namespace pe
