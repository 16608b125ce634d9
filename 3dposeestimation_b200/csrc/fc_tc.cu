/*
 * fc_tc.cu - InnerProduct layers (ip1, ip2 and the bin head) as batched tcgen05 GEMMs with fused bias + activation.
 *
 * Replaces Caffe's InnerProduct + ReLU/ELU/PReLU inside `CNN.Forward()` (reference src/networkevaluator.cpp:45;
 * network/manifold_rgbdnor_16.prototxt:75-108, manifold_rgbdnor_16_bin.prototxt:113-154):
 *     Y[m][o] = act(bias[o] + sum_k X[m][k] W[o][k])
 * M = patches of the chunk (64 per CTA, issued as M = 128 MMAs), N = outputs (128 per CTA), K = fan-in in blocks of 64.
 * Both operands K-major, no swizzle (umma.cuh): a K block is 8 chunks of [128 rows][8 halves].  Weights are
 * packed once on the host (fp16 hi / lo, power-of-two scale) and arrive by one 32 KB cp.async.bulk per K block;
 * activations are fp32 in HBM and are converted (scale 2^4, hi / lo) by the four loader warps, which also drain
 * the accumulator at the end.  fp32 accuracy as in conv_tc.cu: D = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi.
 */
#include <cfloat>
#include <cmath>
#include <cstring>

#include "common.h"
#include "umma.cuh"

namespace pe {

using namespace umma;

namespace {

constexpr int kFcKB = 64;                          /* k values per block = 4 MMA K steps */
constexpr int kFcStages = 3;
constexpr int kFcTerm = 8 * 2048;                  /* one term of one operand block: 8 chunks x 128 rows x 16 B */
constexpr int kFcStageBytes = 4 * kFcTerm;         /* A hi | A lo | B hi | B lo */
constexpr int kFcOffBar = kFcStages * kFcStageBytes;
constexpr int kFcOffBias = kFcOffBar + 256;           /* bias[128] | slope[128] of this CTA's N tile */
constexpr int kFcSmem = kFcOffBias + 1024;
constexpr int kFcThreads = 320;                    /* warp 0 weight copies, warp 1 MMA, warps 2-9 loaders + epilogue */
constexpr int kLog2SxFc = 4;                       /* activation scale 2^4: |x| < 4094 */
constexpr int kFcRows = 64;                        /* batch rows per CTA (the MMA is M = 128, see the kernel) */

__device__ __forceinline__ float fc_act(float v, int act, float slope) {
    const float pos = fmaxf(v, 0.f), neg = fminf(v, 0.f);
    if (act == PE_ACT_RELU) return pos;
    if (act == PE_ACT_ELU) return pos + (expf(neg) - 1.0f);
    if (act == PE_ACT_PRELU) return pos + slope * neg;
    return v;
}
__device__ __forceinline__ uint32_t pk(__half a, __half b) {
    return static_cast<uint32_t>(__half_as_ushort(a)) | (static_cast<uint32_t>(__half_as_ushort(b)) << 16);
}

struct FcArgs {
    const float* x;            /* [M][ldx] fp32, columns >= K are zero or absent */
    int ldx, M, K;             /* K: valid fan-in (multiple of 4) */
    const uint4* wp;           /* [n_tiles][k_blocks][A-like block: hi 16 KB | lo 16 KB] */
    int k_blocks;
    const float* bias;
    const float* slope;
    int act;
    float inv_scale;
    float* y;
    int ldy, N, n_store;       /* columns N..n_store-1 of an intermediate are written as zeros */
    int* err;
};

__global__ void __launch_bounds__(kFcThreads, 1) fc_tc_kernel(const FcArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kFcOffBar);
    uint64_t* full_a = bars;                 /* [stages] activations converted (128 loader threads) */
    uint64_t* full_b = bars + kFcStages;     /* [stages] weight block landed (tx bytes) */
    uint64_t* empty = bars + 2 * kFcStages;  /* [stages] MMAs of the block completed */
    uint64_t* acc_full = bars + 3 * kFcStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    /* a CTA owns kFcRows = 64 rows of the batch but issues M = 128 MMAs: a row of D depends on the same row of A only, so
     * accumulator lanes 64..127 hold whatever the never-written upper half of the A tile produces and are never read.
     * 64-row tiles put (M / 64) x (N / 128) = 128 CTAs on the 148 SMs for the 4096-patch batch (128-row tiles: 64) and
     * halve what a loader thread converts per K block - the kernel is bound by that load -> convert -> store chain,
     * not by the tensor pipe (12 MMAs of 64 cycles per K block). */
    const int m0 = blockIdx.x * kFcRows, n0 = blockIdx.y * 128;
    int n_mma = a.N - n0;
    n_mma = n_mma >= 128 ? 128 : ((n_mma + 15) & ~15);
    float* s_bias = reinterpret_cast<float*>(smem + kFcOffBias);
    float* s_slope = s_bias + 128;
    if (tid < 128) {
        const int o = n0 + tid;
        s_bias[tid] = o < a.N ? a.bias[o] : 0.f;
        s_slope[tid] = (a.slope && o < a.N) ? a.slope[o] : 0.f;
    }
    if (tid == 0) {
        for (int s = 0; s < kFcStages; ++s) { mbar_init(&full_a[s], 256); mbar_init(&full_b[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(acc_full, 1);
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc<256>(tmem_slot);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const uint4* src = a.wp + (size_t)blockIdx.y * a.k_blocks * (2 * kFcTerm / 16);
            for (int kb = 0; kb < a.k_blocks; ++kb) {
                const int s = kb % kFcStages;
                if (!mbar_wait(&empty[s], ((kb / kFcStages) & 1) ^ 1)) { atomicExch(a.err, 11); break; }
                mbar_expect_tx(&full_b[s], 2 * kFcTerm);
                bulk_g2s(smem + s * kFcStageBytes + 2 * kFcTerm, src + (size_t)kb * (2 * kFcTerm / 16), 2 * kFcTerm, &full_b[s]);
            }
        }
    } else if (warp == 1) {
        /* the whole warp runs the loop (warp-uniform operands), one elected lane issues */
        const uint32_t idesc = instr_desc(128, n_mma, FMT_F16, FMT_F16);
        bool ok = true;
        for (int kb = 0; kb < a.k_blocks; ++kb) {
            const int s = kb % kFcStages;
            if (!mbar_wait(&full_a[s], (kb / kFcStages) & 1)) { if (lane == 0) atomicExch(a.err, 12); ok = false; break; }
            if (!mbar_wait(&full_b[s], (kb / kFcStages) & 1)) { if (lane == 0) atomicExch(a.err, 13); ok = false; break; }
            fence_after_sync();
            const uint32_t base = smem_u32(smem + s * kFcStageBytes);
            const uint64_t a_hi = smem_desc(base, 2048, 128), a_lo = smem_desc(base + kFcTerm, 2048, 128);
            const uint64_t b_hi = smem_desc(base + 2 * kFcTerm, 2048, 128), b_lo = smem_desc(base + 3 * kFcTerm, 2048, 128);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint32_t o = ks * 2 * (2048 >> 4);
                    /* two accumulators: the hi.hi chain (columns 0..127) and the cross terms (128..255), summed in the
                     * epilogue with round-to-nearest.  The tensor core truncates every accumulate (round toward zero,
                     * ~2^-24 of the accumulator's magnitude per MMA); kept apart, the 2^-11-sized cross terms neither
                     * suffer nor cause truncation at the main sum's magnitude: 76 instead of 228 chained truncations */
                    mma_f16(tmem, a_hi + o, b_hi + o, idesc, (kb | ks) != 0);
                    mma_f16(tmem + 128, a_hi + o, b_lo + o, idesc, (kb | ks) != 0);
                    mma_f16(tmem + 128, a_lo + o, b_hi + o, idesc, 1);
                }
                mma_commit(&empty[s]);
            }
            __syncwarp();
        }
        if (ok && elect_one()) mma_commit(acc_full);
        __syncwarp();
    } else {
        /* ---- loaders: four threads per row m0 + r (k quarters of 16), one K block at a time.  The loads of block kb + 1 are
         * in flight while block kb is converted (two register buffers): a row is 4.7 KB away from the next one, so every
         * load is its own cache line and its latency would otherwise be paid 19 times in a row ---- */
        const int lt = tid - 64;                  /* 0..255 */
        const int r = lt & 63, kq = lt >> 6;
        const int m = m0 + r;
        const bool row_ok = m < a.M;
        const float sx = static_cast<float>(1 << kLog2SxFc);
        const float* xrow = a.x + (size_t)(row_ok ? m : 0) * a.ldx + 16 * kq;
        float vmax = 0.f;
        bool ok = true;
        auto load = [&](float4 (&v)[4], int kb) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = kb * kFcKB + 16 * kq + 4 * i;
                v[i] = (row_ok && k < a.K) ? *reinterpret_cast<const float4*>(xrow + kb * kFcKB + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        auto convert = [&](const float4 (&v)[4], int kb) {
            const int s = kb % kFcStages;
            if (!mbar_wait(&empty[s], ((kb / kFcStages) & 1) ^ 1)) { atomicExch(a.err, 14); ok = false; return; }
            uint4* dst_hi = reinterpret_cast<uint4*>(smem + s * kFcStageBytes) + (2 * kq) * 128 + r;
            uint4* dst_lo = dst_hi + kFcTerm / 16;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const float e[8] = {v[2 * c].x, v[2 * c].y, v[2 * c].z, v[2 * c].w, v[2 * c + 1].x, v[2 * c + 1].y, v[2 * c + 1].z, v[2 * c + 1].w};
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    __half h0, l0, h1, l1;
                    vmax = fmaxf(vmax, fmaxf(fabsf(e[2 * j]), fabsf(e[2 * j + 1])));
                    split_f16(e[2 * j] * sx, h0, l0);
                    split_f16(e[2 * j + 1] * sx, h1, l1);
                    hi[j] = pk(h0, h1);
                    lo[j] = pk(l0, l1);
                }
                dst_hi[c * 128] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                dst_lo[c * 128] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            fence_async_smem();
            mbar_arrive(&full_a[s]);
        };
        float4 va[4], vb[4];
        load(va, 0);
#pragma unroll 1
        for (int kb = 0; kb < a.k_blocks && ok; kb += 2) {
            if (kb + 1 < a.k_blocks) load(vb, kb + 1);
            convert(va, kb);
            if (kb + 1 < a.k_blocks && ok) {
                if (kb + 2 < a.k_blocks) load(va, kb + 2);
                convert(vb, kb + 1);
            }
        }
        if (!(vmax <= 4000.f)) atomicExch(a.err, 101);
        /* ---- epilogue: accumulator row er -> bias, activation, store.  Rows 0..63 live in TMEM lanes 0..63, which only
         * warps with warp % 4 in {0, 1} may read: warps 4, 5 take the first half of the columns, warps 8, 9 the second ---- */
        if ((warp & 3) < 2) {
            const int er = (warp & 3) * 32 + lane, half = (warp - 2) >> 2;
            const int em = m0 + er;
            const bool erow_ok = em < a.M;
            if (ok && mbar_wait(acc_full, 0)) {
                fence_after_sync();
                const uint32_t tbase = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
                float* yrow = a.y + (size_t)(erow_ok ? em : 0) * a.ldy + n0;
                const bool vec_ok = (a.ldy & 3) == 0 && (reinterpret_cast<uintptr_t>(a.y) & 15) == 0;
                const int groups = n_mma / 16, g_split = (groups + 1) / 2;
                for (int c0 = (half ? g_split : 0) * 16; c0 < (half ? groups : g_split) * 16; c0 += 16) {
                    uint32_t w[16], wx[16];
                    tmem_ld16(tbase + c0, w);
                    tmem_ld16(tbase + 128 + c0, wx);
                    tmem_ld_wait();
                    float r[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        r[j] = fc_act(fmaf(__uint_as_float(w[j]) + __uint_as_float(wx[j]), a.inv_scale, s_bias[c0 + j]), a.act, s_slope[c0 + j]);
                    if (erow_ok) {
                        if (n0 + c0 + 16 <= a.N && vec_ok) {          /* a whole group in range: four 16-byte stores */
#pragma unroll
                            for (int j = 0; j < 16; j += 4)
                                *reinterpret_cast<float4*>(yrow + c0 + j) = make_float4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int o = n0 + c0 + j;
                                if (o < a.N) yrow[c0 + j] = r[j];
                                else if (o < a.n_store) yrow[c0 + j] = 0.f;
                            }
                        }
                    }
                }
            } else if (ok) {
                atomicExch(a.err, 15);
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem);
}

}  // namespace

/* packs one InnerProduct's weights [out][in] into per (N tile, K block) operand blocks (fp16 hi | lo) */
int fc_tc_prepare(pe_net* net, FcPlan* fc) {
    fc->tc_ready = false;
    const std::vector<float>& w = net->blobs[fc->blob_w];
    float wmax = 0.f;
    for (float v : w) wmax = std::fmax(wmax, std::fabs(v));
    if (!(wmax > 0.f) || !std::isfinite(wmax)) return PE_OK;
    const int log2_sw = 10 - std::ilogb(wmax);
    const float sw = std::ldexp(1.f, log2_sw);
    const int n_tiles = (fc->out + 127) / 128, k_blocks = (fc->ld_in + kFcKB - 1) / kFcKB;
    const size_t block_halves = (size_t)2 * kFcTerm / 2;
    std::vector<uint16_t> p((size_t)n_tiles * k_blocks * block_halves, 0);
    for (int o = 0; o < fc->out; ++o)
        for (int i = 0; i < fc->in; ++i) {
            const float v = w[(size_t)o * fc->in + i] * sw;
            const __half h = __float2half_rn(v);
            const __half l = __float2half_rn(v - __half2float(h));
            uint16_t hb, lb;
            memcpy(&hb, &h, 2);
            memcpy(&lb, &l, 2);
            const int nt = o / 128, row = o % 128, kb = i / kFcKB, c = (i % kFcKB) / 8, e = i % 8;
            const size_t at = ((size_t)nt * k_blocks + kb) * block_halves + ((size_t)c * 128 + row) * 8 + e;
            p[at] = hb;
            p[at + kFcTerm / 2] = lb;
        }
    PE_CUDA(cudaSetDevice(net->device));
    if (fc->d_wp) { cudaFree(fc->d_wp); fc->d_wp = nullptr; }
    PE_CUDA(cudaMalloc(&fc->d_wp, p.size() * 2));
    PE_CUDA(cudaMemcpy(fc->d_wp, p.data(), p.size() * 2, cudaMemcpyHostToDevice));
    if (!net->tc_err) {
        PE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&net->tc_err), sizeof(int), cudaHostAllocMapped));
        *net->tc_err = 0;
    }
    fc->k_blocks = k_blocks;
    fc->inv_scale = std::ldexp(1.f, -(log2_sw + kLog2SxFc));
    fc->tc_ready = true;
    return PE_OK;
}

int fc_tc_launch(pe_net* net, const FcPlan& fc, const float* x, int ldx, float* y, int ldy, int n_store, int m) {
    FcArgs a;
    a.x = x; a.ldx = ldx; a.M = m; a.K = fc.ld_in;
    a.wp = reinterpret_cast<const uint4*>(fc.d_wp);
    a.k_blocks = fc.k_blocks;
    a.bias = fc.d_b; a.slope = fc.d_slope; a.act = fc.act;
    a.inv_scale = fc.inv_scale;
    a.y = y; a.ldy = ldy; a.N = fc.out; a.n_store = n_store;
    a.err = net->tc_err;
    static bool attr_done[16] = {};
    if (!attr_done[net->device & 15]) {
        PE_CUDA(cudaFuncSetAttribute(fc_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFcSmem));
        attr_done[net->device & 15] = true;
    }
    dim3 grid((unsigned)((m + kFcRows - 1) / kFcRows), (unsigned)((fc.out + 127) / 128));
    fc_tc_kernel<<<grid, kFcThreads, kFcSmem, net->stream>>>(a);
    PE_LAUNCHED();
    return PE_OK;
}

}  // namespace pe
