/*
 * umma_probe.cu - development probe for the tcgen05 path (not part of the product library).
 *   1. numerics: a 128-pixel x 16-filter x (8x8x8) implicit-GEMM convolution tile computed with
 *      overlapping no-swizzle K-major descriptors straight out of an HWC8 fp16 image, checked on the host;
 *   2. timing: cycles per tcgen05.mma (M=128, K=16, kind::f16, both operands in shared memory) for
 *      N = 16..256, conv-like moving A windows;
 *   3. timing: tcgen05.ld 32x32b.x32 throughput with 4 warps.
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../3dposeestimation_b200/csrc umma_probe.cu -o umma_probe
 */
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "umma.cuh"

using namespace pe::umma;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

constexpr int kImgPix = 704;   /* 11 rows of 64 pixels, 8 fp16 channels each */
constexpr int kTaps = 64;      /* 8 x 8 */

/* ---- 1. numerics ------------------------------------------------------------------------------ */
__global__ void __launch_bounds__(128) conv_tile_kernel(const __half* __restrict__ img, const __half* __restrict__ wgt,
                                                        float* __restrict__ out, int* __restrict__ status) {
    __shared__ __align__(128) __half s_img[kImgPix * 8];
    __shared__ __align__(128) __half s_w[kTaps * 16 * 8];   /* [tap][n][8 ch] : chunk stride 256 B */
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < kImgPix * 8 / 8; i += 128) reinterpret_cast<uint4*>(s_img)[i] = reinterpret_cast<const uint4*>(img)[i];
    for (int i = tid; i < kTaps * 16 * 8 / 8; i += 128) reinterpret_cast<uint4*>(s_w)[i] = reinterpret_cast<const uint4*>(wgt)[i];
    fence_async_smem();
    if (warp == 0) tmem_alloc<32>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = instr_desc(128, 16, FMT_F16, FMT_F16);
        const uint64_t a0 = smem_desc(smem_u32(s_img), 16, 128);     /* LBO = 1 pixel, SBO = 8 pixels */
        const uint64_t b0 = smem_desc(smem_u32(s_w), 256, 128);      /* LBO = next tap, SBO = next 8 filters */
        uint32_t acc = 0;
        for (int ky = 0; ky < 8; ++ky)
            for (int kp = 0; kp < 4; ++kp) {       /* taps (ky, 2kp) and (ky, 2kp+1) */
                mma_f16(tmem, desc_advance(a0, (ky * 64 + 2 * kp) * 16), desc_advance(b0, (ky * 8 + 2 * kp) * 256), idesc, acc);
                acc = 1;
            }
        mma_commit(&bar);
    }
    const bool ok = mbar_wait(&bar, 0);
    if (!ok && tid == 0) atomicExch(status, 1);
    fence_after_sync();
    if (ok) {
        uint32_t r[16];
        tmem_ld16(tmem + (static_cast<uint32_t>(warp * 32) << 16), r);
        tmem_ld_wait();
        for (int j = 0; j < 16; ++j) out[tid * 16 + j] = __uint_as_float(r[j]);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<32>(tmem);
}

/* ---- 2. MMA issue-rate timing ------------------------------------------------------------------ */
template <int N>
__global__ void __launch_bounds__(128) mma_rate_kernel(int reps, int a_mode, long long* __restrict__ cycles, int* __restrict__ status) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __half* s_a = reinterpret_cast<__half*>(dyn);                 /* 64 KB of zeros: an HWC8 image */
    __half* s_b = reinterpret_cast<__half*>(dyn + 65536);         /* N x 16 fp16, [2 chunks][N][8] */
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (65536 + N * 32) / 16; i += 128) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<256>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    long long t0 = 0, t1 = 0;
    bool ok = true;
    if (tid == 0) {
        const uint32_t idesc = instr_desc(128, N, FMT_F16, FMT_F16);
        const uint64_t a0 = smem_desc(smem_u32(s_a), 16, 128);
        const uint64_t b0 = smem_desc(smem_u32(s_b), N * 16, 128);
        t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            /* a_mode 0: conv-like windows walking over the image; 1: always the same window */
            const uint32_t off = a_mode == 0 ? static_cast<uint32_t>(((r & 7) * 64 + ((r >> 3) & 3) * 2 + ((r >> 5) & 15) * 128) * 16) : 0u;
            mma_f16(tmem, desc_advance(a0, off), b0, idesc, r > 0);
        }
        mma_commit(&bar);
        ok = mbar_wait(&bar, 0);
        t1 = clock64();
        cycles[0] = t1 - t0;
        if (!ok) atomicExch(status, 2);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem);
}

/* ---- 2b. accumulator switching: `per` consecutive MMAs per accumulator, round-robin over nbuf accumulators,
 *          a_lbo selects the A layout (16: overlapping pixels, 2048: chunk-major planes) ------------------- */
template <int N>
__global__ void __launch_bounds__(128) mma_switch_kernel(int reps, int per, int nbuf, int a_lbo, int fresh, long long* __restrict__ cycles, int* __restrict__ status) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __half* s_a = reinterpret_cast<__half*>(dyn);
    __half* s_b = reinterpret_cast<__half*>(dyn + 65536);
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (65536 + N * 32) / 16; i += 128) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = instr_desc(128, N, FMT_F16, FMT_F16);
        const uint64_t a0 = smem_desc(smem_u32(s_a), a_lbo, 128);
        const uint64_t b0 = smem_desc(smem_u32(s_b), N * 16, 128);
        const int lper = 31 - __clz(per);
        const long long t0 = clock64();
#pragma unroll 4
        for (int r = 0; r < reps; ++r) {
            /* per and nbuf are powers of two: no integer division on the issuing thread */
            const int buf = (r >> lper) & (nbuf - 1);
            mma_f16(tmem + buf * N, a0, b0, idesc, fresh ? (r & (per - 1)) != 0 : 1);
        }
        mma_commit(&bar);
        const bool ok = mbar_wait(&bar, 0);
        cycles[0] = clock64() - t0;
        if (!ok) atomicExch(status, 3);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

/* ---- 2c. contention: one thread issues MMAs (4 accumulators of 128 columns, round robin) while `nw` warps
 *          read accumulators with tcgen05.ld.x32 + wait + 32 FMNMX, unsynchronised ------------------------- */
__global__ void __launch_bounds__(320) contention_kernel(int mma_reps, int ld_reps, int nw, long long* __restrict__ cycles, float* __restrict__ sink, int* __restrict__ status) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < (8192 + 128 * 32) / 16; i += 320) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (warp == 1) {
        if (lane == 0 && mma_reps > 0) {
            const uint32_t idesc = instr_desc(128, 128, FMT_F16, FMT_F16);
            const uint64_t a0 = smem_desc(smem_u32(dyn), 2048, 128);
            const uint64_t b0 = smem_desc(smem_u32(dyn + 8192), 2048, 128);
            const long long t0 = clock64();
#pragma unroll 4
            for (int r = 0; r < mma_reps; ++r) mma_f16(tmem + ((r >> 2) & 3) * 128, a0, b0, idesc, (r & 3) != 0);
            mma_commit(&bar);
            const bool ok = mbar_wait(&bar, 0);
            cycles[0] = clock64() - t0;
            if (!ok) atomicExch(status, 4);
        }
    } else if (warp >= 2 && warp < 2 + nw) {
        const uint32_t tb = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
        float m = 0.f;
        const long long t0 = clock64();
        for (int r = 0; r < ld_reps; ++r) {
            const uint32_t base = tb + ((r + (warp >> 2)) & 3) * 128;
#pragma unroll
            for (int c = 0; c < 128; c += 32) {
                uint32_t v[32];
                tmem_ld32(base + c, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) m = fminf(m, __uint_as_float(v[j]));
            }
        }
        const long long t1 = clock64();
        if (lane == 0) cycles[warp] = t1 - t0;
        sink[tid] = m;
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

int run_contention(int mma_reps, int ld_reps, int nw, long long* d_cycles, float* d_sink, int* d_status) {
    const size_t smem = 8192 + 128 * 32;
    CK(cudaMemset(d_cycles, 0, 16 * sizeof(long long)));
    for (int i = 0; i < 2; ++i) {
        contention_kernel<<<1, 320, smem>>>(mma_reps, ld_reps, nw, d_cycles, d_sink, d_status);
        CK(cudaDeviceSynchronize());
    }
    long long c[16];
    CK(cudaMemcpy(c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("contention mma_reps=%5d ld_reps=%4d nw=%d: %.1f cycles/MMA, %.1f cycles per 128x128 accumulator drain (warp 2), %.1f (warp %d)\n",
           mma_reps, ld_reps, nw, mma_reps ? (double)c[0] / mma_reps : 0.0, ld_reps ? (double)c[2] / ld_reps : 0.0,
           ld_reps ? (double)c[1 + nw] / ld_reps : 0.0, 1 + nw);
    return 0;
}


/* ---- 2d. drain variants: `nw` warps each read 128x128 accumulators (their lane quarter) and reduce every 32-column
 *          block to its minimum; an optional MMA thread keeps the tensor pipe busy --------------------------- */
__device__ __forceinline__ float probe_min32(const uint32_t (&v)[32]) {
    float m8[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
        m8[j] = fminf(fminf(__uint_as_float(v[j]), __uint_as_float(v[j + 8])), fminf(__uint_as_float(v[j + 16]), __uint_as_float(v[j + 24])));
    return fminf(fminf(fminf(m8[0], m8[1]), fminf(m8[2], m8[3])), fminf(fminf(m8[4], m8[5]), fminf(m8[6], m8[7])));
}
template <int VARIANT>
__global__ void __launch_bounds__(352) drain_kernel(int mma_reps, int ld_reps, int nw, long long* __restrict__ cycles, float* __restrict__ sink, int* __restrict__ status) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < (8192 + 128 * 32) / 16; i += 352) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (warp == 1) {
        if (lane == 0 && mma_reps > 0) {
            const uint32_t idesc = instr_desc(128, 128, FMT_F16, FMT_F16);
            const uint64_t a0 = smem_desc(smem_u32(dyn), 2048, 128);
            const uint64_t b0 = smem_desc(smem_u32(dyn + 8192), 2048, 128);
            const long long t0 = clock64();
#pragma unroll 4
            for (int r = 0; r < mma_reps; ++r) mma_f16(tmem + ((r >> 2) & 3) * 128, a0, b0, idesc, (r & 3) != 0);
            mma_commit(&bar);
            const bool ok = mbar_wait(&bar, 0);
            cycles[0] = clock64() - t0;
            if (!ok) atomicExch(status, 4);
        }
    } else if (warp >= 3 && warp < 3 + nw) {
        const uint32_t tb = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
        float m = 0.f;
        int hits = 0;
        const long long t0 = clock64();
#pragma unroll 1
        for (int r = 0; r < ld_reps; ++r) {
            const uint32_t base = tb + ((r + (warp >> 2)) & 3) * 128;
            uint32_t v0[32], v1[32], v2[32], v3[32];
            float m0, m1, m2, m3;
            if (VARIANT == 0) {          /* serial: load, wait, reduce, four times */
                tmem_ld32(base, v0); tmem_ld_wait(); m0 = probe_min32(v0);
                tmem_ld32(base + 32, v1); tmem_ld_wait(); m1 = probe_min32(v1);
                tmem_ld32(base + 64, v2); tmem_ld_wait(); m2 = probe_min32(v2);
                tmem_ld32(base + 96, v3); tmem_ld_wait(); m3 = probe_min32(v3);
            } else if (VARIANT == 1) {   /* four loads in flight, one wait */
                tmem_ld32(base, v0); tmem_ld32(base + 32, v1); tmem_ld32(base + 64, v2); tmem_ld32(base + 96, v3);
                tmem_ld_wait();
                m0 = probe_min32(v0); m1 = probe_min32(v1); m2 = probe_min32(v2); m3 = probe_min32(v3);
            } else if (VARIANT == 2) {   /* two and two, first half reduced under the second half's loads */
                tmem_ld32(base, v0); tmem_ld32(base + 32, v1);
                tmem_ld_wait();
                tmem_ld32(base + 64, v2); tmem_ld32(base + 96, v3);
                m0 = probe_min32(v0); m1 = probe_min32(v1);
                tmem_ld_wait();
                m2 = probe_min32(v2); m3 = probe_min32(v3);
            } else {                     /* one at a time, each reduced under the next load */
                tmem_ld32(base, v0); tmem_ld_wait();
                tmem_ld32(base + 32, v1); m0 = probe_min32(v0); tmem_ld_wait();
                tmem_ld32(base + 64, v2); m1 = probe_min32(v1); tmem_ld_wait();
                tmem_ld32(base + 96, v3); m2 = probe_min32(v2); tmem_ld_wait();
                m3 = probe_min32(v3);
            }
            const float mm = fminf(fminf(m0, m1), fminf(m2, m3));
            if (__any_sync(0xffffffffu, mm < -1.f)) ++hits;
            m = fminf(m, mm);
        }
        const long long t1 = clock64();
        if (lane == 0) cycles[warp] = t1 - t0;
        sink[tid] = m + hits;
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

template <int VARIANT>
int run_drain(int mma_reps, int ld_reps, int nw, long long* d_cycles, float* d_sink, int* d_status) {
    const size_t smem = 8192 + 128 * 32;
    CK(cudaMemset(d_cycles, 0, 16 * sizeof(long long)));
    for (int i = 0; i < 2; ++i) {
        drain_kernel<VARIANT><<<1, 352, smem>>>(mma_reps, ld_reps, nw, d_cycles, d_sink, d_status);
        CK(cudaDeviceSynchronize());
    }
    long long c[16];
    CK(cudaMemcpy(c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("drain variant %d mma_reps=%5d ld_reps=%4d nw=%d: %.1f cycles/MMA, %.1f cycles per 128x128 drain (warp 3), %.1f (warp %d)\n",
           VARIANT, mma_reps, ld_reps, nw, mma_reps ? (double)c[0] / mma_reps : 0.0, ld_reps ? (double)c[3] / ld_reps : 0.0,
           ld_reps ? (double)c[2 + nw] / ld_reps : 0.0, 2 + nw);
    return 0;
}

template <int N>
int run_switch(int per, int nbuf, int a_lbo, int fresh, long long* d_cycles, int* d_status) {
    const int reps = 2048;
    const size_t smem = 65536 + N * 32;
    CK(cudaFuncSetAttribute(mma_switch_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int i = 0; i < 2; ++i) {
        mma_switch_kernel<N><<<1, 128, smem>>>(reps, per, nbuf, a_lbo, fresh, d_cycles, d_status);
        CK(cudaDeviceSynchronize());
    }
    long long c = 0;
    CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("mma_switch N=%3d per=%2d nbuf=%d a_lbo=%4d fresh=%d: %.1f cycles/MMA\n", N, per, nbuf, a_lbo, fresh, (double)c / reps);
    return 0;
}

/* ---- 3. TMEM load throughput -------------------------------------------------------------------- */
__global__ void __launch_bounds__(128) tmem_ld_rate_kernel(int reps, long long* __restrict__ cycles, float* __restrict__ sink) {
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc<256>(&tmem_slot);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot + (static_cast<uint32_t>(warp * 32) << 16);
    float m = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int c = 0; c < 256; c += 32) {
            uint32_t v[32];
            tmem_ld32(tmem + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) m = fminf(m, __uint_as_float(v[j]));
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) cycles[0] = t1 - t0;
    sink[tid] = m;
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem_slot);
}

/* ---- 2c. roles swapped: A = 128 weight rows (aligned), B = N consecutive pixels of the HWC8 image (sliding, 32-byte
 *          misaligned windows).  b_mode 0: conv-like windows, 1: always the same aligned window -------------- */
template <int N>
__global__ void __launch_bounds__(128) mma_rate_swap_kernel(int reps, int b_mode, int a_var, long long* __restrict__ cycles, int* __restrict__ status) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __half* s_img = reinterpret_cast<__half*>(dyn);               /* 64 KB of zeros: an HWC8 image */
    __half* s_w = reinterpret_cast<__half*>(dyn + 65536);         /* 8 x (128 x 16 fp16, [2 chunks][128][8]) */
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (65536 + 8 * 4096) / 16; i += 128) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = instr_desc(128, N, FMT_F16, FMT_F16);
        const uint64_t a0 = smem_desc(smem_u32(s_w), 128 * 16, 128);
        const uint64_t b0 = smem_desc(smem_u32(s_img), 16, 128);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const uint32_t off = b_mode == 0 ? static_cast<uint32_t>(((r & 7) * 64 + ((r >> 3) & 3) * 2 + ((r >> 5) & 15) * 128) * 16) : 0u;
            const uint32_t aoff = a_var > 0 ? static_cast<uint32_t>(((r >> (a_var - 1)) & 7) << 12) : 0u;   /* a_var 1: new A every MMA, 2: every 2, 3: every 4 */
            mma_f16(tmem + ((r >> 7) & 1) * N, desc_advance(a0, aoff), desc_advance(b0, off), idesc, (r & 127) > 0);
        }
        mma_commit(&bar);
        const bool ok = mbar_wait(&bar, 0);
        cycles[0] = clock64() - t0;
        if (!ok) atomicExch(status, 2);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

template <int N>
int run_rate_swap(int b_mode, int a_var, long long* d_cycles, int* d_status) {
    const int reps = 2048;
    const size_t smem = 65536 + 8 * 4096;
    CK(cudaFuncSetAttribute(mma_rate_swap_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int i = 0; i < 2; ++i) {
        mma_rate_swap_kernel<N><<<1, 128, smem>>>(reps, b_mode, a_var, d_cycles, d_status);
        CK(cudaDeviceSynchronize());
    }
    long long c = 0;
    CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("mma_rate_swap (A = weights, B = image window) N=%3d b_mode=%d a_var=%d: %.1f cycles/MMA\n", N, b_mode, a_var, (double)c / reps);
    return 0;
}

/* ---- 2d. the conv1 issue pattern (A = Toeplitz weight windows, B = 256-pixel image windows, three split products),
 *          with parts of the address pattern frozen by `flags`: 1 A always the hi array, 2 B always the hi image,
 *          4 p = 0 only (aligned image windows), 8 ky = 0 only, 16 A window fixed, 32 p advances the A window only ---- */
__global__ void __launch_bounds__(128) mma_conv_pattern_kernel(int tiles, int flags, long long* __restrict__ cycles, int* __restrict__ status) {
    extern __shared__ __align__(1024) unsigned char dyn[];
    constexpr int kImg = 72 * 1024, kZ = 127 * 256, kZStride = 15 * 256;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (2 * kImg + 2 * kZ) / 16; i += 128) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (warp == 0) {
        const uint32_t idesc = instr_desc(128, 256, FMT_F16, FMT_F16);
        const uint32_t s0 = smem_u32(dyn);
        const uint64_t i_hi = smem_desc(s0, 16, 128), i_lo = (flags & 2) ? i_hi : smem_desc(s0 + kImg, 16, 128);
        const uint64_t z_hi = smem_desc(s0 + 2 * kImg, kZStride, 128), z_lo = (flags & 1) ? z_hi : smem_desc(s0 + 2 * kImg + kZ, kZStride, 128);
        const long long t0 = clock64();
        for (int t = 0; t < tiles; ++t) {
            const uint32_t d = tmem + (t & 1) * 256;
#pragma unroll 1
            for (int ky = 0; ky < 15; ++ky) {
                const int kye = (flags & 8) ? 0 : ky;
                const int y = 32 * (t & 1) + kye;
                const uint32_t i_row = static_cast<uint32_t>((y & 7) * 9 + (y >> 3)) * 64;
                const uint32_t z_row = (flags & 16) ? 0u : static_cast<uint32_t>(14 - kye) * 16;
                if (elect_one()) {
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        const uint32_t io = i_row + ((flags & (4 | 32)) ? 0 : p * 2), zo = z_row + ((flags & 16) ? 0 : p * 2 * (kZStride >> 4));
                        mma_f16(d, z_hi + zo, i_hi + io, idesc, (ky | p) != 0);
                        mma_f16(d, z_lo + zo, i_hi + io, idesc, 1);
                        mma_f16(d, z_hi + zo, i_lo + io, idesc, 1);
                    }
                }
                __syncwarp();
            }
        }
        if (elect_one()) mma_commit(&bar);
        __syncwarp();
        const bool ok = mbar_wait(&bar, 0);
        if (tid == 0) {
            cycles[0] = clock64() - t0;
            if (!ok) atomicExch(status, 2);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

int run_conv_pattern(int flags, long long* d_cycles, int* d_status) {
    const int tiles = 16;
    const size_t smem = 2 * 72 * 1024 + 2 * 127 * 256;
    CK(cudaFuncSetAttribute(mma_conv_pattern_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int i = 0; i < 2; ++i) {
        mma_conv_pattern_kernel<<<1, 128, smem>>>(tiles, flags, d_cycles, d_status);
        CK(cudaDeviceSynchronize());
    }
    long long c = 0;
    CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("conv1 pattern flags=%2d: %.1f cycles/MMA (128x256x16)\n", flags, (double)c / (tiles * 180));
    return 0;
}

template <int N>
int run_rate(int a_mode, long long* d_cycles, int* d_status) {
    const int reps = 2048;
    const size_t smem = 65536 + N * 32;
    CK(cudaFuncSetAttribute(mma_rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mma_rate_kernel<N><<<1, 128, smem>>>(reps, a_mode, d_cycles, d_status);
    CK(cudaDeviceSynchronize());
    mma_rate_kernel<N><<<1, 128, smem>>>(reps, a_mode, d_cycles, d_status);
    CK(cudaDeviceSynchronize());
    long long c = 0;
    CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    printf("mma_rate N=%3d a_mode=%d: %.1f cycles/MMA (%lld cycles / %d)\n", N, a_mode, (double)c / reps, c, reps);
    return 0;
}

int main() {
    int* d_status;
    CK(cudaMalloc(&d_status, sizeof(int)));
    CK(cudaMemset(d_status, 0, sizeof(int)));
    /* ---- numerics */
    std::vector<__half> img(kImgPix * 8), wgt(kTaps * 16 * 8);
    std::vector<float> imgf(kImgPix * 8), wf(kTaps * 16 * 8);
    unsigned s = 12345u;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 65536.0f - 0.5f; };
    for (size_t i = 0; i < img.size(); ++i) { img[i] = __float2half_rn(rnd()); imgf[i] = __half2float(img[i]); }
    for (size_t i = 0; i < wgt.size(); ++i) { wgt[i] = __float2half_rn(rnd() * 0.25f); wf[i] = __half2float(wgt[i]); }
    __half *d_img, *d_w;
    float* d_out;
    CK(cudaMalloc(&d_img, img.size() * 2));
    CK(cudaMalloc(&d_w, wgt.size() * 2));
    CK(cudaMalloc(&d_out, 128 * 16 * 4));
    CK(cudaMemcpy(d_img, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_w, wgt.data(), wgt.size() * 2, cudaMemcpyHostToDevice));
    conv_tile_kernel<<<1, 128>>>(d_img, d_w, d_out, d_status);
    CK(cudaDeviceSynchronize());
    std::vector<float> out(128 * 16);
    CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost));
    int st = 0;
    CK(cudaMemcpy(&st, d_status, sizeof(int), cudaMemcpyDeviceToHost));
    double max_err = 0, max_ref = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 16; ++n) {
            double acc = 0;
            for (int ky = 0; ky < 8; ++ky)
                for (int kx = 0; kx < 8; ++kx)
                    for (int c = 0; c < 8; ++c)
                        acc += (double)imgf[(m + ky * 64 + kx) * 8 + c] * (double)wf[((ky * 8 + kx) * 16 + n) * 8 + c];
            max_err = fmax(max_err, fabs(acc - out[m * 16 + n]));
            max_ref = fmax(max_ref, fabs(acc));
        }
    printf("conv_tile: status=%d max_abs_err=%.3e max_ref=%.3e -> %s\n", st, max_err, max_ref, (st == 0 && max_err < 1e-4 * max_ref) ? "OK" : "MISMATCH");
    /* ---- rates */
    long long* d_cycles;
    CK(cudaMalloc(&d_cycles, sizeof(long long)));
    if (st == 0) {
        for (int a_mode = 0; a_mode < 2; ++a_mode) {
            if (run_rate<16>(a_mode, d_cycles, d_status)) return 1;
            if (run_rate<32>(a_mode, d_cycles, d_status)) return 1;
            if (run_rate<64>(a_mode, d_cycles, d_status)) return 1;
            if (run_rate<128>(a_mode, d_cycles, d_status)) return 1;
            if (run_rate<256>(a_mode, d_cycles, d_status)) return 1;
        }
        for (int flags : {0, 1, 2, 3, 4, 8, 16, 32, 4 | 16, 1 | 2 | 4 | 8 | 16})
            if (run_conv_pattern(flags, d_cycles, d_status)) return 1;
        for (int b_mode = 0; b_mode < 1; ++b_mode) {
            for (int a_var : {0, 1, 2, 3}) {
                if (run_rate_swap<128>(b_mode, a_var, d_cycles, d_status)) return 1;
                if (run_rate_swap<256>(b_mode, a_var, d_cycles, d_status)) return 1;
            }
        }
        for (int a_lbo : {2048})
            for (int per : {4})
                if (run_switch<128>(per, 4, a_lbo, 1, d_cycles, d_status)) return 1;
        if (run_switch<128>(4, 4, 2048, 0, d_cycles, d_status)) return 1;
        if (run_switch<128>(4, 1, 2048, 1, d_cycles, d_status)) return 1;
        if (run_switch<256>(4, 2, 2048, 1, d_cycles, d_status)) return 1;
        if (run_switch<256>(1, 2, 2048, 1, d_cycles, d_status)) return 1;
        if (run_switch<64>(4, 8, 2048, 1, d_cycles, d_status)) return 1;
        if (run_switch<64>(1, 8, 2048, 1, d_cycles, d_status)) return 1;
        float* d_sink;
        CK(cudaMalloc(&d_sink, 352 * 4));
        {
            long long* d_c16;
            CK(cudaMalloc(&d_c16, 16 * sizeof(long long)));
            if (run_contention(4096, 0, 0, d_c16, d_sink, d_status)) return 1;
            if (run_contention(0, 512, 4, d_c16, d_sink, d_status)) return 1;
            if (run_contention(0, 512, 8, d_c16, d_sink, d_status)) return 1;
            if (run_contention(8192, 512, 4, d_c16, d_sink, d_status)) return 1;
            if (run_contention(16384, 512, 8, d_c16, d_sink, d_status)) return 1;
            for (int nw : {1, 4, 8}) {
                if (run_drain<0>(0, 512, nw, d_c16, d_sink, d_status)) return 1;
                if (run_drain<1>(0, 512, nw, d_c16, d_sink, d_status)) return 1;
                if (run_drain<2>(0, 512, nw, d_c16, d_sink, d_status)) return 1;
                if (run_drain<3>(0, 512, nw, d_c16, d_sink, d_status)) return 1;
            }
            if (run_drain<1>(8 * 512, 512, 8, d_c16, d_sink, d_status)) return 1;
            if (run_drain<2>(8 * 512, 512, 8, d_c16, d_sink, d_status)) return 1;
            if (run_drain<3>(8 * 512, 512, 8, d_c16, d_sink, d_status)) return 1;
        }
        tmem_ld_rate_kernel<<<1, 128>>>(256, d_cycles, d_sink);
        CK(cudaDeviceSynchronize());
        long long c = 0;
        CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
        printf("tmem_ld: %.1f cycles per 128x256 fp32 tile read by 4 warps (incl. 256 FMNMX/thread)\n", (double)c / 256);
    }
    CK(cudaMemcpy(&st, d_status, sizeof(int), cudaMemcpyDeviceToHost));
    printf("final status=%d\n", st);
    return 0;
}
