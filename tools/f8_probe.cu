/*
 * f8_probe.cu - development probe (not part of the product library): cycles per tcgen05.mma for kind::f16 (K = 16)
 * against kind::f8f6f4 with e4m3 operands (K = 32), M = 128, both operands in shared memory in the K-major
 * no-swizzle layout of umma.cuh, and for the alternating f16 / f8 issue pattern of conv1_tc_kernel<.., true>.
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../3dposeestimation_b200/csrc f8_probe.cu -o f8_probe
 */
#include <cstdio>
#include <cuda_runtime.h>
#include "umma.cuh"
using namespace pe::umma;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

/* mode 0: f16 only, 1: f8 only, 2: alternate f16, f8;  win 0: fixed aligned operands, 1: conv-like moving windows */
template <int N>
__global__ void __launch_bounds__(128) rate_kernel(int reps, int mode, int win, unsigned fill, long long* cycles, int* status) {
    extern __shared__ __align__(1024) unsigned char dyn[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (2 * 73728 + 2 * 32768) / 16; i += 128) reinterpret_cast<uint4*>(dyn)[i] = make_uint4(fill, fill, fill, fill);
    fence_async_smem();
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (warp == 0) {
        const uint32_t idesc = instr_desc(128, N, FMT_F16, FMT_F16);
        const uint32_t s0 = smem_u32(dyn);
        const uint64_t b_a = smem_desc(s0, 16, 128), b_b = smem_desc(s0 + 73728, 16, 128);
        const uint64_t a_a = smem_desc(s0 + 2 * 73728, 15 * 256, 128), a_b = smem_desc(s0 + 2 * 73728 + 32768, 15 * 256, 128);
        const long long t0 = clock64();
#pragma unroll 1
        for (int r = 0; r < reps; r += 4) {
            const uint32_t io = win ? static_cast<uint32_t>((((r >> 2) % 15) * 9 % 72) * 64) : 0u;
            const uint32_t zo = win ? static_cast<uint32_t>((14 - (r >> 2) % 15) * 16) : 0u;
            if (elect_one()) {
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const uint32_t i2 = io + (win ? p * 2 : 0), z2 = zo + (win ? p * 2 * (15 * 16) : 0);
                    const uint32_t d = tmem + (((r + p) >> 7) & 1) * N;
                    const bool f8 = mode == 1 || (mode == 2 && (p & 1));
                    if (f8) mma_f8(d, a_b + z2, b_b + i2, idesc, 1);
                    else mma_f16(d, a_a + z2, b_a + i2, idesc, 1);
                }
            }
            __syncwarp();
        }
        if (elect_one()) mma_commit(&bar);
        __syncwarp();
        const bool ok = mbar_wait(&bar, 0, 1u << 24);
        if (tid == 0) { cycles[0] = clock64() - t0; if (!ok) atomicExch(status, 2); }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

template <int N>
int run(int mode, int win, unsigned fill, long long* d_cycles, int* d_status) {
    const int reps = 4096;
    const size_t smem = 2 * 73728 + 2 * 32768;
    CK(cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int i = 0; i < 2; ++i) { rate_kernel<N><<<1, 128, smem>>>(reps, mode, win, fill, d_cycles, d_status); CK(cudaDeviceSynchronize()); }
    long long c = 0; int st = 0;
    CK(cudaMemcpy(&c, d_cycles, sizeof(c), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&st, d_status, sizeof(st), cudaMemcpyDeviceToHost));
    printf("N=%3d mode=%d (0 f16, 1 f8, 2 alternate) win=%d fill=%08x: %.1f cycles/MMA status=%d\n", N, mode, win, fill, (double)c / reps, st);
    return 0;
}

int main() {
    long long* d_cycles; int* d_status;
    CK(cudaMalloc(&d_cycles, 8)); CK(cudaMalloc(&d_status, 4)); CK(cudaMemset(d_status, 0, 4));
    for (int win = 0; win < 2; ++win)
        for (int mode = 0; mode < 3; ++mode) {
            if (run<256>(mode, win, 0u, d_cycles, d_status)) return 1;
            if (run<128>(mode, win, 0u, d_cycles, d_status)) return 1;
        }
    /* non-zero data (0x38 = 1.0 in e4m3; as fp16 pairs 0x3838 = 0.527) in case the pipe gates on zeros */
    for (int mode = 0; mode < 3; ++mode) if (run<256>(mode, 1, 0x38383838u, d_cycles, d_status)) return 1;
    return 0;
}
