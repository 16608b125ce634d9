/*
 * synth.cu - seeded synthetic patches generated on the device (bench-scale inputs and
 * full-size property tests), and the matching labels.
 *
 * Same recipe and the same counter-based splitmix64 streams as the numpy generator in
 * 3dposeestimation_b200/synth.py (make_patches / make_labels), evaluated with separately
 * rounded float operations so both produce the same values.  Channel order as the reference
 * builds patches (src/datasetmanager.cpp:327-343): 3 colour, depth, 3 normal; labels
 * {class, qx,qy,qz,qw, tx,ty,tz} (src/datasetmanager.cpp:283-289).
 */
#include <cmath>

#include "common.h"

namespace pe {

__host__ __device__ inline uint64_t splitmix64_hd(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t splitmix64(uint64_t x) { return splitmix64_hd(x); }
uint64_t stream_key(uint64_t seed, uint64_t stream) { return splitmix64_hd(seed ^ (stream * 0xD1342543DE82EF95ull)); }

namespace {

__host__ __device__ inline float u01(uint64_t key, uint64_t ctr) {
    return (float)(splitmix64_hd(key + ctr) >> 40) * 5.9604644775390625e-8f; /* 2^-24 */
}

__device__ __forceinline__ float tri(float x) {
    const float f = __fsub_rn(x, floorf(x));
    return __fsub_rn(1.0f, fabsf(__fsub_rn(__fmul_rn(2.0f, f), 1.0f)));
}

__global__ void synth_patches_kernel(uint64_t key_class, uint64_t key_pose, uint64_t key_pix, long long first,
                                     long long n, int channels, int ppo, float* __restrict__ out) {
    const long long gid = blockIdx.x * (long long)blockDim.x + threadIdx.x; /* one thread per pixel */
    if (gid >= n * 4096) return;
    const long long j = gid >> 12;
    const int pix = (int)(gid & 4095), y = pix >> 6, x = pix & 63;
    const long long i = first + j;
    const long long obj = i / ppo;
    const float pp0 = u01(key_pose, (uint64_t)i * 4), pp1 = u01(key_pose, (uint64_t)i * 4 + 1);
    const float r = __fadd_rn(16.0f, __fmul_rn(12.0f, pp0));
    const float dy = __fsub_rn((float)y, 31.5f), dx = __fsub_rn((float)x, 31.5f);
    const bool inside = __fadd_rn(__fmul_rn(dy, dy), __fmul_rn(dx, dx)) <= __fmul_rn(r, r);
    const uint64_t ubase = ((uint64_t)i * 4096 + (uint64_t)pix) * 7;
    float* o = out + gid * channels;
    const float xf = (float)x, yf = (float)y;
    for (int c = 0; c < channels && c < 3; ++c) {
        const float fx = __fadd_rn(0.5f, __fmul_rn(2.5f, u01(key_class, (uint64_t)obj * 8 + 2 * c)));
        const float fy = __fadd_rn(0.5f, __fmul_rn(2.5f, u01(key_class, (uint64_t)obj * 8 + 2 * c + 1)));
        float ph = __fdiv_rn(__fadd_rn(__fmul_rn(xf, fx), __fmul_rn(yf, fy)), 64.0f);
        ph = __fadd_rn(__fadd_rn(ph, pp1), (float)(0.37 * c));
        o[c] = __fadd_rn(__fmul_rn(0.5f, u01(key_pix, ubase + c)), __fmul_rn(0.5f, tri(ph)));
    }
    if (channels > 3) o[3] = inside ? __fadd_rn(0.3f, __fmul_rn(0.4f, u01(key_pix, ubase + 3))) : 0.f;
    if (channels > 4) {
        float v[3];
        for (int c = 0; c < 3; ++c) v[c] = __fsub_rn(__fmul_rn(u01(key_pix, ubase + 4 + c), 2.0f), 1.0f);
        const float nrm = __fadd_rn(__fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(v[0], v[0]), __fmul_rn(v[1], v[1])),
                                                         __fmul_rn(v[2], v[2]))), 1e-6f);
        for (int c = 0; c < 3 && 4 + c < channels; ++c) o[4 + c] = inside ? __fdiv_rn(v[c], nrm) : 0.f;
    }
    for (int c = 7; c < channels; ++c) o[c] = 0.f;
}

}  // namespace

int synth_patches_dev(cudaStream_t stream, uint64_t seed, int64_t first, int64_t n, int channels,
                      int poses_per_object, float* d_out) {
    PE_REQUIRE(channels >= 1 && channels <= 8 && poses_per_object > 0 && n >= 0 && first >= 0,
               "bad synthetic-patch arguments");
    if (n == 0) return PE_OK;
    const long long tot = n * 4096;
    synth_patches_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(
        stream_key(seed, 7), stream_key(seed, 8), stream_key(seed, 9), first, n, channels, poses_per_object, d_out);
    PE_LAUNCHED();
    return PE_OK;
}

/* synth.py make_labels: class = i / ppo; quaternion = Shoemake from stream 20 */
void synth_labels_host(uint64_t seed, int64_t first, int64_t n, int ppo, float* labels) {
    const uint64_t key = stream_key(seed, 20);
    const double two_pi = 2.0 * 3.14159265358979323846;
    for (int64_t j = 0; j < n; ++j) {
        const int64_t i = first + j;
        const double u0 = (double)u01(key, (uint64_t)i * 3), u1 = (double)u01(key, (uint64_t)i * 3 + 1),
                     u2 = (double)u01(key, (uint64_t)i * 3 + 2);
        const double a = std::sqrt(1.0 - u0), b = std::sqrt(u0);
        const double t1 = two_pi * u1, t2 = two_pi * u2;
        float* l = labels + j * 8;
        l[0] = (float)(i / ppo);
        l[1] = (float)(a * std::sin(t1));
        l[2] = (float)(a * std::cos(t1));
        l[3] = (float)(b * std::sin(t2));
        l[4] = (float)(b * std::cos(t2));
        l[5] = l[6] = l[7] = 0.f;
    }
}

}  // namespace pe
