/* knn_exact.cuh - the matcher's exact fp32 distance, shared by the SIMT and the tensor candidate refinement. */
#pragma once
#include <climits>

namespace pe {

/* The reference's distance, restated operation for operation from the CPU oracle
 * (oracle/oracle_match.c l2_sq): 8 partial sums over dims i%8, multiply and add rounded
 * separately, fixed combine tree, then sqrt. */
__device__ __forceinline__ float exact_l2(const float* __restrict__ a, const float* __restrict__ b, int dp) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    /* dims i .. i + 3 go to acc[i % 8 ..]: two float4 per iteration keep every accumulator index static (registers);
     * the order of operations per accumulator is that of the oracle's loop */
    auto step = [&](int i, int l) {
        const float4 x = *reinterpret_cast<const float4*>(a + i);
        const float4 y = __ldg(reinterpret_cast<const float4*>(b + i));
        const float d0 = __fsub_rn(x.x, y.x), d1 = __fsub_rn(x.y, y.y), d2 = __fsub_rn(x.z, y.z), d3 = __fsub_rn(x.w, y.w);
        acc[l + 0] = __fadd_rn(acc[l + 0], __fmul_rn(d0, d0));
        acc[l + 1] = __fadd_rn(acc[l + 1], __fmul_rn(d1, d1));
        acc[l + 2] = __fadd_rn(acc[l + 2], __fmul_rn(d2, d2));
        acc[l + 3] = __fadd_rn(acc[l + 3], __fmul_rn(d3, d3));
    };
    int i = 0;
    for (; i + 8 <= dp; i += 8) {
        step(i, 0);
        step(i + 4, 4);
    }
    if (i < dp) step(i, 0);          /* dp = 4 (mod 8): the last four dims have i % 8 == 0 */
    const float s = __fadd_rn(__fadd_rn(__fadd_rn(acc[0], acc[4]), __fadd_rn(acc[2], acc[6])),
                              __fadd_rn(__fadd_rn(acc[1], acc[5]), __fadd_rn(acc[3], acc[7])));
    return __fsqrt_rn(s);
}

/* Warp-wide arg-min of (key, id) with distinct ids: returns the lane that holds the smallest pair, -1 when no lane has one.
 * Two redux.sync and a ballot instead of fifteen shuffle steps over three registers.  Float keys compare as floats
 * (-0 = +0: canonicalised before the order-preserving integer map). */
__device__ __forceinline__ int warp_argmin(int key, int id, bool has) {
    const int ko = has ? key : INT_MAX;
    const int kmin = __reduce_min_sync(0xffffffffu, ko);
    const int imin = __reduce_min_sync(0xffffffffu, (has && ko == kmin) ? id : INT_MAX);
    const unsigned win = __ballot_sync(0xffffffffu, has && ko == kmin && id == imin);
    return win ? __ffs(win) - 1 : -1;
}
__device__ __forceinline__ int warp_argmin(float key, int id, bool has) {
    const int b = __float_as_int(key + 0.0f);
    return warp_argmin(b ^ ((b >> 31) & 0x7FFFFFFF), id, has);
}

}  // namespace pe
