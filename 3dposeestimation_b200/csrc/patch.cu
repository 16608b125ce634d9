/*
 * patch.cu - patch extraction on the device: the producer side of the descriptor forward.
 *
 *   depth2normals           reference src/helper.cpp:53-103   (3x3 stencil of step 3, least-squares surface normal)
 *   samplePatchWithScale    reference src/datasetmanager.cpp:299-346 (metric 0.2 m cut around the projected centroid,
 *                           colour / 255, depth de-meaned / clamped / rescaled, cv::resize to 64x64: INTER_LINEAR for
 *                           colour and normals, INTER_NEAREST for depth; channels {B,G,R, depth, nx,ny,nz})
 *
 * The output is exactly the [n][64][64][7] fp32 HWC layout pe_compute_descriptors_dev consumes, so frames go from
 * camera / renderer buffers to descriptors without leaving the GPU.  The sampler can compute the normals it needs
 * on the fly (4 taps per output pixel, 9 depth reads each, same arithmetic as the full-frame kernel): for one
 * 64x64 patch per 640x480 frame that is 16 k normals instead of 307 k and no normals image in HBM.
 *
 * Arithmetic follows the oracle (oracle/oracle_patch.c) operation for operation: products and sums rounded
 * separately (__fmul_rn / __fadd_rn, no FMA contraction), cv::resize's coordinate rules restated in its header.
 */
#include <cstdint>

#include "common.h"

namespace pe {

namespace {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

/* src/helper.cpp:64-97 for one pixel; zero outside the interior, for d <= 0 and for a degenerate system */
__device__ __forceinline__ float3 normal_at(const float* __restrict__ dep, int rows, int cols, int r, int c, float fx, float fy) {
    const int N = 3;
    float3 out = make_float3(0.f, 0.f, 0.f);
    if (r < N || r >= rows - N - 1 || c < N || c >= cols - N - 1) return out;
    const float* p = dep + (size_t)r * cols + c;
    const float d = __ldg(p);
    if (!(d > 0.f)) return out;
    float A0 = 0.f, A1 = 0.f, A3 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll
    for (int i = -N; i <= N; i += N)
#pragma unroll
        for (int j = -N; j <= N; j += N) {
            const float delta = sub(__ldg(p + i + j * cols), d);
            const float f = fabsf(delta) < 0.05f ? 1.f : 0.f;
            const float fi = mul(f, (float)i), fj = mul(f, (float)j);
            A0 = add(A0, mul(fi, (float)i));
            A1 = add(A1, mul(fi, (float)j));
            A3 = add(A3, mul(fj, (float)j));
            b0 = add(b0, mul(fi, delta));
            b1 = add(b1, mul(fj, delta));
        }
    const float n0 = mul(sub(mul(A3, b0), mul(A1, b1)), fx);
    const float n1 = mul(add(mul(-A1, b0), mul(A0, b1)), fy);
    const float n2 = mul(sub(mul(A0, A3), mul(A1, A1)), d);
    const float sq = add(add(mul(n0, n0), mul(n1, n1)), mul(n2, n2));
    if (sq > 0.f) {
        const float s = __fsqrt_rn(sq);
        out = make_float3(__fdiv_rn(n0, s), __fdiv_rn(n1, s), __fdiv_rn(n2, s));
    }
    return out;
}

__global__ void depth2normals_kernel(const float* __restrict__ dep, int n_frames, int rows, int cols, float fx, float fy,
                                     float* __restrict__ nor) {
    const long long px = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long per = (long long)rows * cols;
    if (px >= per * n_frames) return;
    const long long f = px / per;
    const int r = (int)((px - f * per) / cols), c = (int)((px - f * per) % cols);
    const float3 n = normal_at(dep + f * per, rows, cols, r, c, fx, fy);
    float* o = nor + px * 3;
    o[0] = n.x; o[1] = n.y; o[2] = n.z;
}

struct SampleArgs {
    const uint8_t* color;      /* [n_frames][rows][cols][3] BGR */
    const float* depth;        /* [n_frames][rows][cols] metres */
    const float* normals;      /* [n_frames][rows][cols][3] or null: computed on the fly */
    const int32_t* frame_of;   /* [n] or null (patch i from frame i) */
    const int32_t* centers;    /* [n][2] x, y */
    const float* z;            /* [n] */
    int n_frames, rows, cols;
    float fx, fy, m;
    float* out;                /* [n][64][64][7] */
    int32_t* status;           /* [n] 0 ok, 1 cut rectangle outside the frame (the reference asserts) */
};

/* one CTA per patch, 256 threads, 16 output pixels each */
__global__ void __launch_bounds__(256) sample_patches_kernel(const SampleArgs a) {
    __shared__ int s_xo[64], s_yo[64], s_nx[64], s_ny[64];
    __shared__ float s_xa[64], s_ya[64];
    __shared__ int s_rect[5];
    const long long pi = blockIdx.x;
    const int tid = threadIdx.x;
    float* out = a.out + pi * (64 * 64 * 7);
    if (tid == 0) {
        const float z = a.z[pi];
        const int cx0 = a.centers[2 * pi], cy0 = a.centers[2 * pi + 1];
        const int w = (int)__fdiv_rn(mul(a.fx, a.m), z), h = (int)__fdiv_rn(mul(a.fy, a.m), z);
        int x = cx0 - w / 2, y = cy0 - h / 2;
        if (x < 0) x = 0;
        if (y < 0) y = 0;
        if (x > a.cols - w - 1) x = a.cols - w - 1;
        if (y > a.rows - h - 1) y = a.rows - h - 1;
        const bool ok = isfinite(z) && w > 0 && h > 0 && x >= 0 && x < a.cols - w && y >= 0 && y < a.rows - h;
        s_rect[0] = x; s_rect[1] = y; s_rect[2] = w; s_rect[3] = h; s_rect[4] = ok ? 1 : 0;
        if (a.status) a.status[pi] = ok ? 0 : 1;
    }
    __syncthreads();
    const int x = s_rect[0], y = s_rect[1], w = s_rect[2], h = s_rect[3];
    if (!s_rect[4]) {
        for (int i = tid; i < 64 * 64 * 7; i += 256) out[i] = 0.f;
        return;
    }
    if (tid < 128) {
        /* cv::resize coordinate tables: scale = 1. / (64. / size) in double, as resize() computes it */
        const int d = tid & 63;
        const bool is_y = tid >= 64;
        const int ssize = is_y ? h : w;
        const double scale = 1. / (64. / (double)ssize);
        float f = (float)(((double)d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f = sub(f, (float)s);
        int nn = (int)floor((double)d * scale);
        if (nn > ssize - 1) nn = ssize - 1;
        if (is_y) {            /* vertical weights are not clamped; the row indices are clipped at use */
            s_yo[d] = s; s_ya[d] = f; s_ny[d] = nn;
        } else {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
            s_xo[d] = s; s_xa[d] = f; s_nx[d] = nn;
        }
    }
    __syncthreads();
    const long long frame = a.frame_of ? a.frame_of[pi] : pi;
    const size_t per = (size_t)a.rows * a.cols;
    const uint8_t* col = a.color + frame * per * 3;
    const float* dep = a.depth + frame * per;
    const float* nor = a.normals ? a.normals + frame * per * 3 : nullptr;
    const float inv255 = __fdiv_rn(1.f, 255.f), inv_m = __fdiv_rn(1.0f, a.m);
    const float z = a.z[pi];
    const bool area = w == 128 && h == 128;     /* both scale factors exactly 2: cv::resize reroutes INTER_LINEAR to INTER_AREA */
    for (int o = tid; o < 64 * 64; o += 256) {
        const int dy = o >> 6, dx = o & 63;
        float v[7];
        int tx[4], ty[4];
        float wgt_a0, wgt_a1, wgt_b0, wgt_b1;
        bool single_x = false;
        if (area) {
            tx[0] = tx[2] = 2 * dx; tx[1] = tx[3] = 2 * dx + 1;
            ty[0] = ty[1] = 2 * dy; ty[2] = ty[3] = 2 * dy + 1;
            wgt_a0 = wgt_a1 = wgt_b0 = wgt_b1 = 0.f;
        } else {
            const int x0 = s_xo[dx];
            single_x = x0 >= w - 1;
            const int y0 = min(max(s_yo[dy], 0), h - 1), y1 = min(max(s_yo[dy] + 1, 0), h - 1);
            tx[0] = tx[2] = x0; tx[1] = tx[3] = single_x ? x0 : x0 + 1;
            ty[0] = ty[1] = y0; ty[2] = ty[3] = y1;
            wgt_a1 = s_xa[dx]; wgt_a0 = sub(1.f, wgt_a1);
            wgt_b1 = s_ya[dy]; wgt_b0 = sub(1.f, wgt_b1);
        }
        float t[4][6];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int gx = x + tx[k], gy = y + ty[k];
            const uint8_t* c3 = col + ((size_t)gy * a.cols + gx) * 3;
            t[k][0] = mul((float)c3[0], inv255);
            t[k][1] = mul((float)c3[1], inv255);
            t[k][2] = mul((float)c3[2], inv255);
            float3 n;
            if (nor) {
                const float* n3 = nor + ((size_t)gy * a.cols + gx) * 3;
                n = make_float3(__ldg(n3), __ldg(n3 + 1), __ldg(n3 + 2));
            } else {
                n = normal_at(dep, a.rows, a.cols, gy, gx, a.fx, a.fy);
            }
            t[k][3] = n.x; t[k][4] = n.y; t[k][5] = n.z;
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            float r;
            if (area) {
                r = mul(add(add(add(t[0][c], t[1][c]), t[2][c]), t[3][c]), 0.25f);
            } else {
                const float r0 = single_x ? t[0][c] : add(mul(t[0][c], wgt_a0), mul(t[1][c], wgt_a1));
                const float r1 = single_x ? t[2][c] : add(mul(t[2][c], wgt_a0), mul(t[3][c], wgt_a1));
                r = add(mul(r0, wgt_b0), mul(r1, wgt_b1));
            }
            v[c < 3 ? c : c + 1] = r;
        }
        float d = __ldg(dep + (size_t)(y + s_ny[dy]) * a.cols + x + s_nx[dx]);
        d = sub(d, z);
        if (d < -a.m) d = -a.m;
        if (d > a.m) d = a.m;
        d = mul(d, inv_m);
        v[3] = mul(add(d, 1.f), 0.5f);
        float* op = out + (size_t)o * 7;
#pragma unroll
        for (int c = 0; c < 7; ++c) op[c] = v[c];
    }
}

struct Ctx {
    int device;
    cudaStream_t stream;
};
int ctx_of(pe_net* net, int device, Ctx* c) {
    if (net) {
        PE_CHECK(net_ensure_device(net));
        c->device = net->device;
        c->stream = net->stream;
    } else {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
            cudaGetLastError();
            set_error("no CUDA device %d (the library has no CPU fallback)", device);
            return PE_ERR_CUDA;
        }
        c->device = device;
        c->stream = nullptr;
    }
    PE_CUDA(cudaSetDevice(c->device));
    return PE_OK;
}

}  // namespace

int depth2normals_launch(cudaStream_t stream, const float* d_depth, int n_frames, int rows, int cols, float fx, float fy, float* d_normals) {
    const long long total = (long long)n_frames * rows * cols;
    if (total == 0) return PE_OK;
    depth2normals_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(d_depth, n_frames, rows, cols, fx, fy, d_normals);
    PE_LAUNCHED();
    return PE_OK;
}

}  // namespace pe

using namespace pe;

extern "C" {

int pe_depth2normals_dev(pe_net* net, int device, const float* d_depth, int n_frames, int rows, int cols, float fx, float fy,
                         float ox, float oy, float* d_normals) {
    (void)ox; (void)oy;     /* unused by the reference's active branch (src/helper.cpp:91-93) */
    PE_REQUIRE(n_frames >= 0 && rows > 0 && cols > 0, "bad frame geometry");
    PE_REQUIRE(n_frames == 0 || (d_depth && d_normals), "null pointer");
    Ctx c;
    PE_CHECK(ctx_of(net, device, &c));
    return depth2normals_launch(c.stream, d_depth, n_frames, rows, cols, fx, fy, d_normals);
}

int pe_depth2normals(int device, const float* depth, int n_frames, int rows, int cols, float fx, float fy, float ox, float oy,
                     float* normals) {
    PE_REQUIRE(n_frames >= 0 && rows > 0 && cols > 0, "bad frame geometry");
    if (n_frames == 0) return PE_OK;
    PE_REQUIRE(depth && normals, "null pointer");
    Ctx c;
    PE_CHECK(ctx_of(nullptr, device, &c));
    const size_t px = (size_t)n_frames * rows * cols;
    float *d_d = nullptr, *d_n = nullptr;
    PE_CUDA(cudaMalloc(&d_d, px * 4));
    if (cudaMalloc(&d_n, px * 12) != cudaSuccess) { cudaFree(d_d); set_error("out of device memory"); return PE_ERR_CUDA; }
    int st = PE_OK;
    if (cudaMemcpy(d_d, depth, px * 4, cudaMemcpyHostToDevice) != cudaSuccess) st = PE_ERR_CUDA;
    if (st == PE_OK) st = pe_depth2normals_dev(nullptr, device, d_d, n_frames, rows, cols, fx, fy, ox, oy, d_n);
    if (st == PE_OK && cudaMemcpy(normals, d_n, px * 12, cudaMemcpyDeviceToHost) != cudaSuccess) st = PE_ERR_CUDA;
    if (st == PE_ERR_CUDA) set_error("depth2normals: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(d_d);
    cudaFree(d_n);
    return st;
}

int pe_sample_patches_dev(pe_net* net, int device, const uint8_t* d_color, const float* d_depth, const float* d_normals,
                          int n_frames, int rows, int cols, const int32_t* d_frame_of, const int32_t* d_centers_xy,
                          const float* d_z, int64_t n, float fx, float fy, float m, float* d_out_patches, int32_t* d_status) {
    PE_REQUIRE(n >= 0 && n_frames >= 0 && rows > 0 && cols > 0 && m > 0.f, "bad arguments");
    if (n == 0) return PE_OK;
    PE_REQUIRE(d_color && d_depth && d_centers_xy && d_z && d_out_patches, "null pointer");
    PE_REQUIRE(d_frame_of || n <= n_frames, "%lld patches but %d frames and no frame index", (long long)n, n_frames);
    Ctx c;
    PE_CHECK(ctx_of(net, device, &c));
    SampleArgs a;
    a.color = d_color; a.depth = d_depth; a.normals = d_normals; a.frame_of = d_frame_of; a.centers = d_centers_xy; a.z = d_z;
    a.n_frames = n_frames; a.rows = rows; a.cols = cols; a.fx = fx; a.fy = fy; a.m = m; a.out = d_out_patches; a.status = d_status;
    sample_patches_kernel<<<(unsigned)n, 256, 0, c.stream>>>(a);
    PE_LAUNCHED();
    return PE_OK;
}

int pe_sample_patches(int device, const uint8_t* color, const float* depth, const float* normals, int n_frames, int rows,
                      int cols, const int32_t* frame_of, const int32_t* centers_xy, const float* z, int64_t n, float fx,
                      float fy, float m, float* out_patches, int32_t* status) {
    PE_REQUIRE(n >= 0 && n_frames >= 0 && rows > 0 && cols > 0 && m > 0.f, "bad arguments");
    if (n == 0) return PE_OK;
    PE_REQUIRE(color && depth && centers_xy && z && out_patches, "null pointer");
    if (frame_of)
        for (int64_t i = 0; i < n; ++i) PE_REQUIRE(frame_of[i] >= 0 && frame_of[i] < n_frames, "frame_of[%lld] out of range", (long long)i);
    Ctx c;
    PE_CHECK(ctx_of(nullptr, device, &c));
    const size_t px = (size_t)n_frames * rows * cols;
    DevBuf b_col, b_dep, b_nor, b_fr, b_ctr, b_z, b_out, b_st;
    int st = PE_OK;
#define UP(buf, src, bytes) do { if (st == PE_OK) st = (buf).reserve(bytes); \
        if (st == PE_OK && cudaMemcpy((buf).p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) { set_error("H2D copy failed"); st = PE_ERR_CUDA; } } while (0)
    UP(b_col, color, px * 3);
    UP(b_dep, depth, px * 4);
    if (normals) UP(b_nor, normals, px * 12);
    if (frame_of) UP(b_fr, frame_of, (size_t)n * 4);
    UP(b_ctr, centers_xy, (size_t)n * 8);
    UP(b_z, z, (size_t)n * 4);
#undef UP
    if (st == PE_OK) st = b_out.reserve((size_t)n * 64 * 64 * 7 * 4);
    if (st == PE_OK) st = b_st.reserve((size_t)n * 4);
    if (st == PE_OK)
        st = pe_sample_patches_dev(nullptr, device, b_col.as<uint8_t>(), b_dep.as<float>(), normals ? b_nor.as<float>() : nullptr, n_frames,
                                   rows, cols, frame_of ? b_fr.as<int32_t>() : nullptr, b_ctr.as<int32_t>(), b_z.as<float>(), n, fx, fy, m,
                                   b_out.as<float>(), b_st.as<int32_t>());
    if (st == PE_OK && cudaMemcpy(out_patches, b_out.p, (size_t)n * 64 * 64 * 7 * 4, cudaMemcpyDeviceToHost) != cudaSuccess) st = PE_ERR_CUDA;
    if (st == PE_OK && status && cudaMemcpy(status, b_st.p, (size_t)n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) st = PE_ERR_CUDA;
    for (DevBuf* b : {&b_col, &b_dep, &b_nor, &b_fr, &b_ctr, &b_z, &b_out, &b_st}) b->release();
    return st;
}

}  // extern "C"
