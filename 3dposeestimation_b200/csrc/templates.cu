/*
 * templates.cu - the producer of a template set: datasetManager::createSynthSamplesWadim / createSynthSamplesPaul
 * (reference src/datasetmanager.cpp:440-496 and 498-535) for a list of poses of one mesh:
 *     render the view (renderer.renderView / SphereRenderer::createViews)        -> pe_render_views[_tex]_dev
 *     depth2normals + samplePatchWithScale at the principal point, z = pose t_z   -> pe_sample_patches_dev (normals on the fly)
 *     label {index, Quaternionf(pose.linear()).coeffs(), pose.inverse().translation()}  (:484-491, :523-530)
 * The frames of a batch of views live in two temporary device buffers and never reach the host, so
 * mesh -> patches (-> descriptors, with pe_compute_descriptors_dev on the same stream) stays on the GPU.
 * The composition adds no kernel of its own; the label arithmetic is host code (8 floats per view).
 */
#include <cmath>
#include <vector>

#include "common.h"

namespace pe {
namespace {

/* Eigen's Quaternion(rotation matrix) (Quaternion.h quaternionbase_assign_impl<Other,3,3>: Shoemake's branch on the
 * trace) and Isometry3f::inverse().translation() = -(R^T t); same statements as oracle/oracle_render.c orc_pose_label */
void pose_label(const float* rt, int index, float* out8) {
    const float m[3][3] = {{rt[0], rt[1], rt[2]}, {rt[4], rt[5], rt[6]}, {rt[8], rt[9], rt[10]}};
    float q[4];
    float t = m[0][0] + m[1][1] + m[2][2];
    if (t > 0.f) {
        t = std::sqrt(t + 1.0f);
        q[3] = 0.5f * t;
        t = 0.5f / t;
        q[0] = (m[2][1] - m[1][2]) * t;
        q[1] = (m[0][2] - m[2][0]) * t;
        q[2] = (m[1][0] - m[0][1]) * t;
    } else {
        int i = 0;
        if (m[1][1] > m[0][0]) i = 1;
        if (m[2][2] > m[i][i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0f);
        q[i] = 0.5f * t;
        t = 0.5f / t;
        q[3] = (m[k][j] - m[j][k]) * t;
        q[j] = (m[j][i] + m[i][j]) * t;
        q[k] = (m[k][i] + m[i][k]) * t;
    }
    out8[0] = static_cast<float>(index);
    for (int c = 0; c < 4; ++c) out8[1 + c] = q[c];
    for (int r = 0; r < 3; ++r) {
        volatile float s = m[0][r] * rt[3];          /* products and sums rounded separately, as the oracle (no FMA contraction) */
        volatile float p1 = m[1][r] * rt[7];
        s = s + p1;
        volatile float p2 = m[2][r] * rt[11];
        s = s + p2;
        out8[5 + r] = -s;
    }
}

/* extractRealSamplesPaul / Wadim (reference src/datasetmanager.cpp:348-411): projected centroid truncated to int and the
 * centroid's depth; same statements as oracle/oracle_render.c orc_project_centroid (products and sums rounded separately) */
float project_centroid(const float* cam, const float* rt, const float* centroid3, int32_t* out_xy) {
    volatile float c[3];
    if (centroid3) {
        for (int r = 0; r < 3; ++r) {
            volatile float a = rt[4 * r] * centroid3[0], b = rt[4 * r + 1] * centroid3[1], d = rt[4 * r + 2] * centroid3[2];
            volatile float s = a + b;
            s = s + d;
            c[r] = s + rt[4 * r + 3];
        }
    } else {
        c[0] = rt[3]; c[1] = rt[7]; c[2] = rt[11];
    }
    volatile float x0 = cam[0] * c[0], x1 = 0.f * c[1], x2 = cam[2] * c[2];
    volatile float y0 = 0.f * c[0], y1 = cam[1] * c[1], y2 = cam[3] * c[2];
    volatile float z0 = 0.f * c[0], z1 = 0.f * c[1], z2 = 1.f * c[2];
    volatile float px = x0 + x1, py = y0 + y1, pz = z0 + z1;
    px = px + x2; py = py + y2; pz = pz + z2;
    out_xy[0] = (int32_t)(px / pz);
    out_xy[1] = (int32_t)(py / pz);
    return c[2];
}

}  // namespace
}  // namespace pe

using namespace pe;

extern "C" {

int pe_pose_labels(const float* poses, int64_t n, int index, float* out_labels) {
    PE_REQUIRE(n >= 0 && (n == 0 || (poses && out_labels)), "null pointer");
    for (int64_t i = 0; i < n; ++i) pose_label(poses + i * 12, index, out_labels + i * 8);
    return PE_OK;
}

int pe_scene_samples(int device, const uint8_t* color, const float* depth, int n_frames, int rows, int cols, const float* cam4,
                     const float* poses, const float* centroid3, int z_from_depth, int index, float m, float* out_patches,
                     float* out_labels, int32_t* status) {
    PE_REQUIRE(n_frames >= 0 && rows > 0 && cols > 0 && cam4, "bad arguments");
    if (n_frames == 0) return PE_OK;
    PE_REQUIRE(color && depth && poses && out_patches, "null pointer");
    std::vector<int32_t> ctr((size_t)n_frames * 2);
    std::vector<float> z((size_t)n_frames);
    for (int i = 0; i < n_frames; ++i) {
        z[i] = project_centroid(cam4, poses + (size_t)i * 12, centroid3, &ctr[2 * i]);
        if (z_from_depth) {      /* the Wadim variant: the surface point under the projected centroid (:388-389, asserted > 0 there) */
            const int x = ctr[2 * i], y = ctr[2 * i + 1];
            PE_REQUIRE(x >= 0 && y >= 0 && x < cols && y < rows, "frame %d: projected centroid (%d, %d) outside the frame", i, x, y);
            z[i] = depth[((size_t)i * rows + y) * cols + x];
            PE_REQUIRE(z[i] > 0.f, "frame %d: no depth under the projected centroid (the reference asserts z > 0)", i);
        }
    }
    if (out_labels) PE_CHECK(pe_pose_labels(poses, n_frames, index, out_labels));
    return pe_sample_patches(device, color, depth, nullptr, n_frames, rows, cols, nullptr, ctr.data(), z.data(), n_frames, cam4[0], cam4[1], m,
                             out_patches, status);
}

int pe_template_samples_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, const float* d_uv, int n_vertices,
                            const int32_t* d_faces, int n_faces, const uint8_t* d_tex, int tex_rows, int tex_cols, const float* cam4,
                            const float* poses_host, int64_t n_views, int index, const float* z_override, int rows, int cols,
                            float z_near, float z_far, float m, float* d_out_patches, float* out_labels_host, int32_t* d_status) {
    PE_REQUIRE(n_views >= 0 && rows > 0 && cols > 0 && cam4, "bad arguments");
    if (n_views == 0) return PE_OK;
    PE_REQUIRE(poses_host && d_out_patches, "null pointer");
    cudaStream_t stream = nullptr;
    if (net) {
        PE_CHECK(net_ensure_device(net));
        device = net->device;
        stream = net->stream;
    }
    PE_CUDA(cudaSetDevice(device));
    if (out_labels_host) PE_CHECK(pe_pose_labels(poses_host, n_views, index, out_labels_host));
    /* frames of a batch of views: 7 bytes per pixel, batch sized to ~256 MB */
    const size_t frame = (size_t)rows * cols;
    int64_t batch = (int64_t)(((size_t)256 << 20) / (frame * 7));
    if (batch < 1) batch = 1;
    if (batch > n_views) batch = n_views;
    uint8_t* d_col = nullptr;
    float *d_dep = nullptr, *d_pose = nullptr, *d_z = nullptr;
    int32_t* d_ctr = nullptr;
    auto release = [&]() {
        cudaStreamSynchronize(stream);
        for (void* p : {(void*)d_col, (void*)d_dep, (void*)d_pose, (void*)d_z, (void*)d_ctr})
            if (p) cudaFree(p);
    };
    if (cudaMalloc(reinterpret_cast<void**>(&d_col), frame * 3 * batch) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void**>(&d_dep), frame * 4 * batch) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void**>(&d_pose), (size_t)n_views * 48) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void**>(&d_z), (size_t)n_views * 4) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void**>(&d_ctr), (size_t)n_views * 8) != cudaSuccess) {
        cudaGetLastError();
        release();
        set_error("out of device memory for %lld frames of %d x %d", (long long)batch, cols, rows);
        return PE_ERR_CUDA;
    }
    /* patch centre = the principal point truncated to int (cam(0,2), cam(1,2) are passed to int parameters), z = the
     * pose's z translation (:470 trans(2), :516 v.pose.translation()(2)) unless the caller overrides it */
    std::vector<float> z((size_t)n_views);
    std::vector<int32_t> ctr((size_t)n_views * 2);
    for (int64_t i = 0; i < n_views; ++i) {
        z[i] = z_override ? z_override[i] : poses_host[i * 12 + 11];
        ctr[2 * i] = (int32_t)cam4[2];
        ctr[2 * i + 1] = (int32_t)cam4[3];
    }
    int st = PE_OK;
    if (cudaMemcpyAsync(d_pose, poses_host, (size_t)n_views * 48, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
        cudaMemcpyAsync(d_z, z.data(), (size_t)n_views * 4, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
        cudaMemcpyAsync(d_ctr, ctr.data(), (size_t)n_views * 8, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess) {       /* z / ctr are host temporaries */
        set_error("pose upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        st = PE_ERR_CUDA;
    }
    for (int64_t first = 0; first < n_views && st == PE_OK; first += batch) {
        const int64_t nb = n_views - first < batch ? n_views - first : batch;
        st = pe_render_views_tex_dev(net, device, d_vertices, d_colors, d_uv, n_vertices, d_faces, n_faces, d_tex, tex_rows, tex_cols, cam4,
                                     d_pose + first * 12, (int)nb, rows, cols, z_near, z_far, d_col, d_dep);
        if (st == PE_OK)
            st = pe_sample_patches_dev(net, device, d_col, d_dep, nullptr, (int)nb, rows, cols, nullptr, d_ctr + first * 2, d_z + first, nb,
                                       cam4[0], cam4[1], m, d_out_patches + (size_t)first * 64 * 64 * 7, d_status ? d_status + first : nullptr);
    }
    release();
    return st;
}

int pe_template_samples(int device, const float* vertices, const float* colors, const float* uv, int n_vertices, const int32_t* faces,
                        int n_faces, const uint8_t* tex, int tex_rows, int tex_cols, const float* cam4, const float* poses, int64_t n_views,
                        int index, const float* z_override, int rows, int cols, float z_near, float z_far, float m, float* out_patches,
                        float* out_labels, int32_t* status) {
    PE_REQUIRE(n_vertices >= 0 && n_faces >= 0 && n_views >= 0, "bad sizes");
    PE_REQUIRE((uv == nullptr) == (tex == nullptr), "texture coordinates and texture image go together");
    if (n_views == 0) return PE_OK;
    PE_REQUIRE(out_patches && poses && cam4, "null pointer");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) {
        cudaGetLastError();
        set_error("no CUDA device %d (the library has no CPU fallback)", device);
        return PE_ERR_CUDA;
    }
    PE_CUDA(cudaSetDevice(device));
    DevBuf b_v, b_c, b_f, b_uv, b_tex, b_out, b_st;
    int st = PE_OK;
#define UP(buf, src, bytes) do { const size_t nb_ = (bytes); if (st == PE_OK) st = (buf).reserve(nb_ > 0 ? nb_ : 4); \
        if (st == PE_OK && nb_ > 0 && cudaMemcpy((buf).p, src, nb_, cudaMemcpyHostToDevice) != cudaSuccess) { set_error("H2D copy failed"); st = PE_ERR_CUDA; } } while (0)
    UP(b_v, vertices, (size_t)n_vertices * 12);
    UP(b_c, colors, (size_t)n_vertices * 12);
    UP(b_f, faces, (size_t)n_faces * 12);
    if (uv) {
        UP(b_uv, uv, (size_t)n_vertices * 8);
        UP(b_tex, tex, (size_t)tex_rows * tex_cols * 3);
    }
#undef UP
    const size_t patch_bytes = (size_t)n_views * 64 * 64 * 7 * 4;
    if (st == PE_OK) st = b_out.reserve(patch_bytes);
    if (st == PE_OK) st = b_st.reserve((size_t)n_views * 4);
    if (st == PE_OK)
        st = pe_template_samples_dev(nullptr, device, b_v.as<float>(), b_c.as<float>(), uv ? b_uv.as<float>() : nullptr, n_vertices,
                                     b_f.as<int32_t>(), n_faces, uv ? b_tex.as<uint8_t>() : nullptr, tex_rows, tex_cols, cam4, poses, n_views,
                                     index, z_override, rows, cols, z_near, z_far, m, b_out.as<float>(), out_labels, b_st.as<int32_t>());
    if (st == PE_OK && (cudaMemcpy(out_patches, b_out.p, patch_bytes, cudaMemcpyDeviceToHost) != cudaSuccess ||
                        (status && cudaMemcpy(status, b_st.p, (size_t)n_views * 4, cudaMemcpyDeviceToHost) != cudaSuccess))) {
        set_error("template read-back failed: %s", cudaGetErrorString(cudaGetLastError()));
        st = PE_ERR_CUDA;
    }
    for (DevBuf* b : {&b_v, &b_c, &b_f, &b_uv, &b_tex, &b_out, &b_st}) b->release();
    return st;
}

}  // extern "C"
