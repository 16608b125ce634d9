/*
 * render.cu - template renderer as a CUDA rasteriser: what the reference does through Qt / OpenGL
 * (`SphereRenderer::createViews` -> `Painter::paint` with `RealWorldCamera` + `Model`; reference src/sphere.cpp:27-94,
 * 134-226, src/painter.cpp:86-110,186-194,231-248, src/model.cpp:29-54), without a graphics API:
 *
 *   pe_sphere_views      host: icosahedron subdivision in the reference's vertex order, look-at rotation, in-plane
 *                        rotation and distance loops -> one row-major 3x4 [R|t] per view
 *   pe_render_views      device: for a batch of views of one mesh, colour (BGR u8) and metric depth (f32, 0 = background)
 *                        frames in the layout pe_sample_patches reads, so views -> patches -> descriptors stays on the GPU
 *
 * Pipeline (same maths as oracle/oracle_render.c, which it matches bit for bit): camera point Xc = R X + t, window
 * x = fx Xc/Zc + ox, y = fy Yc/Zc + oy, vertices snapped to 1/256 pixel, exact 64-bit integer edge functions at pixel
 * centres with a top-left style tie rule (watertight), screen-space barycentrics, 1/Z interpolated linearly (GL's window
 * depth), perspective-correct per-vertex colour, no lighting, no culling, depth test LESS with ties to the lower
 * triangle index (first drawn wins).
 *
 * Kernels: `render_xform_kernel` (view x vertex), `render_raster_kernel` (view x triangle: near-plane clip into at most
 * two sub-triangles, bounding-box walk, atomicMin of (depth bits << 32 | 2 * triangle + sub) into a 64-bit z-buffer),
 * `render_resolve_kernel` (view x pixel: winner sub-triangle -> colour [x bilinear GL_REPEAT texel, GL_MODULATE, for
 * textured meshes: src/model.cpp:330-350, src/painter.cpp:140-160] + depth).  The 24-bit quantisation of GL's depth buffer
 * is not reproduced.  Meshes come from pe_mesh_load_ply (io.cpp; the reference loads them with cv::viz::Mesh::load,
 * src/model.cpp:304).
 */
#include <cmath>
#include <cstdint>
#include <vector>

#include "common.h"

namespace pe {

namespace {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }

struct Cam { float fx, fy, ox, oy; };

/* per (view, vertex): camera-space point, screen position snapped to 1/256 px (valid when in front of the near plane) */
struct SVert { int x, y; float xc, yc, zc; int ok; };

__device__ __forceinline__ void project_snap(const Cam& cam, float xc, float yc, float zc, int* px, int* py) {
    const float u = __fdiv_rn(add(mul(cam.fx, xc), mul(cam.ox, zc)), zc), w = __fdiv_rn(add(mul(cam.fy, yc), mul(cam.oy, zc)), zc);
    const float lim = 1048576.f * 256.f;
    float us = mul(u, 256.f), vs = mul(w, 256.f);
    us = fminf(fmaxf(us, -lim), lim);
    vs = fminf(fmaxf(vs, -lim), lim);
    *px = (int)floorf(add(us, 0.5f));
    *py = (int)floorf(add(vs, 0.5f));
}

__global__ void render_xform_kernel(const float* __restrict__ vertices, int nv, const float* __restrict__ poses, int n_views,
                                    Cam cam, float z_near, SVert* __restrict__ out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)nv * n_views) return;
    const int view = (int)(i / nv), v = (int)(i - (long long)view * nv);
    const float* rt = poses + (size_t)view * 12;
    const float x = vertices[3 * v], y = vertices[3 * v + 1], z = vertices[3 * v + 2];
    SVert s;
    s.xc = add(add(add(mul(rt[0], x), mul(rt[1], y)), mul(rt[2], z)), rt[3]);
    s.yc = add(add(add(mul(rt[4], x), mul(rt[5], y)), mul(rt[6], z)), rt[7]);
    s.zc = add(add(add(mul(rt[8], x), mul(rt[9], y)), mul(rt[10], z)), rt[11]);
    s.ok = s.zc > z_near ? 1 : 0;
    s.x = s.y = 0;
    if (s.ok) project_snap(cam, s.xc, s.yc, s.zc, &s.x, &s.y);
    out[i] = s;
}

__device__ __forceinline__ long long edge(long long ax, long long ay, long long bx, long long by, long long px, long long py) {
    return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}
__device__ __forceinline__ bool owns_edge(long long ax, long long ay, long long bx, long long by) {
    const long long dx = bx - ax, dy = by - ay;
    return dy < 0 || (dy == 0 && dx > 0);
}

/* a vertex of the clipped polygon: original vertex (i0 == i1) or the cut point i0 + t (i1 - i0), i0 inside the near plane */
struct CV { int x, y; float z; int i0, i1; float t; };

/* Sutherland-Hodgman against Zc = z_near in camera space (oracle_render.c orc_render_tex): 0, 3 or 4 polygon vertices.
 * A vertex exactly on the plane counts as outside; a cut point is always computed from the inside vertex, so the two
 * triangles at a shared edge get the same point (watertight). */
__device__ __forceinline__ int clip_face(const int32_t* __restrict__ faces, int f, const SVert* __restrict__ sv, int nv, const Cam& cam,
                                         float z_near, CV* poly) {
    const int idx[3] = {faces[3 * f], faces[3 * f + 1], faces[3 * f + 2]};
    if (idx[0] < 0 || idx[1] < 0 || idx[2] < 0 || idx[0] >= nv || idx[1] >= nv || idx[2] >= nv) return 0;
    const SVert s[3] = {sv[idx[0]], sv[idx[1]], sv[idx[2]]};
    if (!s[0].ok && !s[1].ok && !s[2].ok) return 0;
    int n = 0;
#pragma unroll
    for (int e = 0; e < 3; ++e) {
        const int p = e, q = (e + 1) % 3;
        if (s[p].ok) {
            CV v = {s[p].x, s[p].y, s[p].zc, idx[p], idx[p], 0.f};
            poly[n++] = v;
        }
        if (s[p].ok != s[q].ok) {
            const int ia = s[p].ok ? p : q, ib = s[p].ok ? q : p;      /* a inside, b outside */
            const float t = __fdiv_rn(__fsub_rn(s[ia].zc, z_near), __fsub_rn(s[ia].zc, s[ib].zc));
            const float xc = add(s[ia].xc, mul(t, __fsub_rn(s[ib].xc, s[ia].xc))), yc = add(s[ia].yc, mul(t, __fsub_rn(s[ib].yc, s[ia].yc)));
            CV v = {0, 0, z_near, idx[ia], idx[ib], t};
            project_snap(cam, xc, yc, z_near, &v.x, &v.y);
            poly[n++] = v;
        }
    }
    return n;
}

struct Tri {
    long long ax, ay, bx, by, cx, cy, area;
    CV va, vb, vc;
    float iza, izb, izc;
    bool oa, ob, oc, ok;
};

__device__ __forceinline__ Tri tri_setup(const CV& a, const CV& b, const CV& c) {
    Tri t;
    t.ok = false;
    t.va = a; t.vb = b; t.vc = c;
    t.ax = a.x; t.ay = a.y; t.bx = b.x; t.by = b.y; t.cx = c.x; t.cy = c.y;
    t.area = edge(t.ax, t.ay, t.bx, t.by, t.cx, t.cy);
    if (t.area == 0) return t;
    if (t.area < 0) {      /* no culling: flip to positive orientation */
        const CV tv = t.vb; t.vb = t.vc; t.vc = tv;
        const long long tx = t.bx, ty = t.by; t.bx = t.cx; t.by = t.cy; t.cx = tx; t.cy = ty;
        t.area = -t.area;
    }
    t.iza = __fdiv_rn(1.f, t.va.z); t.izb = __fdiv_rn(1.f, t.vb.z); t.izc = __fdiv_rn(1.f, t.vc.z);
    t.oa = owns_edge(t.bx, t.by, t.cx, t.cy);
    t.ob = owns_edge(t.cx, t.cy, t.ax, t.ay);
    t.oc = owns_edge(t.ax, t.ay, t.bx, t.by);
    t.ok = true;
    return t;
}

/* coverage + barycentrics of pixel (px, py); false when the pixel is not this triangle's */
__device__ __forceinline__ bool tri_sample(const Tri& t, int px, int py, float* ba, float* bb, float* bc) {
    const long long qx = 256LL * px + 128, qy = 256LL * py + 128;
    const long long wa = edge(t.bx, t.by, t.cx, t.cy, qx, qy), wb = edge(t.cx, t.cy, t.ax, t.ay, qx, qy), wc = edge(t.ax, t.ay, t.bx, t.by, qx, qy);
    if (wa < 0 || wb < 0 || wc < 0) return false;
    if ((wa == 0 && !t.oa) || (wb == 0 && !t.ob) || (wc == 0 && !t.oc)) return false;
    const float fa = (float)t.area;
    *ba = __fdiv_rn((float)wa, fa); *bb = __fdiv_rn((float)wb, fa); *bc = __fdiv_rn((float)wc, fa);
    return true;
}

__global__ void render_raster_kernel(const int32_t* __restrict__ faces, int nf, const SVert* __restrict__ sverts, int nv, int n_views,
                                     Cam cam, int rows, int cols, float z_near, float z_far, unsigned long long* __restrict__ zbuf) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)nf * n_views) return;
    const int view = (int)(i / nf), f = (int)(i - (long long)view * nf);
    CV poly[4];
    const int np = clip_face(faces, f, sverts + (size_t)view * nv, nv, cam, z_near, poly);
    unsigned long long* zb = zbuf + (size_t)view * rows * cols;
    for (int sub = 0; sub + 2 < np; ++sub) {
        const Tri t = tri_setup(poly[0], poly[sub + 1], poly[sub + 2]);
        if (!t.ok) continue;
        long long minx = min(t.ax, min(t.bx, t.cx)), maxx = max(t.ax, max(t.bx, t.cx)), miny = min(t.ay, min(t.by, t.cy)), maxy = max(t.ay, max(t.by, t.cy));
        int x0 = (int)((minx - 128 + 255) >> 8), x1 = (int)((maxx - 128) >> 8), y0 = (int)((miny - 128 + 255) >> 8), y1 = (int)((maxy - 128) >> 8);
        x0 = max(x0, 0); y0 = max(y0, 0); x1 = min(x1, cols - 1); y1 = min(y1, rows - 1);
        for (int py = y0; py <= y1; ++py)
            for (int px = x0; px <= x1; ++px) {
                float ba, bb, bc;
                if (!tri_sample(t, px, py, &ba, &bb, &bc)) continue;
                const float iz = add(add(mul(ba, t.iza), mul(bb, t.izb)), mul(bc, t.izc));
                const float z = __fdiv_rn(1.f, iz);
                if (!(z > z_near) || !(z < z_far)) continue;
                /* positive floats order like their bit patterns; ties go to the lower (triangle, sub-triangle) id: first drawn, GL_LESS */
                atomicMin(zb + (size_t)py * cols + px, ((unsigned long long)__float_as_uint(z) << 32) | (unsigned)(2 * f + sub));
            }
    }
}

__device__ __forceinline__ float attr_at(const float* __restrict__ a, int stride, int ch, const CV& v) {
    const float a0 = a[stride * v.i0 + ch];
    if (v.i0 == v.i1) return a0;
    return add(a0, mul(v.t, __fsub_rn(a[stride * v.i1 + ch], a0)));
}

/* GL_LINEAR + GL_REPEAT fetch of channel ch_rgb from a BGR image, in [0, 1] (oracle_render.c tex_bilinear) */
__device__ __forceinline__ float tex_bilinear(const uint8_t* __restrict__ tex, int th, int tw, float s, float t, int ch_rgb) {
    const float fu = __fsub_rn(mul(s, (float)tw), 0.5f), fv = __fsub_rn(mul(t, (float)th), 0.5f);
    const float x0f = floorf(fu), y0f = floorf(fv);
    const float ax = __fsub_rn(fu, x0f), ay = __fsub_rn(fv, y0f);
    int x0 = (int)fmodf(x0f, (float)tw), y0 = (int)fmodf(y0f, (float)th);
    if (x0 < 0) x0 += tw;
    if (y0 < 0) y0 += th;
    const int x1 = x0 + 1 == tw ? 0 : x0 + 1, y1 = y0 + 1 == th ? 0 : y0 + 1;
    const int c = 2 - ch_rgb;
    const float t00 = tex[((size_t)y0 * tw + x0) * 3 + c], t10 = tex[((size_t)y0 * tw + x1) * 3 + c];
    const float t01 = tex[((size_t)y1 * tw + x0) * 3 + c], t11 = tex[((size_t)y1 * tw + x1) * 3 + c];
    const float top = add(t00, mul(ax, __fsub_rn(t10, t00))), bot = add(t01, mul(ax, __fsub_rn(t11, t01)));
    return __fdiv_rn(add(top, mul(ay, __fsub_rn(bot, top))), 255.f);
}

__global__ void render_resolve_kernel(const int32_t* __restrict__ faces, const float* __restrict__ colors, const float* __restrict__ uv,
                                      const uint8_t* __restrict__ tex, int tex_rows, int tex_cols, const SVert* __restrict__ sverts,
                                      int nv, int n_views, Cam cam, float z_near, int rows, int cols,
                                      const unsigned long long* __restrict__ zbuf, uint8_t* __restrict__ out_color, float* __restrict__ out_depth) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long per = (long long)rows * cols;
    if (i >= per * n_views) return;
    const int view = (int)(i / per);
    const int p = (int)(i - view * per), py = p / cols, px = p - py * cols;
    const unsigned long long e = zbuf[i];
    uint8_t* cp = out_color + i * 3;
    if (e == ~0ULL) {
        cp[0] = cp[1] = cp[2] = 0;      /* painter.setBackground(0,0,0) */
        out_depth[i] = 0.f;             /* convertZBufferToDepth: cleared depth -> 0 */
        return;
    }
    out_depth[i] = __uint_as_float((unsigned)(e >> 32));
    const unsigned id = (unsigned)(e & 0xFFFFFFFFu);
    CV poly[4];
    clip_face(faces, (int)(id >> 1), sverts + (size_t)view * nv, nv, cam, z_near, poly);
    const int sub = (int)(id & 1u);
    const Tri t = tri_setup(poly[0], poly[sub + 1], poly[sub + 2]);
    float ba, bb, bc;
    tri_sample(t, px, py, &ba, &bb, &bc);
    const float iz = add(add(mul(ba, t.iza), mul(bb, t.izb)), mul(bc, t.izc));
    const bool textured = uv != nullptr && tex != nullptr;
    float s = 0.f, tt = 0.f;
    if (textured) {
        s = __fdiv_rn(add(add(mul(ba, mul(attr_at(uv, 2, 0, t.va), t.iza)), mul(bb, mul(attr_at(uv, 2, 0, t.vb), t.izb))),
                          mul(bc, mul(attr_at(uv, 2, 0, t.vc), t.izc))), iz);
        tt = __fdiv_rn(add(add(mul(ba, mul(attr_at(uv, 2, 1, t.va), t.iza)), mul(bb, mul(attr_at(uv, 2, 1, t.vb), t.izb))),
                           mul(bc, mul(attr_at(uv, 2, 1, t.vc), t.izc))), iz);
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const float num = add(add(mul(ba, mul(attr_at(colors, 3, ch, t.va), t.iza)), mul(bb, mul(attr_at(colors, 3, ch, t.vb), t.izb))),
                              mul(bc, mul(attr_at(colors, 3, ch, t.vc), t.izc)));
        float c = __fdiv_rn(num, iz);
        if (textured) c = mul(c, tex_bilinear(tex, tex_rows, tex_cols, s, tt, ch));      /* GL_MODULATE */
        float q = add(mul(c, 255.f), 0.5f);
        q = fminf(fmaxf(q, 0.f), 255.f);
        cp[2 - ch] = (uint8_t)q;        /* BGR read-back (glReadPixels GL_BGR) */
    }
}

/* ---- host: view sphere (restated from the reference in double-checked float arithmetic) ---- */
struct V3 { float x, y, z; };
V3 normalized(V3 a) {
    const float n = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
    return {a.x / n, a.y / n, a.z / n};
}
V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
bool same(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

void add_unique(std::vector<V3>& s, V3 v1, V3 v2, V3 v3) {
    int f1 = -1, f2 = -1, f3 = -1;
    for (size_t i = 0; i < s.size(); ++i) {
        if (same(s[i], v1)) f1 = (int)i;
        if (same(s[i], v2)) f2 = (int)i;
        if (same(s[i], v3)) f3 = (int)i;
        if (f1 != -1 && f2 != -1 && f3 != -1) break;
    }
    if (f1 == -1) s.push_back(v1);
    if (f2 == -1) s.push_back(v2);
    if (f3 == -1) s.push_back(v3);
}
void subdivide(std::vector<V3>& s, V3 v1, V3 v2, V3 v3, int depth) {
    if (depth == 0) { add_unique(s, v1, v2, v3); return; }
    const V3 v12 = normalized({v1.x + v2.x, v1.y + v2.y, v1.z + v2.z});
    const V3 v23 = normalized({v2.x + v3.x, v2.y + v3.y, v2.z + v3.z});
    const V3 v31 = normalized({v3.x + v1.x, v3.y + v1.y, v3.z + v1.z});
    subdivide(s, v1, v12, v31, depth - 1);
    subdivide(s, v2, v23, v12, depth - 1);
    subdivide(s, v3, v31, v23, depth - 1);
    subdivide(s, v12, v23, v31, depth - 1);
}
std::vector<V3> init_sphere(int depth) {
    static const int ind[20][3] = {{0, 4, 1}, {0, 9, 4}, {9, 5, 4}, {4, 5, 8}, {4, 8, 1}, {8, 10, 1}, {8, 3, 10}, {5, 3, 8}, {5, 2, 3}, {2, 7, 3},
                                   {7, 10, 3}, {7, 6, 10}, {7, 11, 6}, {11, 0, 6}, {0, 1, 6}, {6, 1, 10}, {9, 0, 11}, {9, 11, 2}, {9, 2, 5}, {7, 2, 11}};
    const float X = 0.525731112119133606f, Z = 0.850650808352039932f;
    const V3 raw[12] = {{-X, 0, +Z}, {+X, 0, +Z}, {-X, 0, -Z}, {+X, 0, -Z}, {0, +Z, +X}, {0, +Z, -X},
                        {0, -Z, +X}, {0, -Z, -X}, {+Z, +X, 0}, {-Z, +X, 0}, {+Z, -X, 0}, {-Z, -X, 0}};
    V3 v[12];
    for (int i = 0; i < 12; ++i) v[i] = normalized(raw[i]);
    std::vector<V3> s;
    for (int i = 0; i < 20; ++i) subdivide(s, v[ind[i][0]], v[ind[i][1]], v[ind[i][2]], depth);
    for (V3& p : s) p = normalized(p);
    return s;
}
void view_pose(V3 eye, float scale, float angle_deg, float* rt) {
    V3 up = {0, 0, 1};
    if (eye.x == 0 && eye.y == 0 && eye.z != 0) up = {-1, 0, 0};
    const V3 en = normalized(eye);
    const V3 c2 = {-en.x, -en.y, -en.z};
    const V3 c0 = normalized(cross(c2, normalized(up)));
    const V3 c1 = cross(c0, {-c2.x, -c2.y, -c2.z});
    const float rot[3][3] = {{c0.x, c0.y, c0.z}, {c1.x, c1.y, c1.z}, {c2.x, c2.y, c2.z}};
    const float a = (float)((double)angle_deg * 3.1415926535897932384626433832795 / 180.0);
    const float ca = std::cos(a), sa = std::sin(a);
    const float rz[3][3] = {{ca, -sa, 0}, {sa, ca, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) rt[4 * i + j] = rz[i][0] * rot[0][j] + rz[i][1] * rot[1][j] + rz[i][2] * rot[2][j];
        rt[4 * i + 3] = i == 2 ? scale : 0.f;
    }
}

int render_ctx(pe_net* net, int device, cudaStream_t* stream) {
    if (net) {
        PE_CHECK(net_ensure_device(net));
        device = net->device;
        *stream = net->stream;
    } else {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
            cudaGetLastError();
            set_error("no CUDA device %d (the library has no CPU fallback)", device);
            return PE_ERR_CUDA;
        }
        *stream = nullptr;
    }
    PE_CUDA(cudaSetDevice(device));
    return PE_OK;
}

}  // namespace
}  // namespace pe

using namespace pe;

extern "C" {

int pe_sphere_views(int sphere_depth, const float* scale3, const float* rot3, int skip_lower_hemi, float* out_poses, int64_t cap,
                    int64_t* n_views) {
    PE_REQUIRE(sphere_depth >= 0 && sphere_depth <= 6, "sphere depth %d outside 0..6", sphere_depth);
    PE_REQUIRE(scale3 && rot3 && n_views, "null pointer");
    PE_REQUIRE(scale3[1] > 0.f && rot3[1] > 0.f, "scale / rotation step must be positive");
    std::vector<float> sca, rots;
    for (float i = rot3[0]; i <= rot3[2]; i += rot3[1]) rots.push_back(i);          /* src/sphere.cpp:73-74 */
    for (float i = scale3[0]; i <= scale3[2]; i += scale3[1]) sca.push_back(i);
    const std::vector<V3> sphere = init_sphere(sphere_depth);
    int64_t n = 0;
    for (float s : sca)
        for (const V3& pos : sphere) {
            if (pos.z < 0 && skip_lower_hemi) continue;
            for (float r : rots) {
                if (out_poses && n < cap) view_pose(pos, s, r, out_poses + n * 12);
                ++n;
            }
        }
    *n_views = n;
    PE_REQUIRE(!out_poses || n <= cap, "%lld views do not fit the %lld provided", (long long)n, (long long)cap);
    return PE_OK;
}

int pe_render_views_tex_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, const float* d_uv, int n_vertices,
                            const int32_t* d_faces, int n_faces, const uint8_t* d_tex, int tex_rows, int tex_cols, const float* cam4,
                            const float* d_poses, int n_views, int rows, int cols, float z_near, float z_far, uint8_t* d_out_color,
                            float* d_out_depth) {
    PE_REQUIRE(n_vertices >= 0 && n_faces >= 0 && n_views >= 0 && rows > 0 && cols > 0, "bad sizes");
    PE_REQUIRE(n_faces < (1 << 30), "more than 2^30 faces");
    PE_REQUIRE(cam4 && z_near > 0.f && z_far > z_near, "bad camera");
    PE_REQUIRE((d_uv == nullptr) == (d_tex == nullptr), "texture coordinates and texture image go together");
    PE_REQUIRE(d_tex == nullptr || (tex_rows > 0 && tex_cols > 0), "empty texture");
    if (n_views == 0) return PE_OK;
    PE_REQUIRE(d_out_color && d_out_depth && d_poses, "null pointer");
    PE_REQUIRE((n_vertices == 0 || (d_vertices && d_colors)) && (n_faces == 0 || d_faces), "null mesh pointer");
    cudaStream_t stream;
    PE_CHECK(render_ctx(net, device, &stream));
    const Cam cam = {cam4[0], cam4[1], cam4[2], cam4[3]};
    const size_t px = (size_t)rows * cols * n_views;
    SVert* sv = nullptr;
    unsigned long long* zb = nullptr;
    PE_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&sv), sizeof(SVert) * (size_t)(n_vertices > 0 ? n_vertices : 1) * n_views, stream));
    if (cudaMallocAsync(reinterpret_cast<void**>(&zb), px * 8, stream) != cudaSuccess) {
        cudaFreeAsync(sv, stream);
        set_error("out of device memory for the z-buffer of %d views", n_views);
        return PE_ERR_CUDA;
    }
    cudaMemsetAsync(zb, 0xFF, px * 8, stream);
    const long long nvv = (long long)n_vertices * n_views, nfv = (long long)n_faces * n_views;
    if (nvv > 0) { render_xform_kernel<<<(unsigned)((nvv + 255) / 256), 256, 0, stream>>>(d_vertices, n_vertices, d_poses, n_views, cam, z_near, sv); count_launch(); }
    if (nfv > 0) { render_raster_kernel<<<(unsigned)((nfv + 127) / 128), 128, 0, stream>>>(d_faces, n_faces, sv, n_vertices, n_views, cam, rows, cols, z_near, z_far, zb); count_launch(); }
    render_resolve_kernel<<<(unsigned)((px + 255) / 256), 256, 0, stream>>>(d_faces, d_colors, d_uv, d_tex, tex_rows, tex_cols, sv, n_vertices, n_views, cam,
                                                                           z_near, rows, cols, zb, d_out_color, d_out_depth);
    PE_LAUNCHED();
    cudaFreeAsync(sv, stream);
    cudaFreeAsync(zb, stream);
    return PE_OK;
}

int pe_render_views_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, int n_vertices, const int32_t* d_faces,
                        int n_faces, const float* cam4, const float* d_poses, int n_views, int rows, int cols, float z_near, float z_far,
                        uint8_t* d_out_color, float* d_out_depth) {
    return pe_render_views_tex_dev(net, device, d_vertices, d_colors, nullptr, n_vertices, d_faces, n_faces, nullptr, 0, 0, cam4, d_poses, n_views,
                                   rows, cols, z_near, z_far, d_out_color, d_out_depth);
}

int pe_render_views_tex(int device, const float* vertices, const float* colors, const float* uv, int n_vertices, const int32_t* faces, int n_faces,
                        const uint8_t* tex, int tex_rows, int tex_cols, const float* cam4, const float* poses, int n_views, int rows, int cols,
                        float z_near, float z_far, uint8_t* out_color, float* out_depth) {
    PE_REQUIRE(n_vertices >= 0 && n_faces >= 0 && n_views >= 0 && rows > 0 && cols > 0, "bad sizes");
    PE_REQUIRE((uv == nullptr) == (tex == nullptr), "texture coordinates and texture image go together");
    if (n_views == 0) return PE_OK;
    PE_REQUIRE(out_color && out_depth && poses && cam4, "null pointer");
    cudaStream_t stream;
    PE_CHECK(render_ctx(nullptr, device, &stream));
    DevBuf b_v, b_c, b_f, b_p, b_col, b_dep, b_uv, b_tex;
    int st = PE_OK;
#define UP(buf, src, bytes) do { const size_t nb_ = (bytes); if (st == PE_OK) st = (buf).reserve(nb_ > 0 ? nb_ : 4); \
        if (st == PE_OK && nb_ > 0 && cudaMemcpy((buf).p, src, nb_, cudaMemcpyHostToDevice) != cudaSuccess) { set_error("H2D copy failed"); st = PE_ERR_CUDA; } } while (0)
    UP(b_v, vertices, (size_t)n_vertices * 12);
    UP(b_c, colors, (size_t)n_vertices * 12);
    UP(b_f, faces, (size_t)n_faces * 12);
    UP(b_p, poses, (size_t)n_views * 48);
    if (uv) {
        UP(b_uv, uv, (size_t)n_vertices * 8);
        UP(b_tex, tex, (size_t)tex_rows * tex_cols * 3);
    }
#undef UP
    const size_t px = (size_t)rows * cols * n_views;
    if (st == PE_OK) st = b_col.reserve(px * 3);
    if (st == PE_OK) st = b_dep.reserve(px * 4);
    if (st == PE_OK)
        st = pe_render_views_tex_dev(nullptr, device, b_v.as<float>(), b_c.as<float>(), uv ? b_uv.as<float>() : nullptr, n_vertices, b_f.as<int32_t>(),
                                     n_faces, uv ? b_tex.as<uint8_t>() : nullptr, tex_rows, tex_cols, cam4, b_p.as<float>(), n_views, rows, cols,
                                     z_near, z_far, b_col.as<uint8_t>(), b_dep.as<float>());
    if (st == PE_OK && (cudaMemcpy(out_color, b_col.p, px * 3, cudaMemcpyDeviceToHost) != cudaSuccess ||
                        cudaMemcpy(out_depth, b_dep.p, px * 4, cudaMemcpyDeviceToHost) != cudaSuccess)) {
        set_error("render read-back failed: %s", cudaGetErrorString(cudaGetLastError()));
        st = PE_ERR_CUDA;
    }
    for (DevBuf* b : {&b_v, &b_c, &b_f, &b_p, &b_col, &b_dep, &b_uv, &b_tex}) b->release();
    return st;
}

int pe_render_views(int device, const float* vertices, const float* colors, int n_vertices, const int32_t* faces, int n_faces,
                    const float* cam4, const float* poses, int n_views, int rows, int cols, float z_near, float z_far,
                    uint8_t* out_color, float* out_depth) {
    return pe_render_views_tex(device, vertices, colors, nullptr, n_vertices, faces, n_faces, nullptr, 0, 0, cam4, poses, n_views, rows, cols,
                               z_near, z_far, out_color, out_depth);
}

}  // extern "C"
