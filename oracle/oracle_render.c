/*
 * oracle_render.c - CPU restatement of the reference's template renderer (TEST INFRASTRUCTURE, see pose_oracle.h).
 *
 * The reference renders through Qt / OpenGL (src/painter.cpp, src/sphere.cpp, src/model.cpp): unavailable on this box,
 * and it has no CPU rasteriser, so PARITY IS UNPINNED for this file: it restates the maths of the fixed-function
 * pipeline the reference configures, and is pinned only against analytic scenes (tests/test_render_oracle.py).
 *
 *   orc_init_sphere        SphereRenderer::initSphere / subdivide (src/sphere.cpp:165-226): icosahedron subdivision with
 *                          the reference's linear-search de-duplication, in its insertion order
 *   orc_view_pose          SphereRenderer::createTransformation / computeRotation (src/sphere.cpp:27-35,134-146)
 *   orc_render             Painter::paint with RealWorldCamera + Model (src/painter.cpp:86-110,186-194,231-248;
 *                          src/model.cpp:29-54): per-vertex colour, no lighting, no culling, depth test GL_LESS,
 *                          black background, BGR 8-bit colour read-back, z-buffer converted to metric depth
 *                          (0 = background).
 * Pipeline maths: camera point Xc = R X + t; the reference's projection (K [R|t] with the z row replaced, glOrtho over
 * the viewport) gives window x = fx Xc/Zc + ox, y = fy Yc/Zc + oy (image orientation, no flip) and a window depth that
 * is affine in 1/Zc, which convertZBufferToDepth maps back to Zc.  Rasterisation as GPUs do it: vertices snapped to 1/256
 * pixel, exact integer edge functions at pixel centres, top-left style tie rule (each shared edge belongs to exactly one
 * triangle), screen-space barycentrics, 1/Z interpolated linearly (= GL's window depth), colour perspective-correct,
 * round-to-nearest to 8 bits; triangles that cross the near plane are clipped; textured meshes (BigBIRD) modulate the
 * vertex colour with a bilinear GL_REPEAT texture fetch.  Not restated: the 24-bit quantisation of GL's depth buffer
 * (<= 0.06 mm at 1 m with near 0.001 / far 4.0) and the fixed-point weights of a hardware texture unit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "pose_oracle.h"

/* ---------------------------------------------------------------- view sphere ---- */
typedef struct { float x, y, z; } v3;

static v3 v3_normalized(v3 a) {
    const float n = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
    v3 r = {a.x / n, a.y / n, a.z / n};
    return r;
}
static v3 v3_add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static v3 v3_cross(v3 a, v3 b) { v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }
static int v3_eq(v3 a, v3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

typedef struct { v3* p; int n, cap; } vlist;

static void push_unique3(vlist* s, v3 v1, v3 v2, v3 v3_) {
    int f1 = -1, f2 = -1, f3 = -1;
    for (int i = 0; i < s->n; ++i) {
        if (v3_eq(s->p[i], v1)) f1 = i;
        if (v3_eq(s->p[i], v2)) f2 = i;
        if (v3_eq(s->p[i], v3_)) f3 = i;
        if (f1 != -1 && f2 != -1 && f3 != -1) break;
    }
    if (f1 == -1 && s->n < s->cap) s->p[s->n++] = v1;
    if (f2 == -1 && s->n < s->cap) s->p[s->n++] = v2;
    if (f3 == -1 && s->n < s->cap) s->p[s->n++] = v3_;
}

static void subdivide(vlist* s, v3 v1, v3 v2, v3 v3_, int depth) {
    if (depth == 0) { push_unique3(s, v1, v2, v3_); return; }
    const v3 v12 = v3_normalized(v3_add(v1, v2)), v23 = v3_normalized(v3_add(v2, v3_)), v31 = v3_normalized(v3_add(v3_, v1));
    subdivide(s, v1, v12, v31, depth - 1);
    subdivide(s, v2, v23, v12, depth - 1);
    subdivide(s, v3_, v31, v23, depth - 1);
    subdivide(s, v12, v23, v31, depth - 1);
}

/* src/sphere.cpp:186-226; returns the number of vertices written (at most cap) */
int orc_init_sphere(int depth, float* out_xyz, int cap) {
    static const int ind[20][3] = {{0, 4, 1}, {0, 9, 4}, {9, 5, 4}, {4, 5, 8}, {4, 8, 1}, {8, 10, 1}, {8, 3, 10}, {5, 3, 8}, {5, 2, 3}, {2, 7, 3},
                                   {7, 10, 3}, {7, 6, 10}, {7, 11, 6}, {11, 0, 6}, {0, 1, 6}, {6, 1, 10}, {9, 0, 11}, {9, 11, 2}, {9, 2, 5}, {7, 2, 11}};
    const float X = 0.525731112119133606f, Z = 0.850650808352039932f;
    const v3 raw[12] = {{-X, 0, +Z}, {+X, 0, +Z}, {-X, 0, -Z}, {+X, 0, -Z}, {0, +Z, +X}, {0, +Z, -X},
                        {0, -Z, +X}, {0, -Z, -X}, {+Z, +X, 0}, {-Z, +X, 0}, {+Z, -X, 0}, {-Z, -X, 0}};
    v3 v[12];
    for (int i = 0; i < 12; ++i) v[i] = v3_normalized(raw[i]);
    vlist s;
    s.p = (v3*)malloc(sizeof(v3) * (size_t)(cap > 0 ? cap : 1));
    s.n = 0;
    s.cap = cap;
    for (int i = 0; i < 20; ++i) subdivide(&s, v[ind[i][0]], v[ind[i][1]], v[ind[i][2]], depth);
    for (int i = 0; i < s.n; ++i) {
        const v3 p = v3_normalized(s.p[i]);
        out_xyz[3 * i] = p.x; out_xyz[3 * i + 1] = p.y; out_xyz[3 * i + 2] = p.z;
    }
    const int n = s.n;
    free(s.p);
    return n;
}

/* src/sphere.cpp:27-35 + 134-146: pose (row-major 3x4 [R|t]) of the camera looking at the origin from `vertex`,
 * in-plane rotation `angle_deg`, object at distance `scale` */
void orc_view_pose(const float* vertex, float scale, float angle_deg, float* out_rt) {
    const v3 eye = {vertex[0], vertex[1], vertex[2]};
    v3 up = {0, 0, 1};
    if (eye.x == 0 && eye.y == 0 && eye.z != 0) { up.x = -1; up.y = 0; up.z = 0; }
    const v3 en = v3_normalized(eye);
    const v3 c2 = {-en.x, -en.y, -en.z};
    const v3 c0 = v3_normalized(v3_cross(c2, v3_normalized(up)));
    const v3 mc2 = {-c2.x, -c2.y, -c2.z};
    const v3 c1 = v3_cross(c0, mc2);
    /* rot = [c0 c1 c2]^T: rows are c0, c1, c2 */
    const float rot[3][3] = {{c0.x, c0.y, c0.z}, {c1.x, c1.y, c1.z}, {c2.x, c2.y, c2.z}};
    /* AngleAxisf(angle * CV_PI / 180.0, z): the angle is computed in double and narrowed to float */
    const float a = (float)((double)angle_deg * 3.1415926535897932384626433832795 / 180.0);
    const float ca = cosf(a), sa = sinf(a);
    const float rz[3][3] = {{ca, -sa, 0}, {sa, ca, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) out_rt[4 * i + j] = rz[i][0] * rot[0][j] + rz[i][1] * rot[1][j] + rz[i][2] * rot[2][j];
        out_rt[4 * i + 3] = i == 2 ? scale : 0.f;
    }
}

/* -------------------------------------------------------------------- rasteriser ---- */
static int64_t edge(int64_t ax, int64_t ay, int64_t bx, int64_t by, int64_t px, int64_t py) {
    return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}
/* a pixel centre exactly on the edge a->b of a positively oriented triangle belongs to it iff ... (complementary for b->a) */
static int owns_edge(int64_t ax, int64_t ay, int64_t bx, int64_t by) {
    const int64_t dx = bx - ax, dy = by - ay;
    return dy < 0 || (dy == 0 && dx > 0);
}

/* vertices [nv][3], colors [nv][3] RGB in [0,1], faces [nf][3]; cam = {fx, fy, ox, oy}; rt = row-major 3x4;
 * out_color [rows][cols][3] BGR u8, out_depth [rows][cols] f32 metres (0 = background).
 * uv [nv][2] + tex [tex_rows][tex_cols][3] BGR u8 (both or neither): GL_TEXTURE_2D with the fixed-function default
 * GL_MODULATE (fragment = vertex colour x texel), GL_REPEAT wrap, GL_LINEAR filtering of level 0, the image rows in
 * the order glTexImage2D receives them (t = 0 at row 0), as Model::loadModel / bindVBOs set it up
 * (src/model.cpp:330-350, src/painter.cpp:140-160).
 * Near-plane clipping (GL clips every primitive against the frustum; the reference's near is 0.001, src/painter.cpp:9):
 * a triangle with vertices on both sides of Zc = z_near is cut by Sutherland-Hodgman in camera space, every cut point
 * computed from the inside vertex towards the outside one (so both triangles at a shared edge get the same point),
 * attributes interpolated with the same parameter; a quad is split as a fan from its first vertex.  The side and far
 * planes need no geometry: coverage is clamped to the frame and z < z_far is tested per fragment. */
typedef struct { int32_t x, y; float z; int i0, i1; float t; } cvert;

static void project_snap(const float* cam, float xc, float yc, float zc, int32_t* px, int32_t* py) {
    const float u = (cam[0] * xc + cam[2] * zc) / zc, v = (cam[1] * yc + cam[3] * zc) / zc;
    /* snap to 1/256 pixel; clamp far-off vertices so that the integer edge functions cannot overflow */
    float us = u * 256.f, vs = v * 256.f;
    const float lim = 1048576.f * 256.f;
    if (us > lim) us = lim;
    if (us < -lim) us = -lim;
    if (vs > lim) vs = lim;
    if (vs < -lim) vs = -lim;
    *px = (int32_t)floorf(us + 0.5f);
    *py = (int32_t)floorf(vs + 0.5f);
}

static float attr_at(const float* a, int stride, int ch, const cvert* v) {
    const float a0 = a[stride * v->i0 + ch];
    if (v->i0 == v->i1) return a0;
    return a0 + v->t * (a[stride * v->i1 + ch] - a0);
}

/* GL_LINEAR, GL_REPEAT: texel centres at (i + 0.5) / size */
static float tex_bilinear(const uint8_t* tex, int th, int tw, float s, float t, int ch_rgb) {
    const float fu = s * (float)tw - 0.5f, fv = t * (float)th - 0.5f;
    const float x0f = floorf(fu), y0f = floorf(fv);
    const float ax = fu - x0f, ay = fv - y0f;
    int x0 = (int)fmodf(x0f, (float)tw), y0 = (int)fmodf(y0f, (float)th);
    if (x0 < 0) x0 += tw;
    if (y0 < 0) y0 += th;
    const int x1 = x0 + 1 == tw ? 0 : x0 + 1, y1 = y0 + 1 == th ? 0 : y0 + 1;
    const int c = 2 - ch_rgb;                                     /* the image is BGR */
    const float t00 = tex[((size_t)y0 * tw + x0) * 3 + c], t10 = tex[((size_t)y0 * tw + x1) * 3 + c];
    const float t01 = tex[((size_t)y1 * tw + x0) * 3 + c], t11 = tex[((size_t)y1 * tw + x1) * 3 + c];
    const float top = t00 + ax * (t10 - t00), bot = t01 + ax * (t11 - t01);
    return (top + ay * (bot - top)) / 255.f;
}

void orc_render_tex(const float* vertices, const float* colors, const float* uv, int nv, const int32_t* faces, int nf,
                    const uint8_t* tex, int tex_rows, int tex_cols, const float* cam, const float* rt, int rows, int cols,
                    float z_near, float z_far, uint8_t* out_color, float* out_depth) {
    float* cc = (float*)malloc(sizeof(float) * (size_t)(nv > 0 ? nv : 1) * 3);       /* camera-space points */
    int32_t* fx_ = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nv > 0 ? nv : 1) * 2);
    const int textured = uv && tex && tex_rows > 0 && tex_cols > 0;
    for (int i = 0; i < nv; ++i) {
        const float x = vertices[3 * i], y = vertices[3 * i + 1], z = vertices[3 * i + 2];
        const float xc = rt[0] * x + rt[1] * y + rt[2] * z + rt[3];
        const float yc = rt[4] * x + rt[5] * y + rt[6] * z + rt[7];
        const float zc = rt[8] * x + rt[9] * y + rt[10] * z + rt[11];
        cc[3 * i] = xc; cc[3 * i + 1] = yc; cc[3 * i + 2] = zc;
        fx_[2 * i] = fx_[2 * i + 1] = 0;
        if (zc > z_near) project_snap(cam, xc, yc, zc, &fx_[2 * i], &fx_[2 * i + 1]);
    }
    memset(out_color, 0, (size_t)rows * cols * 3);
    for (size_t i = 0; i < (size_t)rows * cols; ++i) out_depth[i] = INFINITY;
    for (int f = 0; f < nf; ++f) {
        const int idx[3] = {faces[3 * f], faces[3 * f + 1], faces[3 * f + 2]};
        if (idx[0] < 0 || idx[1] < 0 || idx[2] < 0 || idx[0] >= nv || idx[1] >= nv || idx[2] >= nv) continue;
        /* ---- clip against Zc = z_near (a vertex exactly on the plane counts as outside) */
        cvert poly[4];
        int np_ = 0;
        const int in[3] = {cc[3 * idx[0] + 2] > z_near, cc[3 * idx[1] + 2] > z_near, cc[3 * idx[2] + 2] > z_near};
        if (!in[0] && !in[1] && !in[2]) continue;
        for (int e = 0; e < 3; ++e) {
            const int p = e, q = (e + 1) % 3;
            if (in[p]) {
                cvert v = {fx_[2 * idx[p]], fx_[2 * idx[p] + 1], cc[3 * idx[p] + 2], idx[p], idx[p], 0.f};
                poly[np_++] = v;
            }
            if (in[p] != in[q]) {
                const int a = in[p] ? idx[p] : idx[q], b = in[p] ? idx[q] : idx[p];      /* a inside, b outside */
                const float t = (cc[3 * a + 2] - z_near) / (cc[3 * a + 2] - cc[3 * b + 2]);
                const float xc = cc[3 * a] + t * (cc[3 * b] - cc[3 * a]), yc = cc[3 * a + 1] + t * (cc[3 * b + 1] - cc[3 * a + 1]);
                cvert v = {0, 0, z_near, a, b, t};
                project_snap(cam, xc, yc, z_near, &v.x, &v.y);
                poly[np_++] = v;
            }
        }
        for (int sub = 0; sub + 2 < np_; ++sub) {
            cvert va = poly[0], vb = poly[sub + 1], vc = poly[sub + 2];
            int64_t ax = va.x, ay = va.y, bx = vb.x, by = vb.y, cx = vc.x, cy = vc.y;
            int64_t area = edge(ax, ay, bx, by, cx, cy);
            if (area == 0) continue;
            if (area < 0) {      /* no culling (src/model.cpp:29-54): flip to positive orientation */
                cvert tv = vb; vb = vc; vc = tv;
                int64_t tx = bx, ty = by; bx = cx; by = cy; cx = tx; cy = ty;
                area = -area;
            }
            int64_t minx = ax < bx ? ax : bx, maxx = ax > bx ? ax : bx, miny = ay < by ? ay : by, maxy = ay > by ? ay : by;
            if (cx < minx) minx = cx;
            if (cx > maxx) maxx = cx;
            if (cy < miny) miny = cy;
            if (cy > maxy) maxy = cy;
            /* pixel px has its centre at 256 px + 128 */
            int x0 = (int)((minx - 128 + 255) >> 8), x1 = (int)((maxx - 128) >> 8), y0 = (int)((miny - 128 + 255) >> 8), y1 = (int)((maxy - 128) >> 8);
            if (x0 < 0) x0 = 0;
            if (y0 < 0) y0 = 0;
            if (x1 > cols - 1) x1 = cols - 1;
            if (y1 > rows - 1) y1 = rows - 1;
            const int oa = owns_edge(bx, by, cx, cy), ob = owns_edge(cx, cy, ax, ay), oc = owns_edge(ax, ay, bx, by);
            const float iza = 1.f / va.z, izb = 1.f / vb.z, izc = 1.f / vc.z;
            const float farea = (float)area;
            for (int py = y0; py <= y1; ++py)
                for (int px = x0; px <= x1; ++px) {
                    const int64_t qx = 256 * (int64_t)px + 128, qy = 256 * (int64_t)py + 128;
                    const int64_t wa = edge(bx, by, cx, cy, qx, qy), wb = edge(cx, cy, ax, ay, qx, qy), wc = edge(ax, ay, bx, by, qx, qy);
                    if (wa < 0 || wb < 0 || wc < 0) continue;
                    if ((wa == 0 && !oa) || (wb == 0 && !ob) || (wc == 0 && !oc)) continue;
                    const float ba = (float)wa / farea, bb = (float)wb / farea, bc = (float)wc / farea;
                    const float iz = ba * iza + bb * izb + bc * izc;
                    const float z = 1.f / iz;
                    if (!(z > z_near) || !(z < z_far)) continue;
                    float* dp = out_depth + (size_t)py * cols + px;
                    if (!(z < *dp)) continue;          /* GL_LESS: the first drawn fragment keeps an equal depth */
                    *dp = z;
                    uint8_t* cp = out_color + ((size_t)py * cols + px) * 3;
                    float s = 0.f, t = 0.f;
                    if (textured) {
                        s = (ba * (attr_at(uv, 2, 0, &va) * iza) + bb * (attr_at(uv, 2, 0, &vb) * izb) + bc * (attr_at(uv, 2, 0, &vc) * izc)) / iz;
                        t = (ba * (attr_at(uv, 2, 1, &va) * iza) + bb * (attr_at(uv, 2, 1, &vb) * izb) + bc * (attr_at(uv, 2, 1, &vc) * izc)) / iz;
                    }
                    for (int ch = 0; ch < 3; ++ch) {
                        float c = (ba * (attr_at(colors, 3, ch, &va) * iza) + bb * (attr_at(colors, 3, ch, &vb) * izb) +
                                   bc * (attr_at(colors, 3, ch, &vc) * izc)) / iz;
                        if (textured) c = c * tex_bilinear(tex, tex_rows, tex_cols, s, t, ch);      /* GL_MODULATE */
                        float q = c * 255.f + 0.5f;
                        if (q < 0.f) q = 0.f;
                        if (q > 255.f) q = 255.f;
                        cp[2 - ch] = (uint8_t)q;        /* BGR read-back */
                    }
                }
        }
    }
    for (size_t i = 0; i < (size_t)rows * cols; ++i)
        if (out_depth[i] == INFINITY) out_depth[i] = 0.f;
    free(cc);
    free(fx_);
}

void orc_render(const float* vertices, const float* colors, int nv, const int32_t* faces, int nf, const float* cam,
                const float* rt, int rows, int cols, float z_near, float z_far, uint8_t* out_color, float* out_depth) {
    orc_render_tex(vertices, colors, 0, nv, faces, nf, 0, 0, 0, cam, rt, rows, cols, z_near, z_far, out_color, out_depth);
}

/* ------------------------------------------------------------------ template labels ---- */
/* The 8-float label datasetManager::createSynthSamplesPaul / createSynthSamplesWadim attach to a rendered view
 * (src/datasetmanager.cpp:484-491, 523-530):  {index, Quaternionf(pose.linear()).coeffs() = (x, y, z, w),
 * pose.inverse().translation()}.  Eigen (3.2 and 3.3 alike) builds a quaternion from a rotation matrix with Shoemake's
 * branch on the trace (Quaternion.h quaternionbase_assign_impl<Other,3,3>); Isometry3f::inverse() is
 * [R^T | -(R^T t)].  rt = row-major 3x4 [R|t]. */
void orc_pose_label(const float* rt, int index, float* out8) {
    const float m[3][3] = {{rt[0], rt[1], rt[2]}, {rt[4], rt[5], rt[6]}, {rt[8], rt[9], rt[10]}};
    float q[4];   /* x y z w */
    float t = m[0][0] + m[1][1] + m[2][2];
    if (t > 0.f) {
        t = sqrtf(t + 1.0f);
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
        t = sqrtf(m[i][i] - m[j][j] - m[k][k] + 1.0f);
        q[i] = 0.5f * t;
        t = 0.5f / t;
        q[3] = (m[k][j] - m[j][k]) * t;
        q[j] = (m[j][i] + m[i][j]) * t;
        q[k] = (m[k][i] + m[i][k]) * t;
    }
    out8[0] = (float)index;
    for (int c = 0; c < 4; ++c) out8[1 + c] = q[c];
    for (int r = 0; r < 3; ++r)      /* -(R^T t): row r of R^T is column r of R */
        out8[5 + r] = -(m[0][r] * rt[3] + m[1][r] * rt[7] + m[2][r] * rt[11]);
}

/* datasetManager::extractRealSamplesPaul / extractRealSamplesWadim (src/datasetmanager.cpp:348-411): where the patch of
 * a real frame is cut.  centroid = pose translation (LineMOD, Washington, and the Wadim variant) or pose * model
 * centroid (BigBIRD: centroid3 != NULL); projCentroid = cam * centroid (Eigen's 3x3 product with the zeros of the
 * intrinsic matrix included), divided by its z; the float coordinates are passed to int parameters (truncation).
 * Returns z = centroid(2) (the Paul variant; the Wadim variant reads the depth frame at the pixel instead). */
float orc_project_centroid(const float* cam /* fx fy ox oy */, const float* rt, const float* centroid3, int32_t* out_xy) {
    float c[3];
    if (centroid3) {
        for (int r = 0; r < 3; ++r)
            c[r] = (rt[4 * r] * centroid3[0] + rt[4 * r + 1] * centroid3[1] + rt[4 * r + 2] * centroid3[2]) + rt[4 * r + 3];
    } else {
        c[0] = rt[3]; c[1] = rt[7]; c[2] = rt[11];
    }
    const float px = (cam[0] * c[0] + 0.f * c[1]) + cam[2] * c[2];
    const float py = (0.f * c[0] + cam[1] * c[1]) + cam[3] * c[2];
    const float pz = (0.f * c[0] + 0.f * c[1]) + 1.f * c[2];
    out_xy[0] = (int32_t)(px / pz);
    out_xy[1] = (int32_t)(py / pz);
    return c[2];
}
