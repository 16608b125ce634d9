/*
 * oracle_patch.c - CPU restatement of the reference's patch extraction (TEST INFRASTRUCTURE: only tests/,
 * __graft_entry__.smoke() and bench.py's CPU legs may load it; see pose_oracle.h).
 *
 *   orc_depth2normals   src/helper.cpp:53-103     3x3 stencil (step 3) least-squares normal per pixel
 *   orc_sample_patch    src/datasetmanager.cpp:299-346 (samplePatchWithScale): metric 0.2 m cut scaled by fx*m/z,
 *                      clamped into the frame, colour / 255, depth de-meaned by z, clamped to +-m, rescaled to [0,1],
 *                      colour + normals resized to 64x64 with cv::resize's default INTER_LINEAR, depth with
 *                      INTER_NEAREST, channels merged as {B,G,R, depth, nx,ny,nz}.
 *
 * cv::resize is restated from OpenCV's imgproc/resize.cpp (3.1 and 4.x agree on this path):
 *   INTER_LINEAR, float data: fx = (float)((dx + 0.5) * scale - 0.5); sx = floor(fx); fx -= sx; sx < 0 -> (0, 0);
 *   sx >= w - 1 -> (w - 1, 0); horizontal pass S[sx]*(1-fx) + S[sx+1]*fx per source row, then vertical pass
 *   r0*(1-fy) + r1*fy; when both scale factors are exactly 2 the call is rerouted to INTER_AREA (mean of the 2x2 block);
 *   INTER_NEAREST: sx = min(floor(dx * scale), w - 1).
 * Products and sums are rounded separately (no FMA contraction: compiled with -ffp-contract=off).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "pose_oracle.h"

/* src/helper.cpp:53-103.  dep: rows x cols f32; nor: rows x cols x 3 f32 (zero where undefined) */
void orc_depth2normals(const float* dep, int rows, int cols, float fx, float fy, float ox, float oy, float* nor) {
    (void)ox;
    (void)oy;
    const int N = 3, stride = cols;
    memset(nor, 0, (size_t)rows * cols * 3 * sizeof(float));
    for (int r = N; r < rows - N - 1; ++r)
        for (int c = N; c < cols - N - 1; ++c) {
            const float* p = dep + (size_t)r * stride + c;
            const float d = *p;
            if (!(d > 0)) continue;
            float A0 = 0, A1 = 0, A3 = 0, b0 = 0, b1 = 0;
            for (int i = -N; i <= N; i += N)
                for (int j = -N; j <= N; j += N) {
                    const float delta = p[i + j * stride] - d;
                    const float f = fabsf(delta) < 0.05f ? 1.f : 0.f;
                    const float fi = f * (float)i, fj = f * (float)j;
                    A0 += fi * (float)i;
                    A1 += fi * (float)j;
                    A3 += fj * (float)j;
                    b0 += fi * delta;
                    b1 += fj * delta;
                }
            const float n0 = (A3 * b0 - A1 * b1) * fx;
            const float n1 = (-A1 * b0 + A0 * b1) * fy;
            const float n2 = (A0 * A3 - A1 * A1) * d;
            const float sq = n0 * n0 + n1 * n1 + n2 * n2;      /* Eigen squaredNorm: x^2 + y^2 + z^2 in this order */
            if (sq > 0) {
                const float s = sqrtf(sq);
                float* o = nor + ((size_t)r * cols + c) * 3;
                o[0] = n0 / s;
                o[1] = n1 / s;
                o[2] = n2 / s;
            }
        }
}

/* the cut rectangle of samplePatchWithScale (:301-316); returns 0 when the reference's asserts would fail */
int orc_patch_rect(int rows, int cols, int center_x, int center_y, float z, float fx, float fy, float m, int* x, int* y, int* w, int* h) {
    const int screenW = (int)(fx * m / z), screenH = (int)(fy * m / z);
    int cx = center_x - screenW / 2, cy = center_y - screenH / 2;
    if (cx < 0) cx = 0;
    if (cy < 0) cy = 0;
    if (cx > cols - screenW - 1) cx = cols - screenW - 1;
    if (cy > rows - screenH - 1) cy = rows - screenH - 1;
    *x = cx; *y = cy; *w = screenW; *h = screenH;
    return screenW > 0 && screenH > 0 && cx >= 0 && cx < cols - screenW && cy >= 0 && cy < rows - screenH;
}

/* horizontal taps: clamped at both borders (a single tap of weight 1 there) */
static void linear_coeff_x(int dsize, int ssize, int* ofs, float* a1) {
    const double scale = 1. / ((double)dsize / ssize);      /* cv::resize: scale_x = 1. / inv_scale_x */
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        ofs[d] = s;
        a1[d] = f;
    }
}
/* vertical taps: the weights are NOT clamped, the two row indices are clipped into the image instead */
static void linear_coeff_y(int dsize, int ssize, int* ofs, float* b1) {
    const double scale = 1. / ((double)dsize / ssize);
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        const int s = (int)floorf(f);
        f -= (float)s;
        ofs[d] = s;
        b1[d] = f;
    }
}
static int clipi(int v, int n) { return v < 0 ? 0 : (v >= n ? n - 1 : v); }

/* cv::resize(src(w x h, cn floats, row pitch `pitch` floats) -> 64 x 64), default interpolation */
static void resize_linear64(const float* src, int pitch, int w, int h, int cn, float* dst, int dpitch, int dofs) {
    if (w == 128 && h == 128) {       /* both scale factors exactly 2: INTER_LINEAR is rerouted to INTER_AREA */
        for (int dy = 0; dy < 64; ++dy)
            for (int dx = 0; dx < 64; ++dx)
                for (int c = 0; c < cn; ++c) {
                    const float* s0 = src + (size_t)(2 * dy) * pitch + (2 * dx) * cn + c;
                    const float* s1 = s0 + pitch;
                    dst[((size_t)dy * 64 + dx) * dpitch + dofs + c] = (s0[0] + s0[cn] + s1[0] + s1[cn]) * 0.25f;
                }
        return;
    }
    int xo[64], yo[64];
    float xa[64], ya[64];
    linear_coeff_x(64, w, xo, xa);
    linear_coeff_y(64, h, yo, ya);
    for (int dy = 0; dy < 64; ++dy) {
        const int y0 = clipi(yo[dy], h), y1 = clipi(yo[dy] + 1, h);
        const float b1 = ya[dy], b0 = 1.f - b1;
        for (int dx = 0; dx < 64; ++dx) {
            const int x0 = xo[dx];
            const float a1 = xa[dx], a0 = 1.f - a1;
            for (int c = 0; c < cn; ++c) {
                float r0, r1;
                if (x0 >= w - 1) {   /* right border: a single tap */
                    r0 = src[(size_t)y0 * pitch + x0 * cn + c];
                    r1 = src[(size_t)y1 * pitch + x0 * cn + c];
                } else {
                    r0 = src[(size_t)y0 * pitch + x0 * cn + c] * a0 + src[(size_t)y0 * pitch + (x0 + 1) * cn + c] * a1;
                    r1 = src[(size_t)y1 * pitch + x0 * cn + c] * a0 + src[(size_t)y1 * pitch + (x0 + 1) * cn + c] * a1;
                }
                dst[((size_t)dy * 64 + dx) * dpitch + dofs + c] = r0 * b0 + r1 * b1;
            }
        }
    }
}

/* src/datasetmanager.cpp:299-346.  color: rows x cols x 3 u8 (BGR as imread gives it); depth: rows x cols f32 (metres);
 * normals: rows x cols x 3 f32; out: 64 x 64 x 7 f32.  Returns 0 on success, 1 when the reference's asserts fail. */
int orc_sample_patch(const uint8_t* color, const float* depth, const float* normals, int rows, int cols, int center_x, int center_y,
                    float z, float fx, float fy, float m, float* out) {
    int x, y, w, h;
    if (!orc_patch_rect(rows, cols, center_x, center_y, z, fx, fy, m, &x, &y, &w, &h)) return 1;
    float* col = (float*)malloc((size_t)w * h * 3 * sizeof(float));
    if (!col) return 2;
    const float inv255 = 1 / 255.f;
    for (int r = 0; r < h; ++r)
        for (int c = 0; c < w * 3; ++c) col[(size_t)r * w * 3 + c] = (float)color[((size_t)(y + r) * cols + x) * 3 + c] * inv255;
    resize_linear64(col, w * 3, w, h, 3, out, 7, 0);
    free(col);
    resize_linear64(normals + ((size_t)y * cols + x) * 3, cols * 3, w, h, 3, out, 7, 4);
    const double sx = 1. / (64. / w), sy = 1. / (64. / h);       /* resizeNN: ifx = 1. / inv_scale_x */
    const float inv_m = 1.0f / m;
    for (int dy = 0; dy < 64; ++dy) {
        int yy = (int)floor(dy * sy);
        if (yy > h - 1) yy = h - 1;
        for (int dx = 0; dx < 64; ++dx) {
            int xx = (int)floor(dx * sx);
            if (xx > w - 1) xx = w - 1;
            float d = depth[(size_t)(y + yy) * cols + x + xx];
            d -= z;
            if (d < -m) d = -m;
            if (d > m) d = m;
            d *= inv_m;
            d = (d + 1.f) * 0.5f;
            out[((size_t)dy * 64 + dx) * 7 + 3] = d;
        }
    }
    return 0;
}
