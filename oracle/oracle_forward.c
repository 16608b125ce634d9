/*
 * oracle_forward.c - CPU restatement of the Caffe TEST-phase forward that
 * networkEvaluator::computeDescriptors drives (reference src/networkevaluator.cpp:9-54).
 * TEST INFRASTRUCTURE ONLY - see pose_oracle.h for the parity statement.
 *
 * Layer semantics restated from BVLC Caffe 1.0 (not vendored by the reference;
 * README.md:21 pins "latest release" + PR #2603):
 *   Convolution  cross-correlation, weight [out][in][kh][kw], + bias, stride 1, pad 0
 *                (network/manifold_rgbdnor_16.prototxt:6-20, 41-55)
 *   Pooling MAX  pooled = ceil((in - k) / s) + 1, window clipped to the input,
 *                running max starts at -FLT_MAX            (:22-31, 57-66)
 *   ReLU         max(x,0)                                  (network/manifold.prototxt:33-38)
 *   ELU          max(x,0) + alpha*(exp(min(x,0)) - 1), alpha = 1
 *                                                          (network/manifold_rgbdnor_16.prototxt:33-38)
 *   PReLU        max(x,0) + a[c]*min(x,0), c = axis-1 index (network/manifold_rgbd_16.prototxt:64-69)
 *   InnerProduct flatten C,H,W; weight [out][in]; + bias   (:75-87)
 * The HWC -> CHW staging copy of src/networkevaluator.cpp:33-42 and the
 * Slice/Concat/Silence channel selection (network/manifold_rgb_16.prototxt:6-33)
 * are folded into the first gather.
 */
#include "pose_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define CLONES __attribute__((target_clones("avx512f", "avx2", "default")))

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

typedef struct {
    int c, h, w;
} shape3;

static int pooled_size(int in, int k, int s) {
    /* Caffe pooling_layer.cpp: static_cast<int>(ceil(static_cast<float>(in - k) / s)) + 1 */
    return (int)ceilf((float)(in - k) / (float)s) + 1;
}

static int layer_out_shape(const orc_layer* L, shape3 in, shape3* out) {
    switch (L->kind) {
    case ORC_CONV:
        if (in.h < L->b || in.w < L->b) return -1;
        out->c = L->a; out->h = in.h - L->b + 1; out->w = in.w - L->b + 1;
        return 0;
    case ORC_POOL:
        if (in.h < L->a || in.w < L->a) return -1;
        out->c = in.c; out->h = pooled_size(in.h, L->a, L->b); out->w = pooled_size(in.w, L->a, L->b);
        return 0;
    case ORC_ACT:
        *out = in;
        return 0;
    case ORC_FC:
        out->c = L->a; out->h = 1; out->w = 1;
        return 0;
    }
    return -1;
}

/* ---- convolution: register-blocked over 4 filters x one output row.  Every
 * output is accumulated as bias, then + w*x over (c, ky, kx) ascending. ---- */
#define ROWP 64 /* padded output-row length handled by the fast path */

CLONES static void conv_rows_fast(const float* inpad, int C, int H, int WP, const float* wt,
                                  const float* bias, int nf, int k, int OH, int OW, float* out) {
    for (int y = 0; y < OH; ++y) {
        for (int f0 = 0; f0 < nf; f0 += 4) {
            const int fb_n = nf - f0 < 4 ? nf - f0 : 4;
            float acc[4][ROWP];
            for (int fb = 0; fb < 4; ++fb) {
                const float b = fb < fb_n ? bias[f0 + fb] : 0.f;
                for (int x = 0; x < ROWP; ++x) acc[fb][x] = b;
            }
            for (int c = 0; c < C; ++c)
                for (int ky = 0; ky < k; ++ky) {
                    const float* ip = inpad + ((size_t)c * H + (y + ky)) * WP;
                    for (int kx = 0; kx < k; ++kx) {
                        float wv[4];
                        for (int fb = 0; fb < 4; ++fb)
                            wv[fb] = fb < fb_n ? wt[(((size_t)(f0 + fb) * C + c) * k + ky) * k + kx] : 0.f;
                        const float* ipx = ip + kx;
                        for (int x = 0; x < ROWP; ++x) {
                            const float v = ipx[x];
                            acc[0][x] += wv[0] * v;
                            acc[1][x] += wv[1] * v;
                            acc[2][x] += wv[2] * v;
                            acc[3][x] += wv[3] * v;
                        }
                    }
                }
            for (int fb = 0; fb < fb_n; ++fb)
                memcpy(out + ((size_t)(f0 + fb) * OH + y) * OW, acc[fb], sizeof(float) * OW);
        }
    }
}

static void conv_forward(const float* in, shape3 s, const float* wt, const float* bias, int nf,
                         int k, float* out, float* pad_scratch) {
    const int OH = s.h - k + 1, OW = s.w - k + 1;
    if (OW <= ROWP) {
        /* zero-padded copy so the fixed-length row loop never reads out of bounds */
        const int WP = ROWP + k;
        memset(pad_scratch, 0, sizeof(float) * (size_t)s.c * s.h * WP);
        for (int c = 0; c < s.c; ++c)
            for (int y = 0; y < s.h; ++y)
                memcpy(pad_scratch + ((size_t)c * s.h + y) * WP, in + ((size_t)c * s.h + y) * s.w,
                       sizeof(float) * s.w);
        conv_rows_fast(pad_scratch, s.c, s.h, WP, wt, bias, nf, k, OH, OW, out);
        return;
    }
    for (int f = 0; f < nf; ++f)
        for (int y = 0; y < OH; ++y)
            for (int x = 0; x < OW; ++x) {
                float a = bias[f];
                for (int c = 0; c < s.c; ++c)
                    for (int ky = 0; ky < k; ++ky)
                        for (int kx = 0; kx < k; ++kx)
                            a += wt[(((size_t)f * s.c + c) * k + ky) * k + kx] *
                                 in[((size_t)c * s.h + y + ky) * s.w + x + kx];
                out[((size_t)f * OH + y) * OW + x] = a;
            }
}

static void pool_forward(const float* in, shape3 s, int k, int st, float* out) {
    const int PH = pooled_size(s.h, k, st), PW = pooled_size(s.w, k, st);
    for (int c = 0; c < s.c; ++c)
        for (int ph = 0; ph < PH; ++ph)
            for (int pw = 0; pw < PW; ++pw) {
                const int h0 = ph * st, w0 = pw * st;
                const int h1 = h0 + k < s.h ? h0 + k : s.h;
                const int w1 = w0 + k < s.w ? w0 + k : s.w;
                float m = -FLT_MAX;
                for (int y = h0; y < h1; ++y)
                    for (int x = w0; x < w1; ++x) {
                        const float v = in[((size_t)c * s.h + y) * s.w + x];
                        if (v > m) m = v;
                    }
                out[((size_t)c * PH + ph) * PW + pw] = m;
            }
}

static void act_forward(float* x, shape3 s, int kind, const float* slope) {
    const int plane = s.h * s.w;
    for (int c = 0; c < s.c; ++c)
        for (int i = 0; i < plane; ++i) {
            float* p = x + (size_t)c * plane + i;
            const float v = *p;
            const float pos = v > 0.f ? v : 0.f, neg = v < 0.f ? v : 0.f;
            if (kind == ORC_ACT_RELU) *p = pos;
            else if (kind == ORC_ACT_ELU) *p = pos + 1.0f * (expf(neg) - 1.0f);
            else *p = pos + slope[c] * neg;
        }
}

CLONES static void fc_forward(const float* in, int n_in, const float* wt, const float* bias,
                              int n_out, float* out) {
    for (int o = 0; o < n_out; ++o) {
        /* 16-lane blocked dot product (a BLAS sgemv would block similarly); fixed order */
        float lane[16];
        for (int l = 0; l < 16; ++l) lane[l] = 0.f;
        const float* wr = wt + (size_t)o * n_in;
        int i = 0;
        for (; i + 16 <= n_in; i += 16)
            for (int l = 0; l < 16; ++l) lane[l] += wr[i + l] * in[i + l];
        for (int l = 0; i < n_in; ++i, ++l) lane[l] += wr[i] * in[i];
        float s = 0.f;
        for (int l = 0; l < 16; ++l) s += lane[l];
        out[o] = s + bias[o];
    }
}

static size_t max_activation(const orc_layer* layers, int n_layers, shape3 in, shape3* final_shape) {
    size_t mx = (size_t)in.c * in.h * in.w;
    shape3 cur = in;
    for (int i = 0; i < n_layers; ++i) {
        shape3 nx;
        if (layer_out_shape(&layers[i], cur, &nx)) return 0;
        size_t n = (size_t)nx.c * nx.h * nx.w;
        if (n > mx) mx = n;
        cur = nx;
    }
    *final_shape = cur;
    return mx;
}

/* floats conv_forward's zero-padded input copy needs: the largest c x h x (ROWP + k) over the Convolution layers */
static size_t pad_floats(const orc_layer* layers, int n_layers, shape3 in) {
    size_t need = 1;
    shape3 cur = in;
    for (int i = 0; i < n_layers; ++i) {
        shape3 nx;
        if (layer_out_shape(&layers[i], cur, &nx)) return 0;
        if (layers[i].kind == ORC_CONV) {
            const size_t n = (size_t)cur.c * cur.h * (ROWP + layers[i].b);
            if (n > need) need = n;
        }
        cur = nx;
    }
    return need;
}

/* runs one patch; returns final count; if stage >= 0 stops after that layer */
static int forward_one(const orc_layer* layers, int n_layers, const float* const* blobs,
                       const int* sel, int n_sel, int h, int w, const float* patch, int patch_c,
                       float* a, float* b, float* pad, int stage, float** result) {
    shape3 cur = {n_sel, h, w};
    /* HWC -> CHW gather with channel selection */
    for (int c = 0; c < n_sel; ++c) {
        const int sc = sel[c];
        for (int i = 0; i < h * w; ++i) a[(size_t)c * h * w + i] = patch[(size_t)i * patch_c + sc];
    }
    float *src = a, *dst = b;
    for (int li = 0; li < n_layers; ++li) {
        const orc_layer* L = &layers[li];
        shape3 nx;
        if (layer_out_shape(L, cur, &nx)) return -1;
        if (L->kind == ORC_CONV) {
            conv_forward(src, cur, blobs[L->blob0], blobs[L->blob1], L->a, L->b, dst, pad);
            float* t = src; src = dst; dst = t;
        } else if (L->kind == ORC_POOL) {
            pool_forward(src, cur, L->a, L->b, dst);
            float* t = src; src = dst; dst = t;
        } else if (L->kind == ORC_ACT) {
            act_forward(src, cur, L->a, L->blob0 >= 0 ? blobs[L->blob0] : 0);
        } else if (L->kind == ORC_FC) {
            fc_forward(src, cur.c * cur.h * cur.w, blobs[L->blob0], blobs[L->blob1], L->a, dst);
            float* t = src; src = dst; dst = t;
        } else return -1;
        cur = nx;
        if (li == stage) break;
    }
    *result = src;
    return cur.c * cur.h * cur.w;
}

int orc_forward(const orc_layer* layers, int n_layers, const float* const* blobs, const int* sel,
                int n_sel, int h, int w, const float* patches_hwc, int64_t n, int patch_c,
                float* out, int threads) {
    shape3 in = {n_sel, h, w}, fin;
    for (int c = 0; c < n_sel; ++c)
        if (sel[c] < 0 || sel[c] >= patch_c) return -1;
    const size_t mx = max_activation(layers, n_layers, in, &fin);
    if (!mx) return -1;
    const int D = fin.c * fin.h * fin.w;
    const size_t pad_n = pad_floats(layers, n_layers, in);
    int err = 0;
#ifdef _OPENMP
    const int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel num_threads(nt)
#endif
    {
        float* a = (float*)malloc(sizeof(float) * mx);
        float* b = (float*)malloc(sizeof(float) * mx);
        float* pad = (float*)malloc(sizeof(float) * pad_n);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 4)
#endif
        for (int64_t i = 0; i < n; ++i) {
            float* res = 0;
            const int cnt = forward_one(layers, n_layers, blobs, sel, n_sel, h, w,
                                        patches_hwc + (size_t)i * h * w * patch_c, patch_c, a, b,
                                        pad, -1, &res);
            if (cnt != D) { err = 1; continue; }
            memcpy(out + (size_t)i * D, res, sizeof(float) * D);
        }
        free(a); free(b); free(pad);
    }
    (void)threads;
    return err ? -1 : D;
}

int orc_forward_stage(const orc_layer* layers, int n_layers, const float* const* blobs,
                      const int* sel, int n_sel, int h, int w, const float* patch_hwc,
                      int patch_c, int stage, float* buf, int cap) {
    shape3 in = {n_sel, h, w}, fin;
    const size_t mx = max_activation(layers, n_layers, in, &fin);
    if (!mx) return -1;
    float* a = (float*)malloc(sizeof(float) * mx);
    float* b = (float*)malloc(sizeof(float) * mx);
    float* pad = (float*)malloc(sizeof(float) * pad_floats(layers, n_layers, in));
    float* res = 0;
    int cnt = forward_one(layers, n_layers, blobs, sel, n_sel, h, w, patch_hwc, patch_c, a, b, pad,
                          stage, &res);
    if (cnt > cap) cnt = -1;
    if (cnt > 0) memcpy(buf, res, sizeof(float) * cnt);
    free(a); free(b); free(pad);
    return cnt;
}

/* ---------------------------------------------------------------------------------------------
 * The same forward in the FORM Caffe-CPU runs it (the "reference-faithful" timing baseline of SURVEY 8d): per sample
 * im2col + SGEMM for the convolutions (caffe/layers/base_conv_layer.cpp: conv_im2col_cpu + caffe_cpu_gemm), one
 * batched SGEMM per InnerProduct (inner_product_layer.cpp: gemm(NoTrans, Trans)), bias added by a rank-1 update as
 * Caffe's bias multiplier does, single-threaded like a stock Caffe-CPU build with a serial BLAS.  No BLAS library is
 * installed here, so the SGEMM is a hand-blocked AVX2/FMA micro-kernel (4 x 16 register tile).  Same arithmetic as
 * orc_forward up to the summation order (tests hold the two within 1e-5 of the row max).
 * --------------------------------------------------------------------------------------------- */
#include <immintrin.h>

/* C[M][N] = A[M][K] * B[K][N] (row-major, C overwritten) */
static void sgemm_nn(int M, int N, int K, const float* A, const float* B, float* C) {
    for (int i0 = 0; i0 < M; i0 += 4) {
        const int mi = M - i0 < 4 ? M - i0 : 4;
        int j0 = 0;
        for (; j0 + 16 <= N; j0 += 16) {
            __m256 acc[4][2];
            for (int i = 0; i < 4; ++i) { acc[i][0] = _mm256_setzero_ps(); acc[i][1] = _mm256_setzero_ps(); }
            for (int k = 0; k < K; ++k) {
                const __m256 b0 = _mm256_loadu_ps(B + (size_t)k * N + j0), b1 = _mm256_loadu_ps(B + (size_t)k * N + j0 + 8);
                for (int i = 0; i < mi; ++i) {
                    const __m256 a = _mm256_broadcast_ss(A + (size_t)(i0 + i) * K + k);
                    acc[i][0] = _mm256_fmadd_ps(a, b0, acc[i][0]);
                    acc[i][1] = _mm256_fmadd_ps(a, b1, acc[i][1]);
                }
            }
            for (int i = 0; i < mi; ++i) {
                _mm256_storeu_ps(C + (size_t)(i0 + i) * N + j0, acc[i][0]);
                _mm256_storeu_ps(C + (size_t)(i0 + i) * N + j0 + 8, acc[i][1]);
            }
        }
        for (; j0 < N; ++j0)
            for (int i = 0; i < mi; ++i) {
                float s = 0.f;
                for (int k = 0; k < K; ++k) s += A[(size_t)(i0 + i) * K + k] * B[(size_t)k * N + j0];
                C[(size_t)(i0 + i) * N + j0] = s;
            }
    }
}

static float hsum8(__m256 v) {
    __m128 lo = _mm256_castps256_ps128(v), hi = _mm256_extractf128_ps(v, 1);
    lo = _mm_add_ps(lo, hi);
    lo = _mm_hadd_ps(lo, lo);
    lo = _mm_hadd_ps(lo, lo);
    return _mm_cvtss_f32(lo);
}

/* C[M][N] = A[M][K] * B[N][K]^T (InnerProduct over a batch: A = bottom, B = weights) */
static void sgemm_nt(int M, int N, int K, const float* A, const float* B, float* C) {
    for (int i0 = 0; i0 < M; i0 += 4)
        for (int j0 = 0; j0 < N; j0 += 2) {
            const int mi = M - i0 < 4 ? M - i0 : 4, nj = N - j0 < 2 ? N - j0 : 2;
            __m256 acc[4][2];
            for (int i = 0; i < 4; ++i) { acc[i][0] = _mm256_setzero_ps(); acc[i][1] = _mm256_setzero_ps(); }
            int k = 0;
            for (; k + 8 <= K; k += 8) {
                const __m256 b0 = _mm256_loadu_ps(B + (size_t)j0 * K + k);
                const __m256 b1 = nj > 1 ? _mm256_loadu_ps(B + (size_t)(j0 + 1) * K + k) : _mm256_setzero_ps();
                for (int i = 0; i < mi; ++i) {
                    const __m256 a = _mm256_loadu_ps(A + (size_t)(i0 + i) * K + k);
                    acc[i][0] = _mm256_fmadd_ps(a, b0, acc[i][0]);
                    acc[i][1] = _mm256_fmadd_ps(a, b1, acc[i][1]);
                }
            }
            for (int i = 0; i < mi; ++i)
                for (int j = 0; j < nj; ++j) {
                    float s = hsum8(acc[i][j]);
                    for (int kk = k; kk < K; ++kk) s += A[(size_t)(i0 + i) * K + kk] * B[(size_t)(j0 + j) * K + kk];
                    C[(size_t)(i0 + i) * N + j0 + j] = s;
                }
        }
}

/* caffe/util/im2col.cpp im2col_cpu for stride 1, pad 0: col[(c*k + ky)*k + kx][y*OW + x] = in[c][y+ky][x+kx] */
static void im2col(const float* in, shape3 s, int k, float* col) {
    const int OH = s.h - k + 1, OW = s.w - k + 1;
    for (int c = 0; c < s.c; ++c)
        for (int ky = 0; ky < k; ++ky)
            for (int kx = 0; kx < k; ++kx) {
                float* dst = col + ((size_t)(c * k + ky) * k + kx) * OH * OW;
                for (int y = 0; y < OH; ++y)
                    memcpy(dst + (size_t)y * OW, in + ((size_t)c * s.h + y + ky) * s.w + kx, sizeof(float) * OW);
            }
}

/* threads > 1: the samples of a batch are spread over OpenMP threads in every layer (own im2col buffer per thread, the
 * InnerProduct GEMM split by rows) - what a threaded BLAS under Caffe amounts to; results do not depend on the count */
int orc_forward_im2col_mt(const orc_layer* layers, int n_layers, const float* const* blobs, const int* sel, int n_sel, int h,
                          int w, const float* patches_hwc, int64_t n, int patch_c, int batch, float* out, int threads) {
#ifdef _OPENMP
    const int nt = threads > 0 ? threads : omp_get_max_threads();
#else
    const int nt = 1;
    (void)threads;
#endif
    shape3 in = {n_sel, h, w}, fin;
    for (int c = 0; c < n_sel; ++c)
        if (sel[c] < 0 || sel[c] >= patch_c) return -1;
    const size_t mx = max_activation(layers, n_layers, in, &fin);
    if (!mx || batch < 1) return -1;
    const int D = fin.c * fin.h * fin.w;
    /* largest im2col buffer over the conv layers */
    size_t col_n = 1;
    {
        shape3 cur = in;
        for (int i = 0; i < n_layers; ++i) {
            shape3 nx;
            layer_out_shape(&layers[i], cur, &nx);
            if (layers[i].kind == ORC_CONV) {
                const size_t need = (size_t)cur.c * layers[i].b * layers[i].b * nx.h * nx.w;
                if (need > col_n) col_n = need;
            }
            cur = nx;
        }
    }
    float* a = (float*)malloc(sizeof(float) * mx * batch);
    float* b = (float*)malloc(sizeof(float) * mx * batch);
    float* col = (float*)malloc(sizeof(float) * col_n * (size_t)nt);
    if (!a || !b || !col) { free(a); free(b); free(col); return -1; }
    for (int64_t first = 0; first < n; first += batch) {
        const int m = (int)(n - first < batch ? n - first : batch);
        shape3 cur = in;
        size_t cnt = (size_t)cur.c * cur.h * cur.w;
        /* the staging copy of src/networkevaluator.cpp:33-42: HWC -> CHW (+ channel select) for the whole batch */
#pragma omp parallel for num_threads(nt) schedule(static) if (nt > 1)
        for (int s = 0; s < m; ++s) {
            const float* patch = patches_hwc + (size_t)(first + s) * h * w * patch_c;
            for (int c = 0; c < n_sel; ++c)
                for (int i = 0; i < h * w; ++i) a[(size_t)s * cnt + (size_t)c * h * w + i] = patch[(size_t)i * patch_c + sel[c]];
        }
        float *src = a, *dst = b;
        for (int li = 0; li < n_layers; ++li) {
            const orc_layer* L = &layers[li];
            shape3 nx;
            if (layer_out_shape(L, cur, &nx)) { free(a); free(b); free(col); return -1; }
            const size_t ncnt = (size_t)nx.c * nx.h * nx.w;
            if (L->kind == ORC_CONV) {
                const int K = cur.c * L->b * L->b, N = nx.h * nx.w;
#pragma omp parallel for num_threads(nt) schedule(dynamic, 1) if (nt > 1)
                for (int s = 0; s < m; ++s) {
#ifdef _OPENMP
                    float* mycol = col + col_n * (size_t)omp_get_thread_num();
#else
                    float* mycol = col;
#endif
                    im2col(src + (size_t)s * cnt, cur, L->b, mycol);
                    float* o = dst + (size_t)s * ncnt;
                    sgemm_nn(L->a, N, K, blobs[L->blob0], mycol, o);
                    for (int f = 0; f < L->a; ++f) {
                        const float bf = blobs[L->blob1][f];
                        for (int i = 0; i < N; ++i) o[(size_t)f * N + i] += bf;
                    }
                }
                float* t = src; src = dst; dst = t;
            } else if (L->kind == ORC_POOL) {
#pragma omp parallel for num_threads(nt) schedule(static) if (nt > 1)
                for (int s = 0; s < m; ++s) pool_forward(src + (size_t)s * cnt, cur, L->a, L->b, dst + (size_t)s * ncnt);
                float* t = src; src = dst; dst = t;
            } else if (L->kind == ORC_ACT) {
#pragma omp parallel for num_threads(nt) schedule(static) if (nt > 1)
                for (int s = 0; s < m; ++s) act_forward(src + (size_t)s * cnt, cur, L->a, L->blob0 >= 0 ? blobs[L->blob0] : 0);
            } else if (L->kind == ORC_FC) {
                {
                    const int rows_per = (m + nt - 1) / nt;       /* row blocks: a row of the product depends on its own row of A only */
#pragma omp parallel for num_threads(nt) schedule(static) if (nt > 1)
                    for (int blk = 0; blk < nt; ++blk) {
                        const int r0 = blk * rows_per, r1 = r0 + rows_per < m ? r0 + rows_per : m;
                        if (r0 < r1) sgemm_nt(r1 - r0, L->a, (int)cnt, src + (size_t)r0 * cnt, blobs[L->blob0], dst + (size_t)r0 * L->a);
                    }
                }
                for (int s = 0; s < m; ++s)
                    for (int o = 0; o < L->a; ++o) dst[(size_t)s * L->a + o] += blobs[L->blob1][o];
                float* t = src; src = dst; dst = t;
            } else { free(a); free(b); free(col); return -1; }
            cur = nx;
            cnt = ncnt;
        }
        memcpy(out + (size_t)first * D, src, sizeof(float) * (size_t)m * D);
    }
    free(a); free(b); free(col);
    return D;
}

/* one thread: a stock Caffe-CPU build without a threaded BLAS */
int orc_forward_im2col(const orc_layer* layers, int n_layers, const float* const* blobs, const int* sel, int n_sel, int h,
                       int w, const float* patches_hwc, int64_t n, int patch_c, int batch, float* out) {
    return orc_forward_im2col_mt(layers, n_layers, blobs, sel, n_sel, h, w, patches_hwc, n, patch_c, batch, out, 1);
}
