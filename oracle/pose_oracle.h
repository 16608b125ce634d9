/*
 * pose_oracle.h - CPU restatement of the reference's evaluation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked, imported or
 * executed by the product (3dposeestimation_b200/, include/); only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * use it, and only as the checker / the timed CPU baseline.
 *
 * Parity status: the path's arithmetic lives in three third-party libraries
 * that are NOT under /root/reference and cannot be built here (BVLC Caffe
 * ~1.0 + PR #2603, OpenCV 3.1 features2d, Eigen 3.x).  The
 * reference ships no tests or golden vectors.  So, strictly, "parity
 * unpinned" against those libraries.  The restatement is instead pinned against
 * independent implementations that are available in the build container
 * (torch-CPU conv2d/max_pool2d(ceil)/elu/prelu/linear, cv2.BFMatcher, scipy
 * Rotation) by tests/golden/make_golden.py -> the .npz files under tests/golden/.
 */
#ifndef POSE_ORACLE_H
#define POSE_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_CONV = 1, ORC_POOL = 2, ORC_ACT = 3, ORC_FC = 4 };
enum { ORC_ACT_RELU = 1, ORC_ACT_ELU = 2, ORC_ACT_PRELU = 3 };

/* One TEST-phase layer.  CONV: a = num_output, b = kernel_size (stride 1, pad 0);
 * POOL: a = kernel, b = stride (MAX, Caffe ceil sizing); ACT: a = ORC_ACT_*;
 * FC: a = num_output.  blob0/blob1 index into the blob array (weight, bias or
 * PReLU slope), -1 when absent. */
typedef struct {
    int kind, a, b;
    int blob0, blob1;
} orc_layer;

/* Forward of n patches (HWC, patch_c channels) through the layer list.
 * sel/n_sel: channels of the input blob fed to the first layer (Slice/Concat/Silence).
 * out: n x D.  threads <= 0 -> all OpenMP threads.  Returns D, or -1 on a shape error. */
int orc_forward(const orc_layer* layers, int n_layers, const float* const* blobs,
                const int* sel, int n_sel, int h, int w,
                const float* patches_hwc, int64_t n, int patch_c, float* out, int threads);

/* the same forward in the form Caffe-CPU executes it: per-sample im2col + SGEMM convolutions, batched SGEMM
 * InnerProducts, `batch` samples per pass (the prototxt's input N), single-threaded - the "reference-faithful"
 * timing baseline of SURVEY 8(d); equal to orc_forward up to the summation order */
int orc_forward_im2col(const orc_layer* layers, int n_layers, const float* const* blobs, const int* sel, int n_sel, int h,
                       int w, const float* patches_hwc, int64_t n, int patch_c, int batch, float* out);
/* ... with the samples of a batch spread over `threads` OpenMP threads (0 = all): Caffe-CPU on a threaded BLAS; same result */
int orc_forward_im2col_mt(const orc_layer* layers, int n_layers, const float* const* blobs, const int* sel, int n_sel, int h,
                          int w, const float* patches_hwc, int64_t n, int patch_c, int batch, float* out, int threads);

/* also return intermediate activations of ONE patch (debug / layer-wise parity):
 * stage index s = output of layer s; buf must hold max activation size; returns count. */
int orc_forward_stage(const orc_layer* layers, int n_layers, const float* const* blobs,
                      const int* sel, int n_sel, int h, int w, const float* patch_hwc,
                      int patch_c, int stage, float* buf, int cap);

/* OpenCV BFMatcher(NORM_L2).knnMatch restated: dist = sqrt(sum (q-t)^2) in fp32
 * (8-lane blocked sum, no FMA contraction; see pose_oracle.c), k smallest
 * ascending, ties -> lower index.  idx/dist: nq x k (-1 / inf padded).
 * gap (may be NULL): nq floats, relative gap between the k-th and (k+1)-th
 * distance (inf when T <= k) for the acceptance rule of SURVEY 8c. */
void orc_knn_l2(const float* q, int64_t nq, const float* t, int64_t nt, int dim, int k,
                int32_t* idx, float* dist, float* gap, int threads);
float orc_l2_distance(const float* a, const float* b, int dim);

/* sign binarisation + Hamming kNN (new behaviour) */
void orc_binarize(const float* desc, int64_t n, int dim, uint32_t* bits);
void orc_knn_hamming(const uint32_t* q, int64_t nq, const uint32_t* t, int64_t nt, int words,
                     int k, int32_t* idx, int32_t* dist, int threads);

/* merge per-shard candidate lists (n_shards x nq x k, global indices) */
void orc_merge_topk_f32(const int32_t* idx, const float* dist, int n_shards, int64_t nq, int k,
                        int32_t* out_idx, float* out_dist);
void orc_merge_topk_i32(const int32_t* idx, const int32_t* dist, int n_shards, int64_t nq, int k,
                        int32_t* out_idx, int32_t* out_dist);

/* Eigen restatements */
float orc_angular_distance(const float* qa_xyzw, const float* qb_xyzw); /* radians; Eigen 3.2 form unless switched */
float orc_angular_distance_atan2(const float* qa_xyzw, const float* qb_xyzw); /* Eigen 3.3+ form */
void orc_set_angular_form(int form);    /* 0 (default): Eigen 3.2 acos form, 1: Eigen 3.3+ atan2 form (process-wide; tests only) */
float orc_rot22(const float* q_xyzw);                                   /* toRotationMatrix()(2,2) */

/* networkEvaluator::computeHistogram body, src/networkevaluator.cpp:269-355 */
void orc_histogram(const int32_t* knn_idx, int64_t nq, int k, const float* q_labels,
                   const float* t_labels, int nr_tmpl_poses, const int* rot_inv,
                   const float* bins, int n_bins, float* out /* n_bins+2 */,
                   float* per_query_err /* nq, may be NULL; -1 = no same-class */);
/* computeConfusionMatrix body, src/networkevaluator.cpp:199-245 */
void orc_confusion(const int32_t* knn_idx, int64_t nq, int k, const float* q_labels,
                   const float* t_labels, const int* class_to_local, int n_local,
                   int nr_test_poses, float* out /* n_local x n_local */);
/* networkSolver::computeKNN body, src/networksolver.cpp:474-504 */
void orc_knn_bootstrap(const int32_t* knn3_idx, int64_t nq, int nr_query_poses,
                       int nr_tmpl_poses, int32_t* out /* nq x 4 */, int32_t* out_len);
/* computeKNNAccuracy, src/networkevaluator.cpp:153-175 */
void orc_knn_accuracy(const int32_t* max_sim_tmpl0, const int32_t* bootstrap, int64_t nq,
                      int nr_query_poses, float* intra_pct, float* inter_pct);

/* patch extraction (oracle_patch.c): depth2normals src/helper.cpp:53-103, samplePatchWithScale
 * src/datasetmanager.cpp:299-346 with cv::resize restated */
void orc_depth2normals(const float* dep, int rows, int cols, float fx, float fy, float ox, float oy, float* nor);
int orc_patch_rect(int rows, int cols, int center_x, int center_y, float z, float fx, float fy, float m,
                   int* x, int* y, int* w, int* h);
int orc_sample_patch(const uint8_t* color, const float* depth, const float* normals, int rows, int cols,
                     int center_x, int center_y, float z, float fx, float fy, float m, float* out /* 64x64x7 */);

/* template renderer (oracle_render.c): view sphere src/sphere.cpp:27-35,134-226; Painter / RealWorldCamera / Model
 * pipeline maths src/painter.cpp:86-110,186-194,231-248, src/model.cpp:29-54.  PARITY UNPINNED (OpenGL unavailable). */
int orc_init_sphere(int depth, float* out_xyz, int cap);
void orc_view_pose(const float* vertex, float scale, float angle_deg, float* out_rt /* row-major 3x4 */);
/* label of a rendered template view (src/datasetmanager.cpp:484-491, 523-530): {index, quaternion xyzw of R, -(R^T t)} */
void orc_pose_label(const float* rt /* row-major 3x4 */, int index, float* out8);
/* patch centre of a real frame (src/datasetmanager.cpp:348-411): projected centroid truncated to int; returns centroid z */
float orc_project_centroid(const float* cam, const float* rt, const float* centroid3 /* or NULL */, int32_t* out_xy);
void orc_render_tex(const float* vertices, const float* colors, const float* uv /* [nv][2] or NULL */, int nv, const int32_t* faces, int nf,
                    const uint8_t* tex /* [tex_rows][tex_cols][3] BGR or NULL */, int tex_rows, int tex_cols, const float* cam, const float* rt,
                    int rows, int cols, float z_near, float z_far, uint8_t* out_color, float* out_depth);
void orc_render(const float* vertices, const float* colors, int nv, const int32_t* faces, int nf, const float* cam /* fx fy ox oy */,
                const float* rt, int rows, int cols, float z_near, float z_far, uint8_t* out_color /* BGR */, float* out_depth);

/* triplet miner (oracle_miner.c): computeMaxSimTmpl[Inplane] src/datasetmanager.cpp:687-783, buildBatch
 * src/networksolver.cpp:52-186, Caffe batch layout :32-40, randomColorFill src/bgfill.cpp:136-158 */
#define ORC_MAX_DRAWS 4096
uint64_t orc_miner_key(uint64_t seed, uint64_t linear_id);
uint64_t orc_miner_draw(uint64_t key, uint64_t j);
float orc_elevation(const float* label);
void orc_max_sim_tmpl(const float* tmpl_labels, const float* train_labels, int n_objects, int nr_template_poses,
                      int nr_training_poses, int inplane, int32_t* out /* [n_objects * nr_training_poses][2] */);
void orc_triplet_indices(const float* tmpl_labels, const int32_t* max_sim, const int32_t* knn_table, const int32_t* knn_len,
                         const int* rot_inv, const int64_t* neg_offsets, int n_objects, int nr_template_poses,
                         int nr_training_poses, int batch_size, int triplet_size, int iter, uint64_t seed,
                         int32_t* out_idx /* [batch_size / triplet_size][5] */, int32_t* out_flag);
void orc_gather_batch(const float* tmpl_pool, const float* train_pool, const float* neg_pool, const int32_t* idx,
                      int n_triplets, int channels, int64_t first_linear_id, uint64_t seed, int random_background,
                      float* out /* [n_triplets * 5][C][64][64] */);

int orc_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
