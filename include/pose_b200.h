/*
 * pose_b200.h - C ABI of the B200-native pose-descriptor evaluation path.
 *
 * This is the drop-in boundary for ONE hot path of mingliangfu/3DPoseEstimation:
 * descriptor forward -> brute-force kNN against the template DB -> pose-error
 * reduction.  Every entry point names the reference interface it replaces
 * (paths relative to the reference checkout).  Plain pointers and sizes only;
 * no C++ / torch / OpenCV / Caffe types cross this boundary.
 *
 * Conventions
 *   - every function returns PE_OK (0) or a PE_ERR_* code and never throws;
 *     pe_last_error() returns a thread-local, human-readable message.
 *   - "host" pointers are ordinary (ideally pinned) host memory owned by the
 *     caller; "_dev" variants take device pointers on the handle's device and
 *     are asynchronous on the handle's stream (call pe_*_sync to wait).
 *   - patches are S x 64 x 64 x C float32, interleaved HWC, exactly the
 *     memory of reference `Sample.data` (include/datatypes.h:25-27);
 *     labels are S x 8 float32 {class, qx,qy,qz,qw, tx,ty,tz}
 *     (src/datasetmanager.cpp:283-289, include/datatypes.h:34-48).
 *   - descriptor matrices are row-major float32, rows x D, the memory of the
 *     reference's `Mat descs(S, D, CV_32F)` (src/networkevaluator.cpp:23).
 *   - there is no CPU fallback: without a CUDA device every compute entry
 *     point fails with PE_ERR_CUDA.
 */
#ifndef POSE_B200_H
#define POSE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PE_API __attribute__((visibility("default")))

enum {
    PE_OK = 0,
    PE_ERR_INVALID = 1,     /* bad argument / shape mismatch                  */
    PE_ERR_CUDA = 2,        /* CUDA runtime error or no device                */
    PE_ERR_PARSE = 3,       /* prototxt could not be parsed / unsupported net */
    PE_ERR_IO = 4,          /* file could not be read / written               */
    PE_ERR_UNSUPPORTED = 5, /* valid request outside the implemented family   */
    PE_ERR_NCCL = 6         /* NCCL error                                     */
};

/* activation kinds reported by pe_net_layer_info (Caffe layer types ReLU/ELU/PReLU) */
enum { PE_ACT_NONE = 0, PE_ACT_RELU = 1, PE_ACT_ELU = 2, PE_ACT_PRELU = 3 };

typedef struct pe_net pe_net;   /* replaces caffe::Net<float> in TEST phase          */
typedef struct pe_db pe_db;     /* replaces cv::DescriptorMatcher("BruteForce")+add  */
typedef struct pe_comm pe_comm; /* NCCL communicator for the row-sharded DB (new)    */

/* ---------------------------------------------------------------- misc --- */
PE_API int pe_version(void);
PE_API const char* pe_last_error(void);
PE_API int pe_device_count(int* n);
/* number of kernels launched by this library in this process (bench "gpu_launches") */
PE_API int64_t pe_launch_count(void);

/* runtime switches (process-wide).  "knn_tensor" (default 1): L2 / Hamming kNN scan on the tcgen05 tensor path (descriptors of
 * up to 128 dimensions, k <= 16, where its shared-memory plan fits), 0 = fp32 SIMT scan.  "forward_tensor" (default 1):
 * conv / FC layers on the tcgen05 split-fp16 path (three fp16 products per term, fp32 accumulate: descriptors within
 * ~1e-6 of the row max of the fp32 reference), 0 = fp32 SIMT kernels.
 * "forward_fp8" (default 0 = off, OPT-IN): conv1 takes its two cross terms (hi.lo + lo.hi) from one e4m3 MMA instead of
 * two fp16 MMAs - conv1 1.45x faster, but the descriptor error grows from ~1e-6 to 1-3e-5 of the row max on the 16-d
 * nets and up to 1.5e-4 where the last layers cancel (manifold.prototxt, D = 3), and inputs must satisfy |x| <= 28.
 * 1 = qualified per set of weights and re-qualified every 32nd batch: a batch of >= 64 patches is forwarded both ways
 * (<= 256 patches) and e4m3 is kept only while it stays within 3e-5 of the all-fp16 split (pe_net_fp8_state reports the
 * outcome; an out-of-range batch withdraws the qualification, host entry points replay it on fp16); 2 = always
 * (out-of-range input: PE_ERR_INVALID).
 * "conv2_tensor" (default 1): 0 pairs the tensor conv1 with the fp32 SIMT conv2 (conv1 then writes fp32 pool1 instead
 * of the packed fp16 image; a test switch).  "knn_certify" (default 1): the L2 matcher proves per query that no
 * unexamined template can beat the reported k-th neighbour (error bound of the split-precision scan) and re-scans the
 * queries it cannot prove with the direct fp32 (q-t)^2 expression; 0 skips the proof (round-1 behaviour).
 * Every path is CUDA; none is a CPU fallback. */
PE_API int pe_set_option(const char* name, int value);
/* conv1 cross terms of this net's current weights: 0 not decided yet (all-fp16 until then), 1 e4m3, 2 fp16 */
PE_API int pe_net_fp8_state(const pe_net* net, int* state);
PE_API int pe_get_option(const char* name, int* value);

/* pinned host memory + raw device memory helpers (harness plumbing) */
PE_API int pe_host_alloc(void** p, size_t bytes);
PE_API int pe_host_free(void* p);
PE_API int pe_dev_alloc(int device, void** p, size_t bytes);
PE_API int pe_dev_free(int device, void* p);
PE_API int pe_memcpy_h2d(int device, void* dst, const void* src, size_t bytes);
PE_API int pe_memcpy_d2h(int device, void* dst, const void* src, size_t bytes);
PE_API int pe_dev_memset(int device, void* dst, int value, size_t bytes);

/* CUDA-event timers on a handle's stream (so kernel time is measured on the
 * stream the kernels are launched on) */
typedef struct pe_timer pe_timer;
PE_API int pe_timer_create(int device, pe_timer** t);
PE_API int pe_timer_start_net(pe_timer* t, pe_net* net);
PE_API int pe_timer_stop_net(pe_timer* t, pe_net* net);
PE_API int pe_timer_start_db(pe_timer* t, pe_db* db);
PE_API int pe_timer_stop_db(pe_timer* t, pe_db* db);
PE_API int pe_timer_elapsed_ms(pe_timer* t, float* ms); /* synchronises on stop */
PE_API void pe_timer_destroy(pe_timer* t);

/* per-kernel device time (bench.py roofline): while enabled, every forward / kNN call brackets its
 * kernels with CUDA events on the launching stream.  *_read synchronises, returns the accumulated
 * milliseconds since the last read and the number of bracketed calls, and resets.
 * net stages: [0] conv1+pool+act, [1] conv2+pool+act, [2] all InnerProducts.
 * db stages:  [0] scan (L2 or Hamming), [1] select / refine (+ merge). */
PE_API int pe_net_profile(pe_net* net, int enable);
PE_API int pe_net_profile_read(pe_net* net, double ms[3], int64_t* calls);
PE_API int pe_db_profile(pe_db* db, int enable);
PE_API int pe_db_profile_read(pe_db* db, double ms[2], int64_t* calls);

/* ----------------------------------------------------------------- net --- */
/* Replaces `caffe::Net<float> testCNN(network_path + net_name + ".prototxt", caffe::TEST)`
 * (src/networksolver.cpp:286).  Parses the TEST-phase subset of a Caffe prototxt
 * (layers Convolution, Pooling, ReLU, ELU, PReLU, InnerProduct, Slice, Concat,
 * Silence; `include {phase: TRAIN}` layers are skipped) and accepts the
 * conv-pool-act-conv-pool-act-FC... family of network/manifold*.prototxt. */
PE_API int pe_net_create_from_prototxt(const char* path, int device, pe_net** out);
PE_API int pe_net_create_from_string(const char* prototxt_text, int device, pe_net** out);
/* Built-in table of the seven reference variants by file stem, e.g.
 * "manifold_rgbdnor_16" (network/manifold_rgbdnor_16.prototxt). */
PE_API int pe_net_create_named(const char* variant, int device, pe_net** out);
PE_API void pe_net_destroy(pe_net* net);

/* input blob N,C,H,W and output dim D: what computeDescriptors reads from
 * CNN.input_blobs()[0] / output_blobs()[0] (src/networkevaluator.cpp:11-19) */
PE_API int pe_net_info(const pe_net* net, int* batch, int* channels, int* height, int* width,
                       int* desc_dim);
/* channel-select list fed to conv1 (Slice/Concat/Silence, network/manifold_rgb_16.prototxt:6-33) */
PE_API int pe_net_selected_channels(const pe_net* net, int* sel, int cap, int* n_sel);
/* human-readable, one TEST-phase layer per line: "conv conv1 16 8", "pool pool1 2 2",
 * "act relu1 PReLU", "fc ip1 256" ... */
PE_API int pe_net_describe(const pe_net* net, char* buf, size_t cap);

/* learnable blobs in Caffe order (per layer: weight, bias; PReLU: slope).
 * Replaces CopyTrainedLayersFrom / ShareTrainedLayersWith (src/networksolver.cpp:307,377). */
PE_API int pe_net_num_blobs(const pe_net* net, int* n);
PE_API int pe_net_blob_info(const pe_net* net, int blob, char* name, size_t name_cap,
                            int64_t* count);
PE_API int pe_net_set_blob(pe_net* net, int blob, const float* host, int64_t count);
PE_API int pe_net_get_blob(const pe_net* net, int blob, float* host, int64_t count);
/* mode 0: Caffe fillers of the prototxts (xavier weights, constant-0 bias, PReLU 0.25);
 * mode 1: parity fill (xavier weights, bias U(-0.1,0.1), PReLU slopes U(-0.25,0.5)). */
PE_API int pe_net_fill(pe_net* net, uint64_t seed, int mode);
/* flat little-endian weight file: magic "PEW1", n_blobs, then per blob {int64 count, floats} */
PE_API int pe_net_save_weights(const pe_net* net, const char* path);
PE_API int pe_net_load_weights(pe_net* net, const char* path);

/* networkEvaluator::computeDescriptors (src/networkevaluator.cpp:9-54).
 * patches: n x H x W x patch_channels (HWC); out: n x D.  patch_channels must
 * equal the net's input channels (the contract noted at :37-40). */
PE_API int pe_compute_descriptors(pe_net* net, const float* patches_hwc, int64_t n,
                                  int patch_channels, float* out_desc);
PE_API int pe_compute_descriptors_dev(pe_net* net, const float* d_patches_hwc, int64_t n,
                                      int patch_channels, float* d_out_desc);
PE_API int pe_net_sync(pe_net* net);

/* deterministic synthetic patches generated on the device (bench / full-size
 * property tests).  Sample i belongs to object i / poses_per_object. */
PE_API int pe_synth_patches_dev(pe_net* net, uint64_t seed, int64_t first_index, int64_t n,
                                int channels, int poses_per_object, float* d_patches_hwc);
/* labels {class, unit quaternion, 0,0,0} for rows [first, first+n) (host output) */
PE_API int pe_synth_labels(uint64_t seed, int64_t first_index, int64_t n, int poses_per_object,
                           float* labels);

/* sign-binarisation of float descriptors to packed bits (bit j of word w is
 * desc[32*w+j] > 0).  New behaviour: the reference never thresholds bin_ip2
 * (SURVEY 0); defined in SURVEY 7. */
PE_API int pe_binarize(pe_net* net, const float* desc, int64_t n, int dim, uint32_t* bits);
PE_API int pe_binarize_dev(pe_net* net, const float* d_desc, int64_t n, int dim, uint32_t* d_bits);

/* ------------------------------------------------------------------ db --- */
/* Replaces `matcher = DescriptorMatcher::create("BruteForce"); matcher->add(DBfeats)`
 * (src/networkevaluator.cpp:193-196, 263-266; src/networksolver.cpp:466-467).
 * desc: t_local x dim rows of this shard; row_offset: global index of row 0
 * (0 when unsharded).  labels_all: t_total x 8 labels of ALL templates (may be
 * NULL when only pe_knn is used).  poses_per_object = template_set[0].size(). */
PE_API int pe_db_create(const float* desc, int64_t t_local, int dim, int64_t row_offset,
                        const float* labels_all, int64_t t_total, int poses_per_object,
                        int device, pe_db** out);
PE_API int pe_db_create_dev(const float* d_desc, int64_t t_local, int dim, int64_t row_offset,
                            const float* labels_all_host, int64_t t_total, int poses_per_object,
                            int device, pe_db** out);
/* Hamming DB over packed codes (words_per_code u32 each). */
PE_API int pe_db_create_hamming(const uint32_t* codes, int64_t t_local, int words_per_code,
                                int64_t row_offset, const float* labels_all, int64_t t_total,
                                int poses_per_object, int device, pe_db** out);
PE_API void pe_db_destroy(pe_db* db);
PE_API int pe_db_info(const pe_db* db, int64_t* t_local, int64_t* t_total, int* dim,
                      int* is_hamming);
PE_API int pe_db_sync(pe_db* db);
/* run the DB's kernels on the net's stream, so forward -> kNN needs no host synchronisation
 * (the net must outlive the DB or be re-shared) */
PE_API int pe_db_share_stream(pe_db* db, pe_net* net);

/* matcher->knnMatch(queries, matches, k) (src/networkevaluator.cpp:197,267;
 * src/networksolver.cpp:472): L2 distance sqrt(sum (q-t)^2), k smallest
 * ascending, ties -> lower index.  out_idx / out_dist are nq x k (global
 * template indices; -1 / +inf pad when t_total < k). */
PE_API int pe_knn(pe_db* db, const float* queries, int64_t nq, int k, int32_t* out_idx,
                  float* out_dist);
PE_API int pe_knn_dev(pe_db* db, const float* d_queries, int64_t nq, int k, int32_t* d_out_idx,
                      float* d_out_dist);
/* exactness certificate of the L2 matcher (pe_set_option("knn_certify"), default on): number of queries of the LAST
 * L2 kNN call on this DB whose neighbours the ranking pass could not prove and that were therefore re-scanned with the
 * direct fp32 (q-t)^2 expression (synchronises the DB's stream).  New: OpenCV's BFMatcher computes that expression for
 * every pair (reference src/networkevaluator.cpp:197); here it is the proof obligation of the tensor / SIMT scans. */
PE_API int pe_db_knn_flagged(pe_db* db, int64_t* n);
/* Hamming variant; distances are bit counts. */
PE_API int pe_knn_hamming(pe_db* db, const uint32_t* qcodes, int64_t nq, int k, int32_t* out_idx,
                          int32_t* out_dist);
PE_API int pe_knn_hamming_dev(pe_db* db, const uint32_t* d_qcodes, int64_t nq, int k,
                              int32_t* d_out_idx, int32_t* d_out_dist);

/* body of networkEvaluator::computeHistogram after knnMatch
 * (src/networkevaluator.cpp:269-355).  knn_idx: nq x k global template indices,
 * q_labels: nq x 8.  rot_inv[class] as `input.rotInv` (config.ini:10).
 * out: n_bins + 2 floats = normalised bins, mean angle, median angle. */
PE_API int pe_histogram(pe_db* db, const int32_t* knn_idx, const float* q_labels, int64_t nq,
                        int k, const int* rot_inv, int n_rot, const float* bins, int n_bins,
                        float* out);
/* same with knn_idx and q_labels resident on the DB's device (out stays a host pointer) */
PE_API int pe_histogram_dev(pe_db* db, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq,
                            int k, const int* rot_inv, int n_rot, const float* bins, int n_bins,
                            float* out);
/* pe_histogram_dev without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, kernels and results as pe_histogram_dev; the scratch (bin counts, error sum, median slot, rotInv
 * table, bin edges, per-query errors) lives in a fixed file-scope device array of 266,448 bytes (260.2 KiB) that every
 * GPU which loads the library holds until the process exits, so the sizes are capped:
 *     nq <= 65,536 queries per call and n_rot <= 1,024 classes - larger values return PE_ERR_INVALID (use
 *     pe_histogram_dev, which allocates).
 * The call makes no cudaMalloc / cudaFree / allocating library call; the first call on a device resolves the array's
 * device address (cudaGetSymbolAddress), so make one call outside a capture first.  All calls on a device share the
 * array, which keeps the previous call's contents (what must start at zero is cleared on the DB's stream at the start of
 * each call): calls that overlap on the GPU - two DBs or two streams of one device - race on it and must be serialised
 * by the caller.  Synchronises the DB's stream before it returns, like pe_histogram_dev. */
PE_API int pe_histogram_dev_noalloc(pe_db* db, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq,
                                    int k, const int* rot_inv, int n_rot, const float* bins, int n_bins,
                                    float* out);
/* body of computeConfusionMatrix after knnMatch (src/networkevaluator.cpp:199-245).
 * class_to_local[c] = local_index[models[c]] (:215-216); out: n_local x n_local,
 * row-normalised by nr_test_poses. */
PE_API int pe_confusion(pe_db* db, const int32_t* knn_idx, const float* q_labels, int64_t nq,
                        int k, const int* class_to_local, int n_classes, int n_local,
                        int nr_test_poses, float* out);
/* body of networkSolver::computeKNN after knnMatch with k = 3
 * (src/networksolver.cpp:474-504).  out: nq x 4 ints {nnObj, nnPose, otherObj,
 * otherPose}, out_len[q] in {2,4} = maxSimKNNTmpl[obj][pose].size(). */
PE_API int pe_knn_bootstrap(pe_db* db, const int32_t* knn3_idx, int64_t nq, int nr_query_poses,
                            int32_t* out, int32_t* out_len);
/* networkEvaluator::computeKNNAccuracy (src/networkevaluator.cpp:153-175):
 * max_sim_tmpl0[q] = maxSimTmpl[obj][pose][0]; returns the two percentages. */
PE_API int pe_knn_accuracy(pe_db* db, const int32_t* max_sim_tmpl0, const int32_t* bootstrap,
                           int64_t nq, int nr_query_poses, float* intra_pct, float* inter_pct);

/* ------------------------------------------------ patch extraction --- */
/* The producer side of computeDescriptors on the device: frames -> 64x64x7 patches in the layout
 * pe_compute_descriptors[_dev] consumes.  `net` (may be null) only supplies the stream / device for the _dev variants,
 * so that extraction and forward queue up without a host round trip; with net == null `device` and the default
 * stream are used.
 *
 * depth2normals (reference src/helper.cpp:53-103, include/helper.h:73-75): depth [n_frames][rows][cols] f32 metres ->
 * normals [n_frames][rows][cols][3] f32, zero where the reference leaves them zero. */
PE_API int pe_depth2normals(int device, const float* depth, int n_frames, int rows, int cols, float fx, float fy,
                            float ox, float oy, float* normals);
PE_API int pe_depth2normals_dev(pe_net* net, int device, const float* d_depth, int n_frames, int rows, int cols,
                                float fx, float fy, float ox, float oy, float* d_normals);
/* datasetManager::samplePatchWithScale (reference src/datasetmanager.cpp:299-346) for n patches: colour
 * [n_frames][rows][cols][3] u8 BGR, depth as above, normals as above or null (then computed on the fly for the pixels
 * the patch touches), frame_of[n] or null (patch i from frame i), centers_xy[n][2] (projected centroid, pixels),
 * z[n] (centroid depth, metres), fx / fy of the camera, m = metric cut size (0.2 in the reference) ->
 * out_patches [n][64][64][7] f32 {B,G,R,depth,nx,ny,nz}; status[n] (may be null): 0 ok, 1 = the cut rectangle does
 * not fit the frame (the reference asserts there; here the patch is zero-filled). */
PE_API int pe_sample_patches(int device, const uint8_t* color, const float* depth, const float* normals, int n_frames,
                             int rows, int cols, const int32_t* frame_of, const int32_t* centers_xy, const float* z,
                             int64_t n, float fx, float fy, float m, float* out_patches, int32_t* status);
PE_API int pe_sample_patches_dev(pe_net* net, int device, const uint8_t* d_color, const float* d_depth,
                                 const float* d_normals, int n_frames, int rows, int cols, const int32_t* d_frame_of,
                                 const int32_t* d_centers_xy, const float* d_z, int64_t n, float fx, float fy, float m,
                                 float* d_out_patches, int32_t* d_status);

/* ------------------------------------------------ template renderer --- */
/* SphereRenderer::createViews' pose list (reference src/sphere.cpp:27-35,69-94,134-226), host only: icosahedron
 * subdivided `sphere_depth` times in the reference's vertex order, lower hemisphere optionally skipped, for every
 * distance in scale3 = {first, step, last} (metres) and in-plane rotation in rot3 = {first, step, last} (degrees) one
 * row-major 3x4 [R|t].  out_poses may be null to query the count. */
PE_API int pe_sphere_views(int sphere_depth, const float* scale3, const float* rot3, int skip_lower_hemi,
                           float* out_poses, int64_t cap, int64_t* n_views);
/* Painter::paint with RealWorldCamera + Model (reference src/painter.cpp:86-110,186-194,231-248,
 * src/model.cpp:29-54) as a CUDA rasteriser, for n_views poses of one mesh: vertices [n_vertices][3], colors
 * [n_vertices][3] RGB in [0,1], faces [n_faces][3], cam4 = {fx, fy, ox, oy}, z_near / z_far (0.001 / 4.0 in the
 * reference) -> out_color [n_views][rows][cols][3] u8 BGR (black background), out_depth [n_views][rows][cols] f32
 * metres (0 = background): the frame layout pe_sample_patches reads.  OpenGL itself is unavailable on this box. */
/* (a triangle that crosses the near plane is clipped at Zc = z_near, not dropped) */
PE_API int pe_render_views(int device, const float* vertices, const float* colors, int n_vertices, const int32_t* faces,
                           int n_faces, const float* cam4, const float* poses, int n_views, int rows, int cols,
                           float z_near, float z_far, uint8_t* out_color, float* out_depth);
PE_API int pe_render_views_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, int n_vertices,
                               const int32_t* d_faces, int n_faces, const float* cam4, const float* d_poses, int n_views,
                               int rows, int cols, float z_near, float z_far, uint8_t* d_out_color, float* d_out_depth);

/* The same for a TEXTURED mesh (BigBIRD models; reference src/model.cpp:330-350, src/painter.cpp:140-160): uv [n_vertices][2]
 * per-vertex texture coordinates, tex [tex_rows][tex_cols][3] u8 BGR in the row order glTexImage2D receives it (t = 0 at
 * row 0; the reference flips, shrinks by 6 and brightens the PNG on the host before the upload - that preparation stays
 * with the caller).  Fixed-function semantics: fragment = vertex colour x texel (GL_MODULATE), GL_REPEAT, GL_LINEAR on
 * level 0.  uv and tex are both null for an untextured mesh (= pe_render_views).  Triangles that cross the near plane are
 * clipped at Zc = z_near, as GL clips them. */
PE_API int pe_render_views_tex(int device, const float* vertices, const float* colors, const float* uv, int n_vertices,
                               const int32_t* faces, int n_faces, const uint8_t* tex, int tex_rows, int tex_cols, const float* cam4,
                               const float* poses, int n_views, int rows, int cols, float z_near, float z_far, uint8_t* out_color,
                               float* out_depth);
PE_API int pe_render_views_tex_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, const float* d_uv,
                                   int n_vertices, const int32_t* d_faces, int n_faces, const uint8_t* d_tex, int tex_rows,
                                   int tex_cols, const float* cam4, const float* d_poses, int n_views, int rows, int cols,
                                   float z_near, float z_far, uint8_t* d_out_color, float* d_out_depth);
/* Model::loadModel (reference src/model.cpp:296-352, there through cv::viz::Mesh::load) as a minimal PLY reader on the
 * host: ascii / binary_little_endian, triangles only (the reference asserts that, :318), per-vertex colours (uchar / 255;
 * 127 grey when absent, :325) and texture coordinates (u v | s t | texture_u texture_v).  pe_mesh_data copies into
 * caller arrays sized by pe_mesh_info (uv only when has_uv; any pointer may be null). */
typedef struct pe_mesh pe_mesh;
PE_API int pe_mesh_load_ply(const char* path, pe_mesh** out);
PE_API int pe_mesh_info(const pe_mesh* mesh, int* n_vertices, int* n_faces, int* has_colors, int* has_uv, float* centroid3);
PE_API int pe_mesh_data(const pe_mesh* mesh, float* vertices, float* colors, float* uv, int32_t* faces);
PE_API void pe_mesh_destroy(pe_mesh* mesh);

/* The producer of a template set: datasetManager::createSynthSamplesWadim / createSynthSamplesPaul (reference
 * src/datasetmanager.cpp:440-496, 498-535) for n_views poses (row-major 3x4 [R|t], e.g. from pe_sphere_views, or camPos x
 * trans as the Paul variant builds them) of one mesh: every view is rendered, its patch is cut at the principal point
 * (cam4[2], cam4[3] truncated to int, as the reference passes them) with z = the pose's z translation (z_override[n_views]
 * replaces it when not null: the Paul variant's trans(2) after its BigBIRD pose adaptation) and metric size m (0.2), and
 * labelled {index, Quaternionf(pose.linear()).coeffs() = x y z w, pose.inverse().translation()} (:484-491, :523-530).
 * out_patches [n_views][64][64][7] (the layout pe_compute_descriptors reads), out_labels [n_views][8] (may be null),
 * status [n_views] as pe_sample_patches (may be null).  uv / tex as pe_render_views_tex (both null: untextured).
 * The _dev variant takes the mesh and writes the patches on the device (poses and labels stay host arrays: 20 floats per
 * view) on the net's stream, so mesh -> patches -> descriptors needs no host round trip; frames never leave the GPU. */
PE_API int pe_template_samples(int device, const float* vertices, const float* colors, const float* uv, int n_vertices,
                               const int32_t* faces, int n_faces, const uint8_t* tex, int tex_rows, int tex_cols,
                               const float* cam4, const float* poses, int64_t n_views, int index, const float* z_override,
                               int rows, int cols, float z_near, float z_far, float m, float* out_patches,
                               float* out_labels, int32_t* status);
PE_API int pe_template_samples_dev(pe_net* net, int device, const float* d_vertices, const float* d_colors, const float* d_uv,
                                   int n_vertices, const int32_t* d_faces, int n_faces, const uint8_t* d_tex, int tex_rows,
                                   int tex_cols, const float* cam4, const float* poses_host, int64_t n_views, int index,
                                   const float* z_override, int rows, int cols, float z_near, float z_far, float m,
                                   float* d_out_patches, float* out_labels_host, int32_t* d_status);
/* The producer of a test / training set from real frames: datasetManager::extractRealSamplesPaul / extractRealSamplesWadim
 * (reference src/datasetmanager.cpp:348-411): frame i (colour u8 BGR, metric depth) with ground-truth pose poses[i]
 * ([R|t]) -> the patch around the projected centroid (cam * centroid / z truncated to int; centroid = the pose's
 * translation, or pose * centroid3 when centroid3 is not null: the BigBIRD branch) at depth z = centroid z (Paul) or, with
 * z_from_depth != 0, the depth under that pixel (Wadim; PE_ERR_INVALID where the reference asserts), normals computed on
 * the fly, and the same label as pe_template_samples. */
PE_API int pe_scene_samples(int device, const uint8_t* color, const float* depth, int n_frames, int rows, int cols,
                            const float* cam4, const float* poses, const float* centroid3, int z_from_depth, int index,
                            float m, float* out_patches, float* out_labels, int32_t* status);
/* the labels alone (host arithmetic): poses [n][12] -> labels [n][8] */
PE_API int pe_pose_labels(const float* poses, int64_t n, int index, float* out_labels);

/* ------------------------------------------------ on-disk sample sets --- */
/* hdf5Handler::write / read (reference src/hdf5handler.cpp:11-132) without HDF5: one flat little-endian file holding
 * the same two arrays as the reference's datasets - "data" [n][C][H][W] f32 and "label" [n][label_dim] f32 - behind a
 * 32-byte header ("PESS0001", int64 n, int32 C, H, W, label_dim).  The calls take / return HWC patches (Sample.data);
 * the HWC <-> CHW transposes of the reference are inside.  tools/h5_to_pess.py converts a reference .h5 file,
 * tools/caffemodel_to_pew.py a trained .caffemodel into the weight file pe_net_load_weights reads. */
PE_API int pe_samples_write(const char* path, const float* data_hwc, const float* labels, int64_t n, int channels,
                            int height, int width, int label_dim);
PE_API int pe_samples_info(const char* path, int64_t* n, int* channels, int* height, int* width, int* label_dim);
PE_API int pe_samples_read(const char* path, int64_t first, int64_t count, float* out_data_hwc, float* out_labels);

/* ------------------------------------------------ triplet miner --- */
/* The training-side consumer of the k-NN: networkSolver's view of the dataset (template_set, training_set, rotInv,
 * inplane; reference include/networksolver.h:60-75) with both sample sets resident on the device.
 * Patches are HWC [n_objects * poses][64][64][channels] f32 in object-major order (template_set[object][pose]), labels
 * [n_objects * poses][8] = {class, quaternion xyzw, translation}; rot_inv[object] = rotInv[model_index[used_models[
 * object]]].  A null patch pointer leaves that pool allocated but unfilled (see pe_miner_pools_dev). */
typedef struct pe_miner pe_miner;
PE_API int pe_miner_create(int device, int channels, int n_objects, int nr_template_poses, int nr_training_poses,
                           const float* tmpl_patches_hwc, const float* tmpl_labels, const float* train_patches_hwc,
                           const float* train_labels, const int* rot_inv, int inplane, pe_miner** out);
PE_API void pe_miner_destroy(pe_miner* m);
/* device addresses of the two pools, for producers that write them in place (pe_render_views_dev ->
 * pe_sample_patches_dev) */
PE_API int pe_miner_pools_dev(pe_miner* m, float** d_tmpl_patches, float** d_train_patches);
PE_API int pe_miner_sync(pe_miner* m);
/* queue the miner's kernels on the net's stream (as pe_db_share_stream); the net must outlive the miner */
PE_API int pe_miner_share_stream(pe_miner* m, pe_net* net);
/* datasetManager::getHardNegatives (used at src/networksolver.cpp:60,95-100): patches of all objects back to back,
 * offsets[n_objects + 1] the prefix offsets per object; offsets == null removes them */
PE_API int pe_miner_set_hard_negatives(pe_miner* m, const float* patches_hwc, const int64_t* offsets);
/* datasetManager::computeMaxSimTmpl / computeMaxSimTmplInplane (src/datasetmanager.cpp:687-783):
 * out [n_objects * nr_training_poses][2] = maxSimTmpl[object][pose][0..1]; -1 where the reference would push an
 * uninitialised int.  out may be null (the table stays on the device for pe_miner_build_batch). */
PE_API int pe_miner_max_sim(pe_miner* m, int32_t* out);
/* networkSolver::computeKNN (src/networksolver.cpp:456-505) over the resident pools: forward of both sets, k = 3
 * matcher, maxSimKNNTmpl; switches `bootstrapping` on as trainNet does after it (:308-309,332-333).
 * out_table [nq][4] / out_len [nq] as pe_knn_bootstrap, both may be null. */
PE_API int pe_miner_compute_knn(pe_miner* m, pe_net* net, int32_t* out_table, int32_t* out_len);
/* the same table from the host (pe_knn_bootstrap's output); table == null switches bootstrapping off */
PE_API int pe_miner_set_knn(pe_miner* m, const int32_t* table, const int32_t* len);
/* networkSolver::buildBatch + the batch_caffe fill (src/networksolver.cpp:32-40,52-186).
 * out_batch: [batch_size / 5 * 5][channels][64][64] f32 (Caffe layout; anchor, puller, pusher0, pusher1, pusher2 per
 * triplet); out_idx: [batch_size / 5][5] = anchor as training linear index, the others as template linear indices, a
 * hard negative as ~(index into the negative pool).  Either may be null.  The reference draws from
 * std::random_device; here draw j of triplet linearId is a hash of (seed, linearId, j), so a call is reproducible and
 * the caller varies `seed` per epoch.  random_background: 0, or 1 = bgfill::randomColorFill on the anchor
 * (src/bgfill.cpp:136-158; 7-channel patches); shapes / fractals / real backgrounds return PE_ERR_INVALID.
 * A rejection loop the reference would never leave fails with PE_ERR_INVALID. */
PE_API int pe_miner_build_batch(pe_miner* m, int batch_size, int triplet_size, int iter, uint64_t seed,
                                int random_background, float* out_batch, int32_t* out_idx);
/* asynchronous on the miner's stream, outputs on the device; check pe_miner_flags afterwards */
PE_API int pe_miner_build_batch_dev(pe_miner* m, int batch_size, int triplet_size, int iter, uint64_t seed,
                                    int random_background, float* d_out_batch, int32_t* d_out_idx);
PE_API int pe_miner_flags(pe_miner* m, int n_triplets, int32_t* out_flags);

/* ----------------------------------------------------------- multi-GPU --- */
/* One process per GPU.  Rank 0 calls pe_comm_unique_id, the harness ships the
 * 128 bytes to all ranks, every rank calls pe_comm_create.  New: the reference
 * is single-process. */
PE_API int pe_comm_unique_id(char* id128);
PE_API int pe_comm_create(int rank, int world, const char* id128, int device, pe_comm** out);
PE_API void pe_comm_destroy(pe_comm* comm);
PE_API int pe_comm_info(const pe_comm* comm, int* rank, int* world);
/* *active = 1 when every rank mapped every peer's window (CUDA IPC over NVLink) AND pe_set_option("p2p", 1) selected the
 * peer-memory all-gather (one kernel: push to every peer's window, publish the epoch, wait, copy out) instead of the
 * NCCL call.  Default 0: measured on B200 (tests/multigpu_check.py, 262 KB per rank) the kernel takes 15 us at 2 GPUs and
 * 21 us at 8 against 12 and 23 us for ncclAllGather - a tie, so NCCL stays the default.  PE_NO_P2P=1 skips the window
 * set-up.  A peer-window all-gather that timed out waiting for a rank makes the sharded entry point that issued it (or the
 * pe_db_sync that completes it) return PE_ERR_NCCL; so does this call. */
PE_API int pe_comm_p2p(const pe_comm* comm, int* active);
/* all-gather n floats per rank (host buffers; descriptor exchange) */
PE_API int pe_comm_allgather_f32(pe_comm* comm, const float* send, int64_t n, float* recv);
PE_API int pe_comm_allreduce_sum_f64(pe_comm* comm, double* buf, int64_t n);
/* device-resident all-gather of n floats per rank, enqueued on the net's stream (descriptor
 * exchange between the forward and the sharded kNN without a host round trip); always NCCL - the
 * peer-window kernel serves the matcher's exchanges on the DB's stream only */
PE_API int pe_comm_allgather_f32_dev(pe_comm* comm, pe_net* net, const float* d_send, int64_t n, float* d_recv);
/* kNN against a row-sharded DB: every rank passes the same nq queries, scans
 * its shard, candidates are all-gathered over NVLink and merged identically on
 * every rank (ties -> lower global index). */
PE_API int pe_knn_sharded(pe_db* db, pe_comm* comm, const float* queries, int64_t nq, int k,
                          int32_t* out_idx, float* out_dist);
PE_API int pe_knn_sharded_dev(pe_db* db, pe_comm* comm, const float* d_queries, int64_t nq, int k,
                              int32_t* d_out_idx, float* d_out_dist);
/* Fused test pass: descriptors of n HOST patches and their k nearest templates in one call, i.e. what
 * networkSolver::computeKNN (reference src/networksolver.cpp:456-472) and the first half of
 * networkEvaluator::computeKNNAccuracy (src/networkevaluator.cpp:182-197) do with CNN.Forward + knnMatch.
 * The batch is staged in 512-patch pieces so that the host->device copy overlaps the forward and the scan.
 * comm may be null (unsharded DB); with a communicator every rank passes its own n patches (same n on all
 * ranks) and gets its own n x k results.  out_desc (n x dim) may be null. */
PE_API int pe_match_patches(pe_net* net, pe_db* db, pe_comm* comm, const float* patches_hwc, int64_t n,
                            int patch_channels, int k, int32_t* out_idx, float* out_dist, float* out_desc);
PE_API int pe_knn_hamming_sharded(pe_db* db, pe_comm* comm, const uint32_t* qcodes, int64_t nq,
                                  int k, int32_t* out_idx, int32_t* out_dist);
PE_API int pe_knn_hamming_sharded_dev(pe_db* db, pe_comm* comm, const uint32_t* d_qcodes, int64_t nq,
                                      int k, int32_t* d_out_idx, int32_t* d_out_dist);
/* Device-resident, asynchronous twin of pe_match_patches, pipelined over two streams: the forward of the n patches is
 * queued on the net's stream, [all-gather of the descriptors ->] scan -> select [-> all-gather of the candidates ->
 * merge] on the DB's stream behind it (event-ordered, two internal descriptor buffers), so the NEXT call's forward runs
 * under THIS call's collectives and matcher tail.  d_out_idx / d_out_dist: n x k neighbours of this rank's
 * patches; d_out_desc (n x dim) may be null.  Results are complete after pe_db_sync(db); pe_net_sync(net) reports
 * forward errors.  Outputs of consecutive calls must not alias while both are in flight.  With pe_db_share_stream the
 * call degenerates to one stream (no overlap). */
PE_API int pe_match_patches_dev(pe_net* net, pe_db* db, pe_comm* comm, const float* d_patches_hwc, int64_t n,
                                int patch_channels, int k, int32_t* d_out_idx, float* d_out_dist, float* d_out_desc);
/* computeHistogram / computeConfusionMatrix bodies (reference src/networkevaluator.cpp:269-355, 199-245) over a query
 * set that is split across the ranks: every rank passes its own nq_local queries (same count on every rank) with their
 * merged global neighbours; bin counts / the error sum / the confusion counts are all-reduced and the per-query errors
 * all-gathered for the median, so every rank returns the result of the whole set.  comm == null: single rank. */
PE_API int pe_histogram_sharded_dev(pe_db* db, pe_comm* comm, const int32_t* d_knn_idx, const float* d_q_labels,
                                    int64_t nq_local, int k, const int* rot_inv, int n_rot, const float* bins,
                                    int n_bins, float* out);
PE_API int pe_confusion_sharded_dev(pe_db* db, pe_comm* comm, const int32_t* d_knn_idx, const float* d_q_labels,
                                    int64_t nq_local, int k, const int* class_to_local, int n_classes, int n_local,
                                    int nr_test_poses, float* out);

#ifdef __cplusplus
}
#endif
#endif /* POSE_B200_H */
