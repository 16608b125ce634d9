/*
 * common.h - internal declarations shared by the translation units of libpose_b200.so.
 * Nothing here is part of the public boundary (include/pose_b200.h).
 */
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/pose_b200.h"

namespace pe {

/* ---------------------------------------------------------------- errors --- */
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
extern std::atomic<int64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define PE_CUDA(expr)                                                                          \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            pe::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,                   \
                          cudaGetErrorString(_e));                                             \
            return PE_ERR_CUDA;                                                                \
        }                                                                                      \
    } while (0)

#define PE_CHECK(expr)                                                                         \
    do {                                                                                       \
        int _s = (expr);                                                                       \
        if (_s != PE_OK) return _s;                                                            \
    } while (0)

#define PE_REQUIRE(cond, ...)                                                                  \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            pe::set_error(__VA_ARGS__);                                                        \
            return PE_ERR_INVALID;                                                             \
        }                                                                                      \
    } while (0)

/* internal status (never crosses the C ABI): the opt-in e4m3 cross terms (forward_fp8 = 1) met an input beyond their
 * range, the net has switched to the all-fp16 split and the call should be replayed */
#define PE_RETRY_FP16 100

/* after a kernel launch: catches launch-configuration errors and counts the launch */
#define PE_LAUNCHED()                                                                          \
    do {                                                                                       \
        pe::count_launch();                                                                    \
        PE_CUDA(cudaGetLastError());                                                           \
    } while (0)

/* a device buffer that grows on demand (never shrinks); bound to one device */
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes, bool zero_new = false);
    void release();
    template <typename T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

/* ------------------------------------------------------------------ net --- */
enum { L_CONV = 1, L_POOL = 2, L_ACT = 3, L_FC = 4 };

struct Layer {
    int kind = 0;
    std::string name;
    int a = 0, b = 0;          /* conv: num_output, kernel; pool: kernel, stride; act: PE_ACT_*; fc: num_output */
    int blob0 = -1, blob1 = -1;
};

struct NetDef {
    int batch = 0, channels = 0, height = 0, width = 0;
    std::vector<int> sel;      /* input-blob channels fed to conv1 */
    std::vector<Layer> layers;
    std::vector<std::string> blob_names;
    std::vector<std::vector<int64_t>> blob_shapes;
    int desc_dim = 0;
};

/* prototxt.cpp: TEST-phase view of a Caffe prototxt; returns PE_OK or PE_ERR_PARSE */
int parse_prototxt(const std::string& text, NetDef* out);
/* built-in emitter for the seven reference variants (network/manifold*.prototxt) */
bool builtin_prototxt(const std::string& variant, std::string* text);

struct FcPlan {
    int in = 0, out = 0;       /* logical sizes */
    int ld_in = 0;             /* padded row stride of the input activation and of the device weight */
    int blob_w = -1, blob_b = -1;
    int act = PE_ACT_NONE, blob_slope = -1;
    float* d_w = nullptr;      /* [out][ld_in] */
    float* d_b = nullptr;      /* [out] */
    float* d_slope = nullptr;  /* [out] or null */
    /* tensor path (fc_tc.cu): packed fp16 hi/lo weight blocks, K blocks of 64, output scale */
    bool tc_ready = false;
    void* d_wp = nullptr;
    int k_blocks = 0;
    float inv_scale = 1.f;
};

}  // namespace pe

struct pe_net {
    int device = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    int num_sms = 0;
    pe::NetDef def;
    std::vector<std::vector<float>> blobs; /* host master copy, Caffe order */
    bool weights_dirty = true;

    /* recognised plan: conv1(16,8) pool(2,2) [act] conv2(7,5) pool(2,2) [act] fc (act fc)*  */
    int cin = 0;                      /* channels fed to conv1 */
    int act1 = 0, act2 = 0;
    int blob_w1 = -1, blob_b1 = -1, blob_s1 = -1, blob_w2 = -1, blob_b2 = -1, blob_s2 = -1;
    std::vector<pe::FcPlan> fcs;
    float *d_w1 = nullptr, *d_b1 = nullptr, *d_s1 = nullptr; /* [cin][8][8][16], [16], [16] */
    float *d_w2 = nullptr, *d_b2 = nullptr, *d_s2 = nullptr; /* [16][5][5][8],  [8],  [8]  */
    int* d_slot = nullptr;            /* patch channel -> conv1 input slot or -1 */
    /* tensor path (conv_tc.cu): Toeplitz fp16 hi/lo conv1 weights, output scale, pinned failure flag */
    bool c1_tc_ready = false;
    int fp8_state = 0;         /* conv1 cross terms for these weights: 0 not calibrated yet, 1 e4m3, 2 fp16 (see net_forward_dev) */
    int fp8_calls = 0;         /* forward calls since the last qualification (forward_fp8 = 1 re-qualifies every 32nd) */
    bool fp8_busy = false;     /* inside the qualification's own forwards */
    void* d_z1 = nullptr;
    float c1_inv_scale = 1.f;
    bool c2_tc_ready = false;
    void* d_z2 = nullptr;
    float c2_inv_scale = 1.f;
    int* tc_err = nullptr;

    bool profile = false;
    std::vector<cudaEvent_t> prof_events; /* groups of 4: before conv1, after conv1, after conv2, after fcs */

    /* per-chunk activations (sized for `chunk` patches) and staging for the host entry point */
    int64_t chunk = 0;
    pe::DevBuf pool1, p1_packed, pool2, fc_a, fc_b, in_stage[2], out_stage;
};

struct pe_db {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    bool profile = false;
    std::vector<cudaEvent_t> prof_events; /* groups of 3: before scan, after scan, after select */
    int num_sms = 0;
    int64_t t_local = 0, t_total = 0, row_offset = 0;
    int dim = 0, dim_pad = 0;         /* descriptor dim and its multiple-of-4 padding */
    int words = 0;                    /* Hamming: u32 words per code */
    bool hamming = false;
    int poses_per_object = 0;
    float* d_desc = nullptr;          /* [t_local][dim_pad], zero padded */
    float* d_norm = nullptr;          /* [t_local] squared norms */
    uint32_t* d_codes = nullptr;      /* [t_local][words] */
    float* d_labels = nullptr;        /* [t_total][8] or null */
    /* tensor-core scan (knn_tc.cu): fp16 hi/lo/norm tiles of 128 templates, power-of-two scales */
    bool tc_ready = false;
    void* d_tc = nullptr;
    int* d_tc_err = nullptr;
    int tc_log2_st = 0, tc_log2_sn = 0;
    /* L2 is translation invariant: both scans rank on descriptors centred on the DB mean, so that the cancellation in
     * |t|^2 - 2 q.t is measured against the spread of the DB, not its distance from the origin */
    float center[128] = {};           /* DB mean (host copy) */
    float* d_center = nullptr;        /* [128] the same on the device */
    float* d_norm_c = nullptr;        /* [t_local] squared norms of the centred rows (SIMT scan) */
    float max_abs_c = 0.f, max_norm2_c = 0.f;   /* max |t-c| over elements, max |t-c|^2 over rows */
    pe::DevBuf tau;                   /* [nq] shared admission thresholds of the tensor scan */
    pe::DevBuf key_scale;             /* [nq] 1 / (s_q s_t): tensor-scan keys -> squared-distance units */
    /* exactness certificate of the L2 matcher: [0] number of queries the select pass could not prove, [1] calls, [4..]
     * their ids; fb_*: per (query, DB chunk) partial top-k of the direct (q-t)^2 re-scan and its arrival counters */
    pe::DevBuf flag, fb_d, fb_i, fb_done;
    long long flagged_total = 0, certified_calls = 0;   /* read back lazily by pe_db_knn_stats */
    pe::DevBuf q, q_in, cand_d, cand_i, out_i, out_d, qlab, knn_idx, scratch, scratch2, gather, boot;
    pe::DevBuf m_q, m_ci, m_cd, m_ri, m_rd;   /* pe_match_patches: gathered chunk queries, chunk and final results */
    cudaEvent_t ev_match = nullptr;           /* forward of a chunk finished (net stream -> DB stream) */
    /* pe_match_patches_dev: two descriptor buffers so that call i + 1's forward (net stream) runs under call i's
     * exchange + scan + merge (DB stream); ev_free[b] = the last reader of buffer b has finished */
    pe::DevBuf m_desc[2];
    cudaEvent_t ev_free[2] = {nullptr, nullptr};
    int64_t match_calls = 0;
    struct pe_comm* last_comm = nullptr;      /* communicator of the last sharded call (pe_db_sync reports its failures) */
};

namespace pe {
/* records a profiling event on `stream` when `on` (appends to `events`) */
int prof_mark(bool on, std::vector<cudaEvent_t>* events, cudaStream_t stream);
int net_ensure_device(pe_net* net);
/* forward.cu */
int net_upload_weights(pe_net* net);
int net_forward_dev(pe_net* net, const float* d_patches, int64_t n, int patch_c, float* d_out,
                    int ld_out);
int net_reserve(pe_net* net, int64_t chunk);
/* conv_tc.cu */
int conv1_tc_prepare(pe_net* net);
bool conv1_tc_usable(const pe_net* net, int patch_c, const void* d_patches);
int conv1_tc_launch(pe_net* net, const float* d_patches, int patch_c, long long n, bool packed_out);
int conv2_tc_prepare(pe_net* net);
int conv2_tc_launch(pe_net* net, long long n);
size_t conv_tc_p1_packed_bytes();
int forward_tc_check(pe_net* net);
/* fc_tc.cu */
int fc_tc_prepare(pe_net* net, FcPlan* fc);
int fc_tc_launch(pe_net* net, const FcPlan& fc, const float* x, int ldx, float* y, int ldy, int n_store, int m);
/* knn.cu */
int db_set_descriptors(pe_db* db, const float* src, bool src_on_device);
int knn_l2_local(pe_db* db, const float* d_queries, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist);
int knn_hamming_local(pe_db* db, const uint32_t* d_qcodes, int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist);
int knn_merge_f32(pe_db* db, const float* d_cd, const int* d_ci, long long stride_l, int n_lists, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist);
int knn_merge_i32(pe_db* db, const int* d_cd, const int* d_ci, long long stride_l, int n_lists, int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist);
/* knn_tc.cu */
int knn_tc_prepare(pe_db* db);
bool knn_tc_usable(const pe_db* db, int k);
void knn_tc_plan(const pe_db* db, int64_t nq, int k, int* n_chunks, int* kc, int* tiles_per_chunk);
int knn_tc_scan(pe_db* db, const void* d_q, int64_t nq, int k, int n_lists, int kc, int tiles_per_chunk,
                float* cand_d, int* cand_i);
int knn_tc_select(pe_db* db, const float* cand_d, const int* cand_i, int n_chunks, int kc, const float* d_q,
                  int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist, int* d_flag);
/* knn.cu: DB mean, centred norms and spread (called once per L2 DB); direct (q-t)^2 re-scan of the flagged queries */
int knn_center_stats(pe_db* db);
int knn_exact_fallback(pe_db* db, const int* d_flag, const float* d_q, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist);
int knn_tc_select_hamming(pe_db* db, const float* cand_d, const int* cand_i, int n_lists, int kc, const uint32_t* d_q,
                          int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist);
int knn_tc_check(pe_db* db);
/* comm.cu: after a synchronisation, reports (and clears) a peer-window all-gather that timed out */
int comm_check(struct pe_comm* comm);
/* runtime switches (pe_set_option): tensor-core paths on/off */
extern std::atomic<int> g_opt_knn_tensor, g_opt_forward_tensor, g_opt_forward_fp8, g_opt_p2p, g_opt_conv2_tensor, g_opt_knn_certify;
/* pose.cu */
int pose_histogram_dev(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                       const int* rot_inv, int n_rot, const float* bins, int n_bins, float* out_host);
int pose_confusion_dev(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                       const int* class_to_local, int n_classes, int n_local, int nr_test_poses,
                       float* out_host);
int pose_hist_local(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k, const int* rot_inv, int n_rot,
                    const float* bins, int n_bins, int64_t err_cap, unsigned long long** d_counts_out, double** d_sum_out,
                    float** d_med_out, float** d_err_out);
int pose_hist_finish(pe_db* db, const unsigned long long* d_counts, const double* d_sum, float* d_med, const float* d_err,
                     int64_t n_err, int64_t nq_total, int n_bins, float* out_host);
int pose_confusion_local(pe_db* db, const int* d_knn_idx, const float* d_q_labels, int64_t nq, int k, const int* class_to_local,
                         int n_classes, int n_local, unsigned int** d_mat_out);
int pose_confusion_finish(pe_db* db, const unsigned int* d_mat, int n_local, int nr_test_poses, float* out_host);
int pose_bootstrap_dev(pe_db* db, const int* d_knn3, int64_t nq, int nr_query_poses, int* d_out, int* d_len);
int pose_accuracy_dev(pe_db* db, const int* d_max_sim0, const int* d_boot, int64_t nq, int nr_query_poses,
                      float* intra_pct, float* inter_pct);
/* synth.cu */
int synth_patches_dev(cudaStream_t stream, uint64_t seed, int64_t first, int64_t n, int channels,
                      int poses_per_object, float* d_out);
void synth_labels_host(uint64_t seed, int64_t first, int64_t n, int poses_per_object, float* labels);
uint64_t splitmix64(uint64_t x);
uint64_t stream_key(uint64_t seed, uint64_t stream);
}  // namespace pe
