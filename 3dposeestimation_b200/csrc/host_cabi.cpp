/*
 * host_cabi.cpp - flat-array doors into the C++ host mirror (include/networkevaluator_b200.hpp) so the
 * Python tests and bench can drive sz::networkEvaluator / sz::networkSolver::computeKNN exactly as a
 * C++ caller of the reference would (vector<vector<Sample>> sets, static methods).  Built into
 * libpose_b200_host.so, which links libpose_b200.so; nothing here computes anything itself.
 */
#include "../../include/networkevaluator_b200.hpp"

#include <csignal>
#include <cstdlib>
#include <execinfo.h>
#include <unistd.h>

namespace {

thread_local std::string g_err;

/* n_obj x poses patches (HWC, `channels` each) + labels (8 floats each) -> vector<vector<Sample>> */
sz::SampleSet make_set(const float* patches, const float* labels, int n_obj, int poses, int channels) {
    sz::SampleSet set(n_obj, std::vector<sz::Sample>(poses));
    const size_t img = (size_t)64 * 64 * channels;
    for (int o = 0; o < n_obj; ++o)
        for (int p = 0; p < poses; ++p) {
            sz::Sample& s = set[o][p];
            const size_t i = (size_t)o * poses + p;
            s.data = sz::Mat(64, 64, channels);
            std::memcpy(s.data.ptr<float>(0), patches + i * img, img * sizeof(float));
            s.label = sz::Mat(1, 8);
            std::memcpy(s.label.ptr<float>(0), labels + i * 8, 8 * sizeof(float));
        }
    return set;
}

struct FlatDataset : sz::DatasetView {
    sz::SampleSet tmpl, train, test;
    const sz::SampleSet& getTemplateSet() const override { return tmpl; }
    const sz::SampleSet& getTrainingSet() const override { return train; }
    const sz::SampleSet& getTestSet() const override { return test; }
};

template <typename F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

}  // namespace

#define HOST_API extern "C" __attribute__((visibility("default")))

/* PE_HOST_BACKTRACE=1: print the native stack on SIGSEGV (debugging aid for callers that load the library into an
 * interpreter) */
namespace {
void on_segv(int) {
    void* bt[64];
    const int n = backtrace(bt, 64);
    backtrace_symbols_fd(bt, n, 2);
    _exit(139);
}
__attribute__((constructor)) void install_backtrace() {
    const char* e = std::getenv("PE_HOST_BACKTRACE");
    if (e && *e == '1') std::signal(SIGSEGV, on_segv);
}
}  // namespace

HOST_API const char* pe_host_last_error() { return g_err.c_str(); }

HOST_API int pe_host_compute_descriptors(pe_net* net, const float* patches, const float* labels, int n, int channels,
                                         float* out) {
    return guarded([&] {
        sz::Net cnn = sz::Net::borrow(net);
        sz::SampleSet set = make_set(patches, labels, 1, n, channels);
        sz::Mat d = sz::networkEvaluator::computeDescriptors(cnn, set[0]);
        std::memcpy(out, d.ptr<float>(0), d.buf.size() * sizeof(float));
    });
}

HOST_API int pe_host_compute_histogram(pe_net* net, const float* t_patches, const float* t_labels, int n_obj, int nr_t,
                                       const float* q_patches, const float* q_labels, int nr_q, int channels,
                                       const int* rot_inv, int n_rot, const float* bins, int n_bins, int knn, float* out) {
    return guarded([&] {
        sz::Net cnn = sz::Net::borrow(net);
        const sz::SampleSet ts = make_set(t_patches, t_labels, n_obj, nr_t, channels);
        const sz::SampleSet qs = make_set(q_patches, q_labels, n_obj, nr_q, channels);
        std::vector<float> h = sz::networkEvaluator::computeHistogram(cnn, ts, qs, std::vector<int>(rot_inv, rot_inv + n_rot),
                                                                     std::vector<float>(bins, bins + n_bins), knn);
        std::memcpy(out, h.data(), h.size() * sizeof(float));
    });
}

/* models: comma-separated names, one per class label; used_models: the subset that gets a row/column */
HOST_API int pe_host_compute_confusion(pe_net* net, const float* t_patches, const float* t_labels, int n_obj, int nr_t,
                                       const float* q_patches, const float* q_labels, int nr_q, int channels,
                                       const char* models_csv, const char* used_models_csv, int knn, float* out) {
    return guarded([&] {
        sz::Net cnn = sz::Net::borrow(net);
        const sz::SampleSet ts = make_set(t_patches, t_labels, n_obj, nr_t, channels);
        const sz::SampleSet qs = make_set(q_patches, q_labels, n_obj, nr_q, channels);
        const std::vector<std::string> models = sz::to_array<std::string>(models_csv);
        const std::vector<std::string> used = sz::to_array<std::string>(used_models_csv);
        std::unordered_map<std::string, int> local_index;
        for (size_t i = 0; i < used.size(); ++i) local_index[used[i]] = (int)i;
        auto cm = sz::networkEvaluator::computeConfusionMatrix(cnn, ts, qs, models, local_index, knn);
        for (size_t r = 0; r < cm.size(); ++r) std::memcpy(out + r * cm.size(), cm[r].data(), cm.size() * sizeof(float));
    });
}

/* networkSolver::computeKNN + networkEvaluator::computeKNNAccuracy; out: nq x 4 (-1 padded), out_len: nq */
HOST_API int pe_host_compute_knn(pe_net* net, const float* t_patches, const float* t_labels, int n_obj, int nr_t,
                                 const float* q_patches, const float* q_labels, int nr_q, int channels,
                                 const int* max_sim_tmpl0, int* out, int* out_len, float* intra, float* inter) {
    return guarded([&] {
        sz::Net cnn = sz::Net::borrow(net);
        const sz::SampleSet ts = make_set(t_patches, t_labels, n_obj, nr_t, channels);
        const sz::SampleSet qs = make_set(q_patches, q_labels, n_obj, nr_q, channels);
        sz::networkSolver solver(ts, qs, qs);
        solver.computeKNN(cnn);
        std::vector<std::vector<std::vector<int>>> maxSimTmpl(n_obj, std::vector<std::vector<int>>(nr_q));
        for (int o = 0; o < n_obj; ++o)
            for (int p = 0; p < nr_q; ++p) {
                const std::vector<int>& v = solver.maxSimKNNTmpl[o][p];
                const size_t i = (size_t)o * nr_q + p;
                for (int j = 0; j < 4; ++j) out[i * 4 + j] = j < (int)v.size() ? v[j] : -1;
                out_len[i] = (int)v.size();
                maxSimTmpl[o][p] = {max_sim_tmpl0 ? max_sim_tmpl0[i] : 0};
            }
        sz::networkEvaluator::computeKNNAccuracy(maxSimTmpl, solver.maxSimKNNTmpl, intra, inter);
    });
}

/* networkEvaluator::saveLog + saveConfusionMatrix with the reference's config-file keys */
HOST_API int pe_host_save_logs(pe_net* net, const float* t_patches, const float* t_labels, int n_obj, int nr_t,
                               const float* tr_patches, const float* tr_labels, int nr_tr, const float* q_patches,
                               const float* q_labels, int nr_q, int channels, const char* config_path, int iter, float time) {
    return guarded([&] {
        sz::Net cnn = sz::Net::borrow(net);
        FlatDataset db;
        db.tmpl = make_set(t_patches, t_labels, n_obj, nr_t, channels);
        db.train = make_set(tr_patches, tr_labels, n_obj, nr_tr, channels);
        db.test = make_set(q_patches, q_labels, n_obj, nr_q, channels);
        sz::networkEvaluator::saveLog(cnn, db, config_path, iter, time);
        sz::networkEvaluator::saveConfusionMatrix(cnn, db, config_path);
    });
}

/* networkSolver::buildBatch / buildBatchQueue's batch_caffe through the mirror: knn_table (nq x 4, -1 padded) + knn_len
 * switch bootstrapping on; out_batch [batch_size / 5 * 5][channels][64][64], out_triplets [batch_size / 5][5],
 * out_max_sim [n_obj * nr_q][2]; also returns the vector<Sample> form's data in out_samples_hwc when non-null */
HOST_API int pe_host_build_batch(const float* t_patches, const float* t_labels, int n_obj, int nr_t, const float* q_patches,
                                 const float* q_labels, int nr_q, int channels, const int* rot_inv, int inplane,
                                 const int* knn_table, const int* knn_len, int batch_size, int iter, uint64_t seed,
                                 int random_background, float* out_batch, int* out_triplets, int* out_max_sim,
                                 float* out_samples_hwc) {
    return guarded([&] {
        const sz::SampleSet ts = make_set(t_patches, t_labels, n_obj, nr_t, channels);
        const sz::SampleSet qs = make_set(q_patches, q_labels, n_obj, nr_q, channels);
        sz::networkSolver solver(ts, qs, qs, std::vector<int>(rot_inv, rot_inv + n_obj), inplane != 0);
        solver.seed = seed;
        solver.random_background = random_background;
        if (knn_table) {
            solver.maxSimKNNTmpl.assign(n_obj, std::vector<std::vector<int>>(nr_q));
            for (int q = 0; q < n_obj * nr_q; ++q)
                solver.maxSimKNNTmpl[q / nr_q][q % nr_q].assign(knn_table + (size_t)q * 4, knn_table + (size_t)q * 4 + knn_len[q]);
            solver.bootstrapping = true;
        }
        const std::vector<float> b = solver.buildBatchCaffe(batch_size, 5, iter);
        std::memcpy(out_batch, b.data(), b.size() * sizeof(float));
        std::memcpy(out_triplets, solver.last_triplets.data(), solver.last_triplets.size() * sizeof(int));
        for (int q = 0; q < n_obj * nr_q; ++q) {
            out_max_sim[q * 2] = solver.maxSimTmpl[q / nr_q][q % nr_q][0];
            out_max_sim[q * 2 + 1] = solver.maxSimTmpl[q / nr_q][q % nr_q][1];
        }
        if (out_samples_hwc) {
            sz::networkSolver again(ts, qs, qs, std::vector<int>(rot_inv, rot_inv + n_obj), inplane != 0);
            again.seed = seed;
            again.maxSimKNNTmpl = solver.maxSimKNNTmpl;
            again.bootstrapping = solver.bootstrapping;
            const std::vector<sz::Sample> v = again.buildBatch(batch_size, 5, iter);
            const size_t img = (size_t)64 * 64 * channels;
            for (size_t i = 0; i < v.size(); ++i) std::memcpy(out_samples_hwc + i * img, v[i].data.buf.data(), img * sizeof(float));
        }
    });
}

/* ---- sample producers through the mirror (include/datasetmanager_b200.hpp) ---- */
#include "../../include/datasetmanager_b200.hpp"

namespace {
sz::Matrix3f make_cam(const float* cam9) {
    sz::Matrix3f cam;
    std::copy(cam9, cam9 + 9, cam.v.begin());
    return cam;
}
/* out_patches / out_labels may be null to query the count; cap = samples the arrays hold */
void copy_samples(const std::vector<sz::Sample>& s, float* out_patches, float* out_labels, int64_t cap, int64_t* n) {
    *n = (int64_t)s.size();
    if (!out_patches && !out_labels) return;
    if (cap < *n) throw std::runtime_error("output arrays hold " + std::to_string(cap) + " samples, " + std::to_string(*n) + " produced");
    const size_t img = (size_t)64 * 64 * 7;
    for (size_t i = 0; i < s.size(); ++i) {
        if (out_patches) std::memcpy(out_patches + i * img, s[i].data.buf.data(), img * sizeof(float));
        if (out_labels) std::memcpy(out_labels + i * 8, s[i].label.buf.data(), 8 * sizeof(float));
    }
}
}  // namespace

/* Model::loadModel + datasetManager::createSynthSamplesWadim (paul == 0) or createSynthSamplesPaul (paul != 0, reads
 * dataset_path + "paul/camPositionsElAz.txt"); tex (may be null) [tex_rows][tex_cols][3] for a textured model */
HOST_API int pe_host_create_synth_samples(const char* ply_path, const float* cam9, int index, int subdiv, int inplane, int paul,
                                          const char* dataset_path, const char* dataset_name, const uint8_t* tex, int tex_rows,
                                          int tex_cols, float* out_patches, float* out_labels, int64_t cap, int64_t* n) {
    return guarded([&] {
        sz::Model model;
        if (!model.loadModel(ply_path)) throw std::runtime_error(std::string("pose_b200: ") + pe_last_error());
        if (tex) model.setTexture(tex, tex_rows, tex_cols);
        sz::Matrix3f cam = make_cam(cam9);
        sz::datasetManager db;
        db.inplane = inplane != 0;
        if (dataset_path) db.dataset_path = dataset_path;
        if (dataset_name) db.dataset_name = dataset_name;
        copy_samples(paul ? db.createSynthSamplesPaul(model, cam, index) : db.createSynthSamplesWadim(model, cam, index, subdiv),
                     out_patches, out_labels, cap, n);
    });
}

/* datasetManager::extractRealSamplesPaul (wadim == 0; ply_path gives Model::centroid for dataset_name "BigBIRD") or
 * extractRealSamplesWadim (wadim != 0) over n_frames frames with one ground-truth pose each */
HOST_API int pe_host_extract_real_samples(const uint8_t* color, const float* depth, int n_frames, int rows, int cols,
                                          const float* poses, const float* cam9, int index, int wadim, const char* ply_path,
                                          const char* dataset_name, float* out_patches, float* out_labels, int64_t cap, int64_t* n) {
    return guarded([&] {
        const size_t px = (size_t)rows * cols;
        std::vector<sz::Frame> frames(n_frames);
        for (int i = 0; i < n_frames; ++i) {
            sz::Frame& f = frames[i];
            f.nr = i; f.rows = rows; f.cols = cols;
            f.color.assign(color + i * px * 3, color + (i + 1) * px * 3);
            f.depth.assign(depth + i * px, depth + (i + 1) * px);
            sz::Isometry3f pose;
            std::copy(poses + (size_t)i * 12, poses + (size_t)i * 12 + 12, pose.m.begin());
            f.gt.push_back({"object", pose});
        }
        sz::Model model;
        if (ply_path && !model.loadModel(ply_path)) throw std::runtime_error(std::string("pose_b200: ") + pe_last_error());
        sz::Matrix3f cam = make_cam(cam9);
        sz::datasetManager db;
        if (dataset_name) db.dataset_name = dataset_name;
        copy_samples(wadim ? db.extractRealSamplesWadim(frames, cam, index) : db.extractRealSamplesPaul(frames, cam, index, model),
                     out_patches, out_labels, cap, n);
    });
}

/* SphereRenderer::createViews (clip = false) keeping the frames: out_color [n][480][640][3], out_depth [n][480][640],
 * out_poses [n][12]; all three null to query the count */
HOST_API int pe_host_create_views(const char* ply_path, const float* cam9, int sphere_dep, const float* scale3, const float* rot3,
                                  int skip_lower_hemi, uint8_t* out_color, float* out_depth, float* out_poses, int64_t cap, int64_t* n) {
    return guarded([&] {
        const sz::Vector3f sc(scale3[0], scale3[1], scale3[2]), ro(rot3[0], rot3[1], rot3[2]);
        if (!out_color && !out_depth && !out_poses) {
            *n = (int64_t)sz::SphereRenderer::createPoses(sphere_dep, sc, ro, skip_lower_hemi != 0).size();
            return;
        }
        sz::Model model;
        if (!model.loadModel(ply_path)) throw std::runtime_error(std::string("pose_b200: ") + pe_last_error());
        sz::SphereRenderer sphere(make_cam(cam9));
        const std::vector<sz::RenderView> views = sphere.createViews(model, sphere_dep, sc, ro, skip_lower_hemi != 0, false);
        *n = (int64_t)views.size();
        if (cap < *n) throw std::runtime_error("output arrays hold " + std::to_string(cap) + " views, " + std::to_string(*n) + " produced");
        const size_t px = (size_t)sz::kFrameRows * sz::kFrameCols;
        for (size_t i = 0; i < views.size(); ++i) {
            if (out_color) std::memcpy(out_color + i * px * 3, views[i].col.data(), px * 3);
            if (out_depth) std::memcpy(out_depth + i * px, views[i].dep.data(), px * sizeof(float));
            if (out_poses) std::memcpy(out_poses + i * 12, views[i].pose.m.data(), 12 * sizeof(float));
        }
    });
}
