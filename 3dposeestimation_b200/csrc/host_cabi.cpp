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
