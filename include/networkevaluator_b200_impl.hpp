/* networkevaluator_b200_impl.hpp - included by networkevaluator_b200.hpp; see the notes there. */
#pragma once

namespace sz {

namespace detail {

/* owns a pe_db for the duration of one evaluator call */
struct DbGuard {
    pe_db* db = nullptr;
    ~DbGuard() { if (db) pe_db_destroy(db); }
};

inline std::vector<float> flatten_labels(const SampleSet& set) {
    std::vector<float> out;
    for (auto& obj : set)
        for (auto& s : obj) out.insert(out.end(), s.label.ptr<float>(0), s.label.ptr<float>(0) + 8);
    return out;
}

inline void check_rectangular(const SampleSet& set, const char* what) {
    if (set.empty() || set[0].empty()) throw std::runtime_error(std::string(what) + " is empty");
    for (auto& obj : set)
        if (obj.size() != set[0].size())
            throw std::runtime_error(std::string(what) + " must hold the same number of poses per object");
}

}  // namespace detail

class networkEvaluator {
public:
    networkEvaluator() {}

    /* src/networkevaluator.cpp:9-54.  Batching, the HWC->CHW staging copy and the forward are one
     * call into the CUDA library; rows come back in sample order. */
    static Mat computeDescriptors(Net& CNN, std::vector<Sample> samples) {
        int batch, channels, height, width, desc_dim;
        CNN.shape(&batch, &channels, &height, &width, &desc_dim);
        Mat descs((int)samples.size(), desc_dim);
        if (samples.empty()) return descs;
        const size_t img = (size_t)height * width * channels;
        std::vector<float> staging(samples.size() * img);
        for (size_t i = 0; i < samples.size(); ++i) {
            const Mat& patch = samples[i].data;
            if (patch.rows != height || patch.cols != width || patch.channels() != channels)
                throw std::runtime_error("computeDescriptors: patch shape does not match the net input");
            std::memcpy(staging.data() + i * img, patch.ptr<float>(0), img * sizeof(float));
        }
        pe_check(pe_compute_descriptors(CNN.handle(), staging.data(), (int64_t)samples.size(), channels, descs.ptr<float>(0)));
        return descs;
    }

    /* src/networkevaluator.cpp:153-175 (prints the two accuracies like the reference) */
    static void computeKNNAccuracy(std::vector<std::vector<std::vector<int>>>& maxSimTmpl,
                                   std::vector<std::vector<std::vector<int>>>& maxSimKNNTmpl) {
        float intra, inter;
        computeKNNAccuracy(maxSimTmpl, maxSimKNNTmpl, &intra, &inter);
        std::cout << "Intra-class accuracy: " << intra << std::endl;
        std::cout << "Inter-class accuracy: " << inter << std::endl;
    }
    static void computeKNNAccuracy(const std::vector<std::vector<std::vector<int>>>& maxSimTmpl,
                                   const std::vector<std::vector<std::vector<int>>>& maxSimKNNTmpl, float* intra, float* inter) {
        const int nr_objects = (int)maxSimTmpl.size();
        const int nr_training_poses = (int)maxSimTmpl[0].size();
        std::vector<int32_t> ms, boot;
        for (int o = 0; o < nr_objects; ++o)
            for (int p = 0; p < nr_training_poses; ++p) {
                ms.push_back(maxSimTmpl[o][p][0]);
                const std::vector<int>& b = maxSimKNNTmpl[o][p];
                for (int j = 0; j < 4; ++j) boot.push_back(j < (int)b.size() ? b[j] : -1);
            }
        detail::DbGuard g;
        const float zero[4] = {0, 0, 0, 0};
        pe_check(pe_db_create(zero, 1, 4, 0, nullptr, 1, 1, 0, &g.db));
        pe_check(pe_knn_accuracy(g.db, ms.data(), boot.data(), (int64_t)ms.size(), nr_training_poses, intra, inter));
    }

    /* src/networkevaluator.cpp:248-356 */
    static std::vector<float> computeHistogram(Net& CNN, const SampleSet& template_set, const SampleSet& test_set,
                                               std::vector<int> rotInv, std::vector<float> bins, int knn) {
        detail::DbGuard g;
        Mat DBtest;
        std::vector<int32_t> idx = match(CNN, template_set, test_set, knn, &g, &DBtest);
        const std::vector<float> ql = detail::flatten_labels(test_set);
        std::vector<float> histo(bins.size() + 2);
        pe_check(pe_histogram(g.db, idx.data(), ql.data(), DBtest.rows, knn, rotInv.data(), (int)rotInv.size(), bins.data(),
                              (int)bins.size(), histo.data()));
        return histo;
    }

    /* src/networkevaluator.cpp:178-246 */
    static std::vector<std::vector<float>> computeConfusionMatrix(Net& CNN, const SampleSet& template_set,
                                                                  const SampleSet& test_set, std::vector<std::string> models,
                                                                  std::unordered_map<std::string, int> local_index, int knn) {
        detail::DbGuard g;
        Mat DBtest;
        std::vector<int32_t> idx = match(CNN, template_set, test_set, knn, &g, &DBtest);
        const std::vector<float> ql = detail::flatten_labels(test_set);
        /* class label -> local_index[models[label]] (:215-216); models absent from local_index map to -1 */
        std::vector<int> class_to_local(models.size(), -1);
        for (size_t c = 0; c < models.size(); ++c) {
            auto it = local_index.find(models[c]);
            if (it != local_index.end()) class_to_local[c] = it->second;
        }
        const int n_local = (int)local_index.size();
        std::vector<float> flat((size_t)n_local * n_local);
        pe_check(pe_confusion(g.db, idx.data(), ql.data(), DBtest.rows, knn, class_to_local.data(), (int)class_to_local.size(),
                              n_local, (int)test_set[0].size(), flat.data()));
        std::vector<std::vector<float>> conf(n_local, std::vector<float>(n_local));
        for (int r = 0; r < n_local; ++r)
            for (int c = 0; c < n_local; ++c) conf[r][c] = flat[(size_t)r * n_local + c];
        return conf;
    }

    /* src/networkevaluator.cpp:416-486: appends the "test" and "train" rows to <output_path>log<log_name> */
    static void saveLog(Net& CNN, DatasetView& db, std::string config, int iter, float time) {
        IniFile pt(config);
        const int knn = pt.get<int>("output.kNN");
        const std::vector<int> rotInv = to_array<int>(pt.get<std::string>("input.rotInv"));
        (void)pt.get<int>("input.random_background"); /* read, unused: the fill is commented out at :434-446 */
        const SampleSet& template_set = db.getTemplateSet();
        const SampleSet& training_set = db.getTrainingSet();
        const SampleSet& test_set = db.getTestSet();
        const std::vector<float> bins = {-1, 0, 10, 20, 40, 180};
        const std::vector<float> test_hist = computeHistogram(CNN, template_set, test_set, rotInv, bins, knn);
        const std::vector<float> train_hist = computeHistogram(CNN, template_set, training_set, rotInv, bins, knn);

        const std::string log_name = pt.get<std::string>("output.log_name");
        const std::string output_path = pt.get<std::string>("paths.output_path");
        const std::vector<std::string> used_models = to_array<std::string>(pt.get<std::string>("input.used_models"));
        const size_t per_epoch = used_models.size() * training_set[0].size() / 60;
        const int epoch = per_epoch ? (int)(iter / per_epoch) : 0;

        std::ifstream check(output_path + "log" + log_name);
        const bool need_header = !check.good() || check.peek() == std::ifstream::traits_type::eof();
        std::ofstream log_file(output_path + "log" + log_name, std::ios::app);
        if (!log_file) throw std::runtime_error("cannot write " + output_path + "log" + log_name);
        if (need_header) {
            log_file << "mod" << "\t" << "iter" << "\t" << "epoch" << "\t" << "time" << "\t";
            for (size_t i = 0; i < bins.size(); ++i) log_file << "<" << bins[i] << "\t";
            log_file << "mean" << "\t" << "median" << std::endl;
        }
        auto row = [&](const char* mod, const std::vector<float>& h) {
            log_file << mod << "\t" << iter << "\t" << epoch << "\t" << time << "\t";
            for (size_t i = 0; i < bins.size(); ++i) log_file << std::setprecision(3) << h[i] * 100 << "\t";
            log_file << h[h.size() - 2] << "\t" << h.back() << std::endl;
        };
        row("test", test_hist);
        row("train", train_hist);
    }

    /* src/networkevaluator.cpp:358-414: confusion matrix of the TRAINING set (:390) in percent */
    static void saveConfusionMatrix(Net& CNN, DatasetView& db, std::string config) {
        IniFile pt(config);
        const int knn = pt.get<int>("output.kNN");
        (void)pt.get<int>("input.random_background");
        const std::vector<std::string> models = to_array<std::string>(pt.get<std::string>("input.models"));
        const std::vector<std::string> used_models = to_array<std::string>(pt.get<std::string>("input.used_models"));
        std::unordered_map<std::string, int> local_index;
        for (size_t i = 0; i < used_models.size(); ++i) local_index[used_models[i]] = (int)i;
        const std::vector<std::vector<float>> conf =
            computeConfusionMatrix(CNN, db.getTemplateSet(), db.getTrainingSet(), models, local_index, knn);
        const std::string log_name = pt.get<std::string>("output.log_name");
        const std::string output_path = pt.get<std::string>("paths.output_path");
        std::ofstream log_file(output_path + "cm" + log_name);
        if (!log_file) throw std::runtime_error("cannot write " + output_path + "cm" + log_name);
        for (size_t m = 0; m + 1 < used_models.size(); ++m) log_file << used_models[m] << "\t";
        log_file << used_models.back() << std::endl;
        for (size_t m = 0; m < used_models.size(); ++m) {
            log_file << used_models[m] << "\t";
            for (size_t v = 0; v + 1 < used_models.size(); ++v) log_file << std::setprecision(3) << conf[m][v] * 100 << "\t";
            log_file << std::setprecision(3) << conf[m].back() * 100 << std::endl;
        }
    }

    /* shared front half of computeHistogram / computeConfusionMatrix / computeKNN (:257-267):
     * descriptors per object pushed into DBfeats / DBtest, BruteForce matcher, knnMatch */
    static std::vector<int32_t> match(Net& CNN, const SampleSet& template_set, const SampleSet& query_set, int knn,
                                      detail::DbGuard* g, Mat* DBquery) {
        detail::check_rectangular(template_set, "template_set");
        detail::check_rectangular(query_set, "query set");
        if (query_set.size() != template_set.size()) throw std::runtime_error("template and query sets differ in object count");
        Mat DBfeats;
        for (size_t object = 0; object < template_set.size(); ++object)
            DBfeats.push_back(computeDescriptors(CNN, template_set[object]));
        const std::vector<float> tl = detail::flatten_labels(template_set);
        pe_check(pe_db_create(DBfeats.ptr<float>(0), DBfeats.rows, DBfeats.cols, 0, tl.data(), DBfeats.rows,
                              (int)template_set[0].size(), 0, &g->db));
        /* the query side (DBtest.push_back(computeDescriptors(...)) per object, then knnMatch) as one fused call:
         * the patches of all objects in object order, descriptors and neighbours back together */
        int batch, channels, height, width, desc_dim;
        CNN.shape(&batch, &channels, &height, &width, &desc_dim);
        const size_t img = (size_t)height * width * channels;
        size_t nq = 0;
        for (const std::vector<Sample>& obj : query_set) nq += obj.size();
        std::vector<float> staging(nq * img);
        size_t at = 0;
        for (const std::vector<Sample>& obj : query_set)
            for (const Sample& smp : obj) {
                if (smp.data.rows != height || smp.data.cols != width || smp.data.ch != channels)
                    throw std::runtime_error("match: patch shape does not match the net input");
                std::memcpy(staging.data() + at * img, smp.data.buf.data(), img * sizeof(float));
                ++at;
            }
        *DBquery = Mat((int)nq, desc_dim);
        std::vector<int32_t> idx(nq * knn);
        std::vector<float> dist(nq * knn);
        pe_check(pe_match_patches(CNN.handle(), g->db, nullptr, staging.data(), (int64_t)nq, channels, knn, idx.data(),
                                  dist.data(), DBquery->ptr<float>(0)));
        return idx;
    }
};

/* the evaluation-side member of networkSolver (include/networksolver.h:49-57) and its batch builder (:38-43) */
class networkSolver {
public:
    /* rot_inv_per_object[object] = rotInv[model_index[used_models[object]]] (src/networksolver.cpp:78); `inplane`
     * selects computeMaxSimTmplInplane over computeMaxSimTmpl as datasetManager does (src/datasetmanager.cpp:683-684) */
    networkSolver(const SampleSet& templates, const SampleSet& training, const SampleSet& test,
                  std::vector<int> rot_inv_per_object = {}, bool inplane = true, int device = 0)
        : rotInv(std::move(rot_inv_per_object)), template_set(templates), training_set(training), test_set(test),
          inplane_(inplane), device_(device) {
        detail::check_rectangular(template_set, "template_set");
        detail::check_rectangular(training_set, "training_set");
        nr_objects = (unsigned)template_set.size();
        nr_training_poses = (unsigned)training_set.front().size();
        nr_template_poses = (unsigned)template_set.front().size();
        nr_test_poses = test_set.empty() ? 0u : (unsigned)test_set.front().size();
        if (rotInv.empty()) rotInv.assign(nr_objects, 0);
        if (rotInv.size() != nr_objects) throw std::runtime_error("rotInv must hold one entry per used object");
    }
    networkSolver(const networkSolver&) = delete;
    networkSolver& operator=(const networkSolver&) = delete;
    ~networkSolver() { if (miner_) pe_miner_destroy(miner_); }

    /* src/networksolver.cpp:456-505: k = 3 matcher over the training set -> maxSimKNNTmpl; both sets stay on the
     * device between calls.  As in trainNet (:308-309,332-333) the caller sets `bootstrapping` afterwards. */
    void computeKNN(Net& CNN) {
        ensureMiner();
        const size_t nq = (size_t)nr_objects * nr_training_poses;
        std::vector<int32_t> out(nq * 4), len(nq);
        pe_check(pe_miner_compute_knn(miner_, CNN.handle(), out.data(), len.data()));
        maxSimKNNTmpl.assign(nr_objects, std::vector<std::vector<int>>(nr_training_poses, std::vector<int>()));
        for (size_t linearId = 0; linearId < nq; ++linearId) {
            std::vector<int>& v = maxSimKNNTmpl[linearId / nr_training_poses][linearId % nr_training_poses];
            v.assign(out.begin() + linearId * 4, out.begin() + linearId * 4 + len[linearId]);
        }
        knn_on_device_ = true;
    }

    /* datasetManager::getHardNegatives as buildBatch reads it (:60,95-100): one (possibly empty) list per used object */
    void setHardNegatives(const std::vector<std::vector<Sample>>& negs) {
        ensureMiner();
        if (negs.size() != nr_objects) throw std::runtime_error("hard negatives: one list per used object");
        std::vector<int64_t> off(nr_objects + 1, 0);
        std::vector<float> flat;
        for (size_t o = 0; o < nr_objects; ++o) {
            off[o + 1] = off[o] + (int64_t)negs[o].size();
            for (const Sample& smp : negs[o]) flat.insert(flat.end(), smp.data.buf.begin(), smp.data.buf.end());
        }
        hard_negs_ = negs;
        pe_check(pe_miner_set_hard_negatives(miner_, flat.data(), off.data()));
    }

    /* the batch_caffe buffer buildBatchQueue pushes (src/networksolver.cpp:12-50) / trainNet fills (:353-368):
     * [batch_size / 5 * 5][channels][64][64].  Every call draws with a fresh seed (the reference draws from
     * std::random_device); `seed` makes a run repeatable. */
    std::vector<float> buildBatchCaffe(int batch_size, unsigned int triplet_size, int iter) {
        ensureMiner();
        syncBootstrapping();
        const size_t n_triplets = triplet_size ? (size_t)batch_size / triplet_size : 0;
        std::vector<float> batch_caffe(n_triplets * 5 * img_, 0.f);
        last_triplets.assign(n_triplets * 5, 0);
        pe_check(pe_miner_build_batch(miner_, batch_size, (int)triplet_size, iter, seed + batches_built_++, random_background,
                                      batch_caffe.data(), last_triplets.data()));
        return batch_caffe;
    }

    /* src/networksolver.cpp:52-186 with the reference's return type: anchor, puller, pusher0-2 per triplet (the
     * selection runs on the device, the Samples are copied from the host sets; no background fill here) */
    std::vector<Sample> buildBatch(int batch_size, unsigned int triplet_size, int iter) {
        ensureMiner();
        syncBootstrapping();
        const size_t n_triplets = triplet_size ? (size_t)batch_size / triplet_size : 0;
        last_triplets.assign(n_triplets * 5, 0);
        pe_check(pe_miner_build_batch(miner_, batch_size, (int)triplet_size, iter, seed + batches_built_++, 0, nullptr,
                                      last_triplets.data()));
        std::vector<Sample> batch;
        batch.reserve(n_triplets * 5);
        for (size_t t = 0; t < n_triplets; ++t)
            for (int slot = 0; slot < 5; ++slot) {
                const int32_t id = last_triplets[t * 5 + slot];
                Sample smp;
                if (slot == 0) smp.copySample(training_set[id / nr_training_poses][id % nr_training_poses]);
                else if (id >= 0) smp.copySample(template_set[id / nr_template_poses][id % nr_template_poses]);
                else smp.copySample(negSample(~id));
                batch.push_back(std::move(smp));
            }
        return batch;
    }

    std::vector<std::vector<std::vector<int>>> maxSimTmpl;      /* db.getMaxSimTmpl() (filled with the first batch) */
    std::vector<std::vector<std::vector<int>>> maxSimKNNTmpl;
    std::vector<int32_t> last_triplets;                         /* [n_triplets][5] indices of the last batch */
    std::vector<int> rotInv;
    bool bootstrapping = false;
    int random_background = 0;
    uint64_t seed = 0x5eed;
    unsigned int nr_objects, nr_training_poses, nr_template_poses, nr_test_poses;
    const SampleSet &template_set, &training_set, &test_set;

private:
    static std::vector<float> flatten_patches(const SampleSet& set, size_t img) {
        std::vector<float> out;
        out.reserve(set.size() * set[0].size() * img);
        for (auto& obj : set)
            for (auto& smp : obj) {
                if (smp.data.buf.size() != img) throw std::runtime_error("samples differ in patch size");
                out.insert(out.end(), smp.data.buf.begin(), smp.data.buf.end());
            }
        return out;
    }
    void ensureMiner() {
        if (miner_) return;
        const Mat& first = template_set[0][0].data;
        if (first.rows != 64 || first.cols != 64) throw std::runtime_error("patches must be 64x64");
        img_ = (size_t)64 * 64 * first.ch;
        const std::vector<float> tp = flatten_patches(template_set, img_), qp = flatten_patches(training_set, img_);
        const std::vector<float> tl = detail::flatten_labels(template_set), ql = detail::flatten_labels(training_set);
        pe_check(pe_miner_create(device_, first.ch, (int)nr_objects, (int)nr_template_poses, (int)nr_training_poses, tp.data(),
                                 tl.data(), qp.data(), ql.data(), rotInv.data(), inplane_ ? 1 : 0, &miner_));
        std::vector<int32_t> ms((size_t)nr_objects * nr_training_poses * 2);
        pe_check(pe_miner_max_sim(miner_, ms.data()));
        maxSimTmpl.assign(nr_objects, std::vector<std::vector<int>>(nr_training_poses));
        for (size_t q = 0; q < ms.size() / 2; ++q)
            maxSimTmpl[q / nr_training_poses][q % nr_training_poses] = {ms[q * 2], ms[q * 2 + 1]};
    }
    /* `bootstrapping` is a plain member in the reference: follow it, and push a table the caller filled by hand */
    void syncBootstrapping() {
        if (!bootstrapping) { pe_check(pe_miner_set_knn(miner_, nullptr, nullptr)); knn_on_device_ = false; return; }
        if (knn_on_device_) return;
        const size_t nq = (size_t)nr_objects * nr_training_poses;
        if (maxSimKNNTmpl.size() != nr_objects) throw std::runtime_error("bootstrapping without maxSimKNNTmpl (call computeKNN)");
        std::vector<int32_t> table(nq * 4, -1), len(nq);
        for (size_t q = 0; q < nq; ++q) {
            const std::vector<int>& v = maxSimKNNTmpl[q / nr_training_poses][q % nr_training_poses];
            len[q] = (int32_t)v.size();
            for (size_t j = 0; j < v.size() && j < 4; ++j) table[q * 4 + j] = v[j];
        }
        pe_check(pe_miner_set_knn(miner_, table.data(), len.data()));
        knn_on_device_ = true;
    }
    const Sample& negSample(int64_t linear) const {
        for (const std::vector<Sample>& list : hard_negs_) {
            if (linear < (int64_t)list.size()) return list[linear];
            linear -= (int64_t)list.size();
        }
        throw std::runtime_error("hard negative index out of range");
    }
    pe_miner* miner_ = nullptr;
    std::vector<std::vector<Sample>> hard_negs_;
    bool inplane_, knn_on_device_ = false;
    int device_;
    size_t img_ = 0;
    uint64_t batches_built_ = 0;
};

}  // namespace sz

typedef sz::networkEvaluator eval;
