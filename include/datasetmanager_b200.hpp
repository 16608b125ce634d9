/*
 * datasetmanager_b200.hpp - C++ host mirror of the reference's SAMPLE PRODUCERS (the callers that feed
 * the descriptor path), on top of the C ABI in pose_b200.h.  Header-only, C++17, no OpenCV / Eigen / OpenGL.
 *
 * Mirrors, with the same names, argument order and meaning:
 *   sz::Model::loadModel                       include/model.h:49, src/model.cpp:295-352 (PLY; OBJ is not read)
 *   sz::SphereRenderer::createViews            include/sphere.h:44, src/sphere.cpp:69-94 (clip = false only)
 *   sz::datasetManager::createSynthSamplesWadim / createSynthSamplesPaul     src/datasetmanager.cpp:498-535, 425-496
 *   sz::datasetManager::extractRealSamplesPaul / extractRealSamplesWadim     src/datasetmanager.cpp:348-411
 *   sz::Frame                                  include/datatypes.h:18-23 (colour, depth and gt only)
 *   sz::hdf5Handler::read / write              include/hdf5handler.h:15-16 (on PESS0001 files, not HDF5)
 *   sz::datasetManager(config), generateDatasets, getTemplateSet / getTrainingSet / getTestSet
 *                                              src/datasetmanager.cpp:802-831, 602-676, include/datasetmanager.h:59-61
 * Substitutions: Eigen::Matrix3f -> sz::Matrix3f (row-major 9 floats, cam(r, c)); Eigen::Isometry3f -> sz::Isometry3f
 * (row-major 3x4 [R|t]); cv::Mat frames -> flat arrays in the layout pe_sample_patches reads.  Rendering, normals, patch
 * cut and labels all run inside the C ABI (pe_template_samples / pe_scene_samples): rendered frames never reach the host.
 * The datasetManager mirror holds what the producers and the evaluator read; the benchmark loaders (image files through
 * OpenCV), generateAndStoreSamples, the background set and the hard negatives are outside this path.
 */
#pragma once
#include <array>
#include <limits>
#include <random>
#include <utility>

#include "networkevaluator_b200.hpp"

namespace sz {

/* camera intrinsics, row-major: cam(0,0) = fx, cam(1,1) = fy, cam(0,2) = ox, cam(1,2) = oy */
struct Matrix3f {
    std::array<float, 9> v{};
    float& operator()(int r, int c) { return v[3 * r + c]; }
    float operator()(int r, int c) const { return v[3 * r + c]; }
    std::array<float, 4> cam4() const { return {v[0], v[4], v[2], v[5]}; }
};

/* rigid transform, row-major 3x4 [R|t] (the layout of every `poses` argument in pose_b200.h) */
struct Isometry3f {
    std::array<float, 12> m{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    float& operator()(int r, int c) { return m[4 * r + c]; }
    float operator()(int r, int c) const { return m[4 * r + c]; }
    Isometry3f inverse() const { /* [R^T | -R^T t] */
        Isometry3f o;
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) o(r, c) = (*this)(c, r);
            o(r, 3) = -((*this)(0, r) * (*this)(0, 3) + (*this)(1, r) * (*this)(1, 3) + (*this)(2, r) * (*this)(2, 3));
        }
        return o;
    }
};

using Instance = std::pair<std::string, Isometry3f>;

/* include/datatypes.h:18-23: colour u8 BGR [rows][cols][3], metric depth f32 [rows][cols], gt[0] = the object's pose */
struct Frame {
    int nr = 0, rows = 0, cols = 0;
    std::vector<uint8_t> color;
    std::vector<float> depth;
    std::vector<Instance> gt;
};

/* the data members of Model the producers read (include/model.h:40-62); m_tex is the texture as glTexImage2D receives it */
class Model {
public:
    bool loadModel(const std::string& filename, int type = 1) {
        if (type != 1) throw std::runtime_error("Model::loadModel: only PLY (type 1) is read");
        pe_mesh* h = nullptr;
        if (pe_mesh_load_ply(filename.c_str(), &h) != PE_OK) return false;       /* the reference returns false too */
        int nv = 0, nf = 0, hc = 0, hu = 0;
        pe_mesh_info(h, &nv, &nf, &hc, &hu, centroid.data());
        m_points.resize((size_t)nv * 3);
        m_colors.resize((size_t)nv * 3);
        m_tcoords.resize(hu ? (size_t)nv * 2 : 0);
        m_faces.resize((size_t)nf * 3);
        const int st = pe_mesh_data(h, m_points.data(), m_colors.data(), hu ? m_tcoords.data() : nullptr, m_faces.data());
        pe_mesh_destroy(h);
        pe_check(st);
        return true;
    }
    void setTexture(const uint8_t* bgr, int rows, int cols) { m_tex.assign(bgr, bgr + (size_t)rows * cols * 3); tex_rows = rows; tex_cols = cols; }
    int numPoints() const { return (int)(m_points.size() / 3); }
    int numFaces() const { return (int)(m_faces.size() / 3); }
    std::vector<float>& getPoints() { return m_points; }
    std::vector<float>& getColors() { return m_colors; }
    std::vector<int32_t>& getFaces() { return m_faces; }
    bool textured() const { return !m_tcoords.empty() && !m_tex.empty(); }

    std::array<float, 3> centroid{};
    std::vector<float> m_points, m_colors, m_tcoords;      /* xyz, rgb in [0,1], uv per vertex */
    std::vector<int32_t> m_faces;
    std::vector<uint8_t> m_tex;
    int tex_rows = 0, tex_cols = 0;
};

/* include/sphere.h:18-26; col / dep are filled only when createViews is asked to keep the frames */
struct RenderView {
    Isometry3f pose;
    std::vector<uint8_t> col;
    std::vector<float> dep;
    int x_off = 0, y_off = 0;
};

struct Vector3f {
    float x = 0, y = 0, z = 0;
    Vector3f() = default;
    Vector3f(float a, float b, float c) : x(a), y(b), z(c) {}
};

constexpr int kFrameRows = 480, kFrameCols = 640;          /* SingletonPainter(CAM_NEAR, CAM_FAR, 640, 480), src/painter.cpp:216 */
constexpr float kCamNear = 0.001f, kCamFar = 4.0f;

class SphereRenderer {
public:
    SphereRenderer() = default;
    explicit SphereRenderer(const Matrix3f& cam) : m_camera(cam) {}
    void init(const Matrix3f& cam) { m_camera = cam; }

    /* the pose list alone (src/sphere.cpp:74-94 without renderView) */
    static std::vector<Isometry3f> createPoses(int sphereDep, Vector3f scale, Vector3f rotation, bool skipLowerHemi = true) {
        const float sc[3] = {scale.x, scale.y, scale.z}, ro[3] = {rotation.x, rotation.y, rotation.z};
        int64_t n = 0;
        pe_check(pe_sphere_views(sphereDep, sc, ro, skipLowerHemi ? 1 : 0, nullptr, 0, &n));
        std::vector<Isometry3f> poses((size_t)n);
        if (n) pe_check(pe_sphere_views(sphereDep, sc, ro, skipLowerHemi ? 1 : 0, poses[0].m.data(), n, &n));
        return poses;
    }

    /* src/sphere.cpp:69-94.  clip = true (a view cut to the projected bounding box) is not offered: the producers pass
     * false (src/datasetmanager.cpp:510).  Frames come back to the host here; createSynthSamples* do not go through this. */
    std::vector<RenderView> createViews(Model& model, int sphereDep, Vector3f scale, Vector3f rotation, bool skipLowerHemi = true,
                                        bool clip = true, int device = 0) const {
        if (m_camera(0, 0) == 0.f) throw std::runtime_error("SphereRenderer: camera not set");
        if (clip) throw std::runtime_error("SphereRenderer::createViews: clip = true is not supported, pass false");
        const std::vector<Isometry3f> poses = createPoses(sphereDep, scale, rotation, skipLowerHemi);
        std::vector<RenderView> out(poses.size());
        const size_t px = (size_t)kFrameRows * kFrameCols;
        const auto cam = m_camera.cam4();
        const bool tex = model.textured();
        constexpr size_t kChunk = 64;
        std::vector<uint8_t> col(kChunk * px * 3);
        std::vector<float> dep(kChunk * px);
        for (size_t first = 0; first < poses.size(); first += kChunk) {
            const size_t n = std::min(kChunk, poses.size() - first);
            pe_check(pe_render_views_tex(device, model.m_points.data(), model.m_colors.data(), tex ? model.m_tcoords.data() : nullptr,
                                         model.numPoints(), model.m_faces.data(), model.numFaces(), tex ? model.m_tex.data() : nullptr,
                                         tex ? model.tex_rows : 0, tex ? model.tex_cols : 0, cam.data(), poses[first].m.data(), (int)n,
                                         kFrameRows, kFrameCols, kCamNear, kCamFar, col.data(), dep.data()));
            for (size_t i = 0; i < n; ++i) {
                RenderView& v = out[first + i];
                v.pose = poses[first + i];
                v.col.assign(col.begin() + i * px * 3, col.begin() + (i + 1) * px * 3);
                v.dep.assign(dep.begin() + i * px, dep.begin() + (i + 1) * px);
            }
        }
        std::clog << "    - sphereRenderer - rendered " << out.size() << " views" << std::endl;
        return out;
    }

private:
    Matrix3f m_camera;
};

static_assert(sizeof(Isometry3f) == 12 * sizeof(float), "poses are passed to the C ABI as one flat array");

/* hdf5Handler::read / write (include/hdf5handler.h:15-16, src/hdf5handler.cpp:11-132) on the flat PESS0001 files of
 * pe_samples_read / pe_samples_write: the same two arrays as the reference's HDF5 datasets ("data" [n][C][H][W], "label"
 * [n][8]); tools/h5_to_pess.py converts a reference .h5 file.  HDF5 itself is not linked. */
class hdf5Handler {
public:
    std::vector<Sample> read(const std::string& filename) const {
        int64_t n = 0;
        int c = 0, h = 0, w = 0, ld = 0;
        pe_check(pe_samples_info(filename.c_str(), &n, &c, &h, &w, &ld));
        std::vector<Sample> samples((size_t)n);
        if (n == 0) return samples;
        const size_t img = (size_t)h * w * c;
        std::vector<float> data((size_t)n * img), labels((size_t)n * ld);
        pe_check(pe_samples_read(filename.c_str(), 0, n, data.data(), labels.data()));
        for (int64_t i = 0; i < n; ++i) {
            samples[i].data = Mat(h, w, c);
            std::memcpy(samples[i].data.ptr<float>(0), data.data() + i * img, img * sizeof(float));
            samples[i].label = Mat(1, ld);
            std::memcpy(samples[i].label.ptr<float>(0), labels.data() + (size_t)i * ld, ld * sizeof(float));
        }
        return samples;
    }
    void write(const std::string& filename, std::vector<Sample>& samples) const {
        if (samples.empty()) throw std::runtime_error("hdf5Handler::write: no samples for " + filename);
        const Mat& d0 = samples[0].data;
        const int ld = samples[0].label.cols;
        const size_t img = (size_t)d0.rows * d0.cols * d0.ch;
        std::vector<float> data(samples.size() * img), labels(samples.size() * ld);
        for (size_t i = 0; i < samples.size(); ++i) {
            const Sample& s = samples[i];
            if (s.data.rows != d0.rows || s.data.cols != d0.cols || s.data.ch != d0.ch || s.label.cols != ld)
                throw std::runtime_error("hdf5Handler::write: sample " + std::to_string(i) + " has another shape");
            std::memcpy(data.data() + i * img, s.data.buf.data(), img * sizeof(float));
            std::memcpy(labels.data() + i * ld, s.label.buf.data(), ld * sizeof(float));
        }
        pe_check(pe_samples_write(filename.c_str(), data.data(), labels.data(), (int64_t)samples.size(), d0.ch, d0.rows, d0.cols, ld));
    }
};

class datasetManager : public DatasetView {
public:
    std::string dataset_path, hdf5_path, dataset_name = "LineMOD";
    std::string sample_ext = ".pess";                   /* the reference's ".h5" files, converted by tools/h5_to_pess.py */
    bool inplane = false, use_real = true;
    int random_background = 0;
    std::vector<std::string> models, used_models;
    std::vector<int> rotInv;
    std::unordered_map<std::string, int> model_index;
    uint64_t shuffle_seed = 5489;                       /* the reference shuffles with the C library's rand(): order is not portable */
    int device = 0;
    float patch_metric_size = 0.2f;                     /* samplePatchWithScale's m, src/datasetmanager.cpp:302-306 */
    hdf5Handler h5;

    datasetManager() = default;
    /* src/datasetmanager.cpp:802-831: the keys of config.ini this mirror uses (background_path and the simulated-data keys
     * belong to parts that are not mirrored and are not required) */
    explicit datasetManager(const std::string& config) {
        IniFile pt(config);
        dataset_path = pt.get<std::string>("paths.dataset_path");
        hdf5_path = pt.get<std::string>("paths.hdf5_path");
        dataset_name = pt.get<std::string>("input.dataset_name");
        random_background = pt.get<int>("input.random_background");
        use_real = pt.get<bool>("input.use_real");
        inplane = pt.get<bool>("input.inplane");
        models = to_array<std::string>(pt.get<std::string>("input.models"));
        used_models = to_array<std::string>(pt.get<std::string>("input.used_models"));
        rotInv = to_array<int>(pt.get<std::string>("input.rotInv"));
        if (dataset_name != "LineMOD" && dataset_name != "BigBIRD" && dataset_name != "Washington")
            throw std::runtime_error("Unknown dataset: " + dataset_name + "!");
        for (size_t i = 0; i < models.size(); ++i) model_index[models[i]] = (int)i;
    }

    const SampleSet& getTrainingSet() const override { return training_set; }
    const SampleSet& getTemplateSet() const override { return template_set; }
    const SampleSet& getTestSet() const override { return test_set; }
    const std::vector<std::string>& getModels() const { return used_models; }
    std::string getDatasetPath() const { return dataset_path; }
    std::string getHDF5Path() const { return hdf5_path; }

    /* src/datasetmanager.cpp:602-676 without generateAndStoreSamples (the three files per model must exist), the
     * background set and the hard negatives: per used model read realSamples_ / synthSamples_ / templates_, hang every real
     * sample on its closest template (angularDistance, first minimum), send the first ceil(n / 2) of each template's
     * samples to the training set (when use_real) and the rest to the test set, crop every object to the smallest count,
     * shuffle.  maxSimTmpl is the miner's business (pe_miner_max_sim). */
    void generateDatasets() {
        training_set.clear();
        template_set.clear();
        test_set.clear();
        for (const std::string& seq : used_models) {
            std::vector<Sample> train_real = h5.read(hdf5_path + "realSamples_" + seq + sample_ext);
            std::vector<Sample> train_synth = h5.read(hdf5_path + "synthSamples_" + seq + sample_ext);
            template_set.push_back(h5.read(hdf5_path + "templates_" + seq + sample_ext));
            test_set.push_back(std::vector<Sample>());
            const size_t nr_template_poses = template_set[0].size(), nr_real_poses = train_real.size();
            if (template_set.back().size() < nr_template_poses)
                throw std::runtime_error("templates_" + seq + ": fewer templates than the first model");
            std::vector<std::vector<int>> maxSimTmpl(nr_template_poses);
            for (size_t real_sample = 0; real_sample < nr_real_poses; ++real_sample) {
                float best_dist = std::numeric_limits<float>::max(), q[4];
                size_t sim_tmpl = 0;
                train_real[real_sample].getQuat(q);
                for (size_t tmpl = 0; tmpl < nr_template_poses; ++tmpl) {
                    float t[4];
                    template_set.back()[tmpl].getQuat(t);
                    const float temp_dist = angularDistance(q, t);
                    if (temp_dist >= best_dist) continue;
                    best_dist = temp_dist;
                    sim_tmpl = tmpl;
                }
                if (nr_template_poses) maxSimTmpl[sim_tmpl].push_back((int)real_sample);
            }
            for (size_t tmpl = 0; tmpl < nr_template_poses; ++tmpl) {
                const std::vector<int>& of = maxSimTmpl[tmpl];
                const size_t half = (size_t)std::ceil(of.size() / 2.0);
                if (use_real)
                    for (size_t i = 0; i < half; ++i) train_synth.push_back(train_real[of[i]]);
                for (size_t i = half; i < of.size(); ++i) test_set.back().push_back(train_real[of[i]]);
            }
            training_set.push_back(std::move(train_synth));
        }
        size_t min_training = std::numeric_limits<size_t>::max(), min_test = min_training;
        for (size_t object = 0; object < used_models.size(); ++object) {
            min_training = std::min(min_training, training_set[object].size());
            min_test = std::min(min_test, test_set[object].size());
        }
        std::mt19937_64 rng(shuffle_seed);
        for (size_t object = 0; object < used_models.size(); ++object) {
            training_set[object].resize(min_training);
            test_set[object].resize(min_test);
            std::shuffle(training_set[object].begin(), training_set[object].end(), rng);
            std::shuffle(test_set[object].begin(), test_set[object].end(), rng);
        }
    }

    /* Eigen 3.2 Quaternionf::angularDistance as DESIGN 2 states it: 2 acos(min(1, |a.b|)) */
    static float angularDistance(const float a[4], const float b[4]) {
        float d = std::fabs(a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3]);
        if (d > 1.f) d = 1.f;
        return 2.f * std::acos(d);
    }

    /* src/datasetmanager.cpp:498-535 */
    std::vector<Sample> createSynthSamplesWadim(Model& model, Matrix3f& cam, int index, int subdiv) const {
        const Vector3f scales(0.4f, 1.1f, 1.0f);
        const Vector3f in_plane_rots = inplane ? Vector3f(-45, 15, 45) : Vector3f(0, 15, 10);
        const std::vector<Isometry3f> poses = SphereRenderer::createPoses(subdiv, scales, in_plane_rots, true);
        return synth(model, cam, index, poses, nullptr);
    }

    /* src/datasetmanager.cpp:425-496: camera positions (elevation, azimuth) from paul/camPositionsElAz.txt */
    std::vector<Sample> createSynthSamplesPaul(Model& model, Matrix3f& cam, int index) const {
        std::ifstream file(dataset_path + "paul/camPositionsElAz.txt");
        if (!file.good()) throw std::runtime_error("cannot open " + dataset_path + "paul/camPositionsElAz.txt");
        const float tz = dataset_name == "BigBIRD" ? 0.72f : 0.4f;
        std::vector<Isometry3f> poses;
        std::vector<float> z;
        float el, az;
        while (poses.size() < 1542 && (file >> el >> az)) {
            /* camRot0 * (Rx(el) * Rz(-az)), camRot0 = Rx(pi/2) */
            const float r0[9] = {1, 0, 0, 0, std::cos((float)M_PI / 2), -std::sin((float)M_PI / 2), 0, std::sin((float)M_PI / 2),
                                 std::cos((float)M_PI / 2)};
            const float ce = std::cos(el), se = std::sin(el), ca = std::cos(-az), sa = std::sin(-az);
            const float rx[9] = {1, 0, 0, 0, ce, -se, 0, se, ce}, rz[9] = {ca, -sa, 0, sa, ca, 0, 0, 0, 1};
            float a[9], b[9];
            mul3(rx, rz, a);
            mul3(r0, a, b);
            Isometry3f pose;
            for (int r = 0; r < 3; ++r) {
                for (int c = 0; c < 3; ++c) pose(r, c) = b[3 * r + c];
                pose(r, 3) = r == 2 ? tz : 0.f;
            }
            if (dataset_name == "BigBIRD") { /* pose_object = pose^-1; translation += centroid; pose = pose_object^-1 */
                Isometry3f po = pose.inverse();
                for (int r = 0; r < 3; ++r) po(r, 3) += model.centroid[r];
                pose = po.inverse();
            }
            poses.push_back(pose);
            z.push_back(tz);                            /* the patch is cut at trans(2), not at the adapted pose's z (:478) */
        }
        return synth(model, cam, index, poses, z.data());
    }

    /* src/datasetmanager.cpp:348-379 */
    std::vector<Sample> extractRealSamplesPaul(std::vector<Frame>& frames, Matrix3f& cam, int index, Model& model) const {
        return real(frames, cam, index, dataset_name == "BigBIRD" ? model.centroid.data() : nullptr, 0);
    }
    /* src/datasetmanager.cpp:381-411 */
    std::vector<Sample> extractRealSamplesWadim(std::vector<Frame>& frames, Matrix3f& cam, int index) const {
        return real(frames, cam, index, nullptr, 1);
    }

private:
    SampleSet template_set, training_set, test_set;
    static void mul3(const float* a, const float* b, float* o) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) o[3 * r + c] = a[3 * r] * b[c] + a[3 * r + 1] * b[3 + c] + a[3 * r + 2] * b[6 + c];
    }
    static std::vector<Sample> to_samples(const std::vector<float>& patches, const std::vector<float>& labels, size_t n) {
        std::vector<Sample> samples(n);
        const size_t img = (size_t)64 * 64 * 7;
        for (size_t i = 0; i < n; ++i) {
            samples[i].data = Mat(64, 64, 7);
            std::memcpy(samples[i].data.ptr<float>(0), patches.data() + i * img, img * sizeof(float));
            samples[i].label = Mat(1, 8);
            std::memcpy(samples[i].label.ptr<float>(0), labels.data() + i * 8, 8 * sizeof(float));
        }
        return samples;
    }
    std::vector<Sample> synth(Model& model, const Matrix3f& cam, int index, const std::vector<Isometry3f>& poses, const float* z) const {
        const size_t n = poses.size();
        if (!n) return {};
        std::vector<float> patches(n * 64 * 64 * 7), labels(n * 8);
        std::vector<int32_t> status(n);
        const auto c4 = cam.cam4();
        const bool tex = model.textured();
        pe_check(pe_template_samples(device, model.m_points.data(), model.m_colors.data(), tex ? model.m_tcoords.data() : nullptr,
                                     model.numPoints(), model.m_faces.data(), model.numFaces(), tex ? model.m_tex.data() : nullptr,
                                     tex ? model.tex_rows : 0, tex ? model.tex_cols : 0, c4.data(), poses[0].m.data(), (int64_t)n, index, z,
                                     kFrameRows, kFrameCols, kCamNear, kCamFar, patch_metric_size, patches.data(), labels.data(),
                                     status.data()));
        check_status(status);
        return to_samples(patches, labels, n);
    }
    std::vector<Sample> real(std::vector<Frame>& frames, const Matrix3f& cam, int index, const float* centroid3, int z_from_depth) const {
        const size_t n = frames.size();
        if (!n) return {};
        const int rows = frames[0].rows, cols = frames[0].cols;
        const size_t px = (size_t)rows * cols;
        std::vector<uint8_t> color(n * px * 3);
        std::vector<float> depth(n * px), poses(n * 12);
        for (size_t i = 0; i < n; ++i) {
            const Frame& f = frames[i];
            if (f.rows != rows || f.cols != cols || f.color.size() != px * 3 || f.depth.size() != px || f.gt.empty())
                throw std::runtime_error("extractRealSamples: frame " + std::to_string(i) + " has another size or no ground truth");
            std::memcpy(color.data() + i * px * 3, f.color.data(), px * 3);
            std::memcpy(depth.data() + i * px, f.depth.data(), px * sizeof(float));
            std::memcpy(poses.data() + i * 12, f.gt[0].second.m.data(), 12 * sizeof(float));
        }
        std::vector<float> patches(n * 64 * 64 * 7), labels(n * 8);
        std::vector<int32_t> status(n);
        const auto c4 = cam.cam4();
        pe_check(pe_scene_samples(device, color.data(), depth.data(), (int)n, rows, cols, c4.data(), poses.data(), centroid3, z_from_depth,
                                  index, patch_metric_size, patches.data(), labels.data(), status.data()));
        check_status(status);
        return to_samples(patches, labels, n);
    }
    /* the reference asserts that the clamped cut rectangle lies inside the frame (src/datasetmanager.cpp:317-318) */
    static void check_status(const std::vector<int32_t>& status) {
        for (size_t i = 0; i < status.size(); ++i)
            if (status[i] != 0) throw std::runtime_error("samplePatchWithScale: the patch of sample " + std::to_string(i) + " leaves the frame");
    }
};

}  // namespace sz
