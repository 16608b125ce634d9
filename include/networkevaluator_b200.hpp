/*
 * networkevaluator_b200.hpp - C++ host mirror of the reference's evaluation interface, on top of
 * the C ABI in pose_b200.h.  Header-only, C++17, no OpenCV / Caffe / Eigen / Boost.
 *
 * Mirrors, with the same names, argument order and meaning:
 *   sz::networkEvaluator   include/networkevaluator.h:25-41 (static methods; the two GUI helpers
 *                          computeManifold / visualizeKNN need a window and are left out)
 *   sz::networkSolver::computeKNN   include/networksolver.h:49, src/networksolver.cpp:456-505
 *   sz::Sample             include/datatypes.h:25-49
 * Substitutions: caffe::Net<float>& -> sz::Net& (handle on a pe_net); cv::Mat -> sz::Mat (row-major
 * float matrix with the few members the path uses); datasetManager& -> sz::DatasetView (the three
 * getters the evaluator calls, include/datasetmanager.h:59-61); boost::property_tree INI -> a small
 * reader of the same `[section] key = value` files.  Errors surface as std::runtime_error carrying
 * pe_last_error().
 */
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "pose_b200.h"

namespace sz {

inline void pe_check(int status) {
    if (status != PE_OK) throw std::runtime_error(std::string("pose_b200: ") + pe_last_error());
}

/* minimal stand-in for cv::Mat of type CV_32FC(ch): rows x cols x ch floats, row-major, owning */
struct Mat {
    int rows = 0, cols = 0, ch = 1;
    std::vector<float> buf;
    Mat() = default;
    Mat(int r, int c, int channels = 1) : rows(r), cols(c), ch(channels), buf((size_t)r * c * channels, 0.f) {}
    int channels() const { return ch; }
    bool empty() const { return buf.empty(); }
    template <typename T = float>
    T* ptr(int r = 0) { return reinterpret_cast<T*>(buf.data() + (size_t)r * cols * ch); }
    template <typename T = float>
    const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(buf.data() + (size_t)r * cols * ch); }
    template <typename T = float>
    T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T = float>
    const T& at(int r, int c) const { return ptr<T>(r)[c]; }
    /* cv::Mat::push_back(const Mat&): append rows */
    void push_back(const Mat& m) {
        if (m.rows == 0) return;
        if (rows == 0) { cols = m.cols; ch = m.ch; }
        if (m.cols != cols || m.ch != ch) throw std::runtime_error("Mat::push_back: shape mismatch");
        buf.insert(buf.end(), m.buf.begin(), m.buf.end());
        rows += m.rows;
    }
};

/* include/datatypes.h:25-49; the quaternion is kept in Eigen coeffs() order x,y,z,w */
struct Sample {
    Mat data, label;
    void copySample(const Sample& s) { data = s.data; label = s.label; }
    void getQuat(float q[4]) const { for (int j = 0; j < 4; ++j) q[j] = label.at<float>(0, 1 + j); }
    void getTrans(float t[3]) const { for (int j = 0; j < 3; ++j) t[j] = label.at<float>(0, 5 + j); }
};

using SampleSet = std::vector<std::vector<Sample>>;

/* the part of datasetManager the evaluator uses (include/datasetmanager.h:59-61) */
struct DatasetView {
    virtual ~DatasetView() = default;
    virtual const SampleSet& getTemplateSet() const = 0;
    virtual const SampleSet& getTrainingSet() const = 0;
    virtual const SampleSet& getTestSet() const = 0;
};

/* stands in for caffe::Net<float>(prototxt, caffe::TEST) (src/networksolver.cpp:286) */
class Net {
public:
    Net(const std::string& prototxt_path, int device = 0) { pe_check(pe_net_create_from_prototxt(prototxt_path.c_str(), device, &h_)); }
    static Net named(const std::string& variant, int device = 0) {
        pe_net* h = nullptr;
        pe_check(pe_net_create_named(variant.c_str(), device, &h));
        return Net(h, true);
    }
    static Net borrow(pe_net* h) { return Net(h, false); }
    Net(Net&& o) noexcept : h_(o.h_), own_(o.own_) { o.h_ = nullptr; }
    Net(const Net&) = delete;
    Net& operator=(const Net&) = delete;
    ~Net() { if (h_ && own_) pe_net_destroy(h_); }
    pe_net* handle() const { return h_; }
    void fill(uint64_t seed, int mode) { pe_check(pe_net_fill(h_, seed, mode)); }
    void loadWeights(const std::string& path) { pe_check(pe_net_load_weights(h_, path.c_str())); }
    /* learnable blobs in Caffe order (net.params()): stands in for CopyTrainedLayersFrom /
     * ShareTrainedLayersWith (src/networksolver.cpp:307,377) */
    int numBlobs() const { int n = 0; pe_check(pe_net_num_blobs(h_, &n)); return n; }
    void setBlob(int blob, const float* data, int64_t count) { pe_check(pe_net_set_blob(h_, blob, data, count)); }
    /* CNN.input_blobs()[0] / output_blobs()[0] shapes (src/networkevaluator.cpp:11-19) */
    void shape(int* batch, int* channels, int* height, int* width, int* desc_dim) const {
        pe_check(pe_net_info(h_, batch, channels, height, width, desc_dim));
    }

private:
    Net(pe_net* h, bool own) : h_(h), own_(own) {}
    pe_net* h_ = nullptr;
    bool own_ = true;
};

/* include/helper.h to_array<T>: comma-separated list */
template <typename T>
std::vector<T> to_array(const std::string& s) {
    std::vector<T> out;
    std::stringstream ss(s);
    std::string item;
    while (std::getline(ss, item, ',')) {
        size_t a = item.find_first_not_of(" \t\r"), b = item.find_last_not_of(" \t\r");
        if (a == std::string::npos) continue;
        std::stringstream conv(item.substr(a, b - a + 1));
        T v;
        conv >> v;
        out.push_back(v);
    }
    return out;
}

/* reader for the reference's config files (config.ini): `[section]`, `key = value`, # and ; comments;
 * get() throws on a missing key like boost::property_tree::ptree::get<T>(path) */
class IniFile {
public:
    explicit IniFile(const std::string& path) {
        std::ifstream f(path);
        if (!f) throw std::runtime_error("cannot open config " + path);
        std::string line, section;
        while (std::getline(f, line)) {
            size_t a = line.find_first_not_of(" \t\r");
            if (a == std::string::npos || line[a] == '#' || line[a] == ';') continue;
            if (line[a] == '[') { section = line.substr(a + 1, line.find(']') - a - 1); continue; }
            size_t eq = line.find('=');
            if (eq == std::string::npos) continue;
            kv_[section + "." + trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
        }
    }
    template <typename T>
    T get(const std::string& path) const {
        auto it = kv_.find(path);
        if (it == kv_.end()) throw std::runtime_error("No such node (" + path + ")");
        if constexpr (std::is_same<T, std::string>::value) return it->second;
        else {
            std::stringstream ss(it->second);
            T v;
            ss >> std::boolalpha >> v;
            if (ss.fail()) { std::stringstream s2(it->second); s2 >> v; }
            return v;
        }
    }

private:
    static std::string trim(const std::string& s) {
        size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r");
        return a == std::string::npos ? "" : s.substr(a, b - a + 1);
    }
    std::map<std::string, std::string> kv_;
};

}  // namespace sz

#include "networkevaluator_b200_impl.hpp"
