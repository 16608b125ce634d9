/*
 * api.cu - the extern "C" boundary declared in include/pose_b200.h (everything except the
 * multi-GPU entry points, which live in comm.cu).  Thin: argument checks, host<->device
 * plumbing, and calls into forward.cu / knn.cu / pose.cu / synth.cu.
 */
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>

#include "common.h"

namespace pe {

std::atomic<int64_t> g_launches{0};
std::atomic<int> g_opt_knn_tensor{1}, g_opt_forward_tensor{1}, g_opt_forward_fp8{0}, g_opt_p2p{0}, g_opt_conv2_tensor{1}, g_opt_knn_certify{1};
static thread_local std::string t_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    t_error = buf;
}

int DevBuf::reserve(size_t bytes, bool zero_new) {
    if (bytes <= cap && p) return PE_OK;
    if (bytes == 0) bytes = 16;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    PE_CUDA(cudaMalloc(&p, bytes));
    cap = bytes;
    if (zero_new) PE_CUDA(cudaMemset(p, 0, bytes));
    return PE_OK;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}

int prof_mark(bool on, std::vector<cudaEvent_t>* events, cudaStream_t stream) {
    if (!on) return PE_OK;
    cudaEvent_t e;
    PE_CUDA(cudaEventCreate(&e));
    PE_CUDA(cudaEventRecord(e, stream));
    events->push_back(e);
    return PE_OK;
}

/* sums elapsed times of consecutive events in groups of `group`, destroys them */
static int prof_collect(std::vector<cudaEvent_t>* events, int group, double* ms, int64_t* calls) {
    for (int i = 0; i < group - 1; ++i) ms[i] = 0.0;
    *calls = 0;
    const size_t n_groups = events->size() / group;
    for (size_t g = 0; g < n_groups; ++g) {
        PE_CUDA(cudaEventSynchronize((*events)[g * group + group - 1]));
        for (int i = 0; i + 1 < group; ++i) {
            float t = 0.f;
            PE_CUDA(cudaEventElapsedTime(&t, (*events)[g * group + i], (*events)[g * group + i + 1]));
            ms[i] += t;
        }
        ++*calls;
    }
    for (cudaEvent_t e : *events) cudaEventDestroy(e);
    events->clear();
    return PE_OK;
}

static int set_device_checked(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error("no CUDA device available (%s); this library has no CPU fallback",
                  e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return PE_ERR_CUDA;
    }
    if (device < 0 || device >= n) { set_error("device %d outside 0..%d", device, n - 1); return PE_ERR_INVALID; }
    PE_CUDA(cudaSetDevice(device));
    return PE_OK;
}

int net_ensure_device(pe_net* net) {
    if (net->stream) { PE_CUDA(cudaSetDevice(net->device)); return PE_OK; }
    PE_CHECK(set_device_checked(net->device));
    PE_CUDA(cudaDeviceGetAttribute(&net->num_sms, cudaDevAttrMultiProcessorCount, net->device));
    PE_CUDA(cudaStreamCreateWithFlags(&net->stream, cudaStreamNonBlocking));
    PE_CUDA(cudaStreamCreateWithFlags(&net->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        PE_CUDA(cudaEventCreateWithFlags(&net->ev_in[i], cudaEventDisableTiming));
        PE_CUDA(cudaEventCreateWithFlags(&net->ev_done[i], cudaEventDisableTiming));
    }
    return PE_OK;
}

/* match the layer list against conv(16,8) pool(2,2) [act] conv(7,5) pool(2,2) [act] fc ([act] fc)* [act] */
static int plan_net(pe_net* net) {
    const NetDef& d = net->def;
    auto unsupported = [](const char* why) { set_error("net outside the manifold family: %s", why); return PE_ERR_UNSUPPORTED; };
    if (d.height != 64 || d.width != 64) return unsupported("input must be 64x64");
    if (d.channels < 1 || d.channels > 8) return unsupported("1..8 input channels");
    if (d.sel.empty() || d.sel.size() > 8) return unsupported("1..8 conv1 channels");
    size_t i = 0;
    const auto& L = d.layers;
    auto is = [&](size_t at, int kind) { return at < L.size() && L[at].kind == kind; };
    if (!is(i, L_CONV) || L[i].a != 16 || L[i].b != 8) return unsupported("conv1 must be 16 filters of 8x8");
    net->blob_w1 = L[i].blob0; net->blob_b1 = L[i].blob1; ++i;
    if (!is(i, L_POOL) || L[i].a != 2 || L[i].b != 2) return unsupported("pool1 must be 2x2 stride 2");
    ++i;
    net->act1 = PE_ACT_NONE;
    if (is(i, L_ACT)) { net->act1 = L[i].a; net->blob_s1 = L[i].blob0; ++i; }
    if (!is(i, L_CONV) || L[i].a != 7 || L[i].b != 5) return unsupported("conv2 must be 7 filters of 5x5");
    net->blob_w2 = L[i].blob0; net->blob_b2 = L[i].blob1; ++i;
    if (!is(i, L_POOL) || L[i].a != 2 || L[i].b != 2) return unsupported("pool2 must be 2x2 stride 2");
    ++i;
    net->act2 = PE_ACT_NONE;
    if (is(i, L_ACT)) { net->act2 = L[i].a; net->blob_s2 = L[i].blob0; ++i; }
    if (!is(i, L_FC)) return unsupported("an InnerProduct must follow pool2");
    int in = 7 * 13 * 13;
    net->fcs.clear();
    while (i < L.size()) {
        if (!is(i, L_FC)) return unsupported("only InnerProduct/activation layers may follow pool2");
        FcPlan fc;
        fc.in = in; fc.out = L[i].a; fc.ld_in = (in + 3) & ~3;
        fc.blob_w = L[i].blob0; fc.blob_b = L[i].blob1; ++i;
        if (is(i, L_ACT)) { fc.act = L[i].a; fc.blob_slope = L[i].blob0; ++i; }
        in = fc.out;
        net->fcs.push_back(fc);
    }
    if (d.desc_dim != in) return unsupported("descriptor dim mismatch");
    net->cin = (int)d.sel.size();
    return PE_OK;
}

static int create_net(const std::string& text, int device, pe_net** out) {
    PE_REQUIRE(out != nullptr, "null output pointer");
    *out = nullptr;
    pe_net* net = new pe_net();
    net->device = device;
    int st = parse_prototxt(text, &net->def);
    if (st == PE_OK) st = plan_net(net);
    if (st != PE_OK) { delete net; return st; }
    net->blobs.resize(net->def.blob_names.size());
    for (size_t b = 0; b < net->blobs.size(); ++b) {
        int64_t cnt = 1;
        for (auto v : net->def.blob_shapes[b]) cnt *= v;
        net->blobs[b].assign((size_t)cnt, 0.f);
    }
    *out = net;
    return PE_OK;
}

static float uniform_f32(uint64_t key, uint64_t ctr, double lo, double hi) {
    const float u = (float)(splitmix64(key + ctr) >> 40) * 5.9604644775390625e-8f;
    const float flo = (float)lo, span = (float)(hi - lo);
    volatile float prod = u * span; /* separately rounded, as numpy evaluates it */
    return flo + prod;
}

}  // namespace pe

using namespace pe;

struct pe_timer {
    int device = 0;
    cudaEvent_t a = nullptr, b = nullptr;
};

extern "C" {

int pe_version(void) { return 100; }
const char* pe_last_error(void) { return t_error.c_str(); }
int64_t pe_launch_count(void) { return g_launches.load(); }
int pe_set_option(const char* name, int value) {
    PE_REQUIRE(name != nullptr, "null option name");
    const std::string n(name);
    if (n == "knn_tensor") { g_opt_knn_tensor.store(value); return PE_OK; }
    if (n == "forward_tensor") { g_opt_forward_tensor.store(value); return PE_OK; }
    if (n == "forward_fp8") { g_opt_forward_fp8.store(value); return PE_OK; }
    if (n == "p2p") { g_opt_p2p.store(value); return PE_OK; }
    if (n == "conv2_tensor") { g_opt_conv2_tensor.store(value); return PE_OK; }
    if (n == "knn_certify") { g_opt_knn_certify.store(value); return PE_OK; }
    set_error("unknown option '%s'", name);
    return PE_ERR_INVALID;
}
int pe_get_option(const char* name, int* value) {
    PE_REQUIRE(name && value, "null pointer");
    const std::string n(name);
    if (n == "knn_tensor") { *value = g_opt_knn_tensor.load(); return PE_OK; }
    if (n == "forward_tensor") { *value = g_opt_forward_tensor.load(); return PE_OK; }
    if (n == "forward_fp8") { *value = g_opt_forward_fp8.load(); return PE_OK; }
    if (n == "p2p") { *value = g_opt_p2p.load(); return PE_OK; }
    if (n == "conv2_tensor") { *value = g_opt_conv2_tensor.load(); return PE_OK; }
    if (n == "knn_certify") { *value = g_opt_knn_certify.load(); return PE_OK; }
    set_error("unknown option '%s'", name);
    return PE_ERR_INVALID;
}

int pe_device_count(int* n) {
    PE_REQUIRE(n != nullptr, "null pointer");
    *n = 0;
    cudaError_t e = cudaGetDeviceCount(n);
    if (e != cudaSuccess) { cudaGetLastError(); *n = 0; set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return PE_ERR_CUDA; }
    return PE_OK;
}

int pe_host_alloc(void** p, size_t bytes) {
    PE_REQUIRE(p != nullptr, "null pointer");
    PE_CUDA(cudaMallocHost(p, bytes ? bytes : 16));
    return PE_OK;
}
int pe_host_free(void* p) { PE_CUDA(cudaFreeHost(p)); return PE_OK; }
int pe_dev_alloc(int device, void** p, size_t bytes) {
    PE_REQUIRE(p != nullptr, "null pointer");
    PE_CHECK(set_device_checked(device));
    PE_CUDA(cudaMalloc(p, bytes ? bytes : 16));
    return PE_OK;
}
int pe_dev_free(int device, void* p) {
    PE_CHECK(set_device_checked(device));
    PE_CUDA(cudaFree(p));
    return PE_OK;
}
int pe_memcpy_h2d(int device, void* dst, const void* src, size_t bytes) {
    PE_CHECK(set_device_checked(device));
    PE_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
    return PE_OK;
}
int pe_memcpy_d2h(int device, void* dst, const void* src, size_t bytes) {
    PE_CHECK(set_device_checked(device));
    PE_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return PE_OK;
}
int pe_dev_memset(int device, void* dst, int value, size_t bytes) {
    PE_CHECK(set_device_checked(device));
    PE_CUDA(cudaMemset(dst, value, bytes));
    return PE_OK;
}

/* ---------------------------------------------------------------- timers --- */
int pe_timer_create(int device, pe_timer** t) {
    PE_REQUIRE(t != nullptr, "null pointer");
    PE_CHECK(set_device_checked(device));
    pe_timer* x = new pe_timer();
    x->device = device;
    if (cudaEventCreate(&x->a) != cudaSuccess || cudaEventCreate(&x->b) != cudaSuccess) {
        delete x;
        set_error("cudaEventCreate failed");
        return PE_ERR_CUDA;
    }
    *t = x;
    return PE_OK;
}
int pe_timer_start_net(pe_timer* t, pe_net* net) {
    PE_REQUIRE(t && net, "null pointer");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaEventRecord(t->a, net->stream));
    return PE_OK;
}
int pe_timer_stop_net(pe_timer* t, pe_net* net) {
    PE_REQUIRE(t && net, "null pointer");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaEventRecord(t->b, net->stream));
    return PE_OK;
}
int pe_timer_start_db(pe_timer* t, pe_db* db) {
    PE_REQUIRE(t && db, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CUDA(cudaEventRecord(t->a, db->stream));
    return PE_OK;
}
int pe_timer_stop_db(pe_timer* t, pe_db* db) {
    PE_REQUIRE(t && db, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CUDA(cudaEventRecord(t->b, db->stream));
    return PE_OK;
}
int pe_timer_elapsed_ms(pe_timer* t, float* ms) {
    PE_REQUIRE(t && ms, "null pointer");
    PE_CUDA(cudaSetDevice(t->device));
    PE_CUDA(cudaEventSynchronize(t->b));
    PE_CUDA(cudaEventElapsedTime(ms, t->a, t->b));
    return PE_OK;
}
void pe_timer_destroy(pe_timer* t) {
    if (!t) return;
    cudaSetDevice(t->device);
    if (t->a) cudaEventDestroy(t->a);
    if (t->b) cudaEventDestroy(t->b);
    delete t;
}

/* ------------------------------------------------------------------- net --- */
int pe_net_create_from_string(const char* prototxt_text, int device, pe_net** out) {
    PE_REQUIRE(prototxt_text != nullptr, "null prototxt text");
    return create_net(prototxt_text, device, out);
}
int pe_net_create_from_prototxt(const char* path, int device, pe_net** out) {
    PE_REQUIRE(path != nullptr, "null path");
    std::ifstream f(path);
    if (!f) { set_error("cannot open %s", path); return PE_ERR_IO; }
    std::stringstream ss;
    ss << f.rdbuf();
    return create_net(ss.str(), device, out);
}
int pe_net_create_named(const char* variant, int device, pe_net** out) {
    PE_REQUIRE(variant != nullptr, "null variant");
    std::string text;
    if (!builtin_prototxt(variant, &text)) { set_error("unknown net variant `%s`", variant); return PE_ERR_INVALID; }
    return create_net(text, device, out);
}
void pe_net_destroy(pe_net* net) {
    if (!net) return;
    if (net->stream) {
        cudaSetDevice(net->device);
        cudaStreamSynchronize(net->stream);
        cudaStreamSynchronize(net->copy_stream);
        for (cudaEvent_t e : net->prof_events) cudaEventDestroy(e);
        for (float* p : {net->d_w1, net->d_b1, net->d_s1, net->d_w2, net->d_b2, net->d_s2})
            if (p) cudaFree(p);
        if (net->d_slot) cudaFree(net->d_slot);
        if (net->d_z1) cudaFree(net->d_z1);
        if (net->d_z2) cudaFree(net->d_z2);
        if (net->tc_err) cudaFreeHost(net->tc_err);
        for (auto& fc : net->fcs)
        {
            for (float* p : {fc.d_w, fc.d_b, fc.d_slope})
                if (p) cudaFree(p);
            if (fc.d_wp) cudaFree(fc.d_wp);
        }
        net->pool1.release(); net->p1_packed.release(); net->pool2.release(); net->fc_a.release(); net->fc_b.release();
        net->in_stage[0].release(); net->in_stage[1].release(); net->out_stage.release();
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(net->ev_in[i]); cudaEventDestroy(net->ev_done[i]); }
        cudaStreamDestroy(net->stream);
        cudaStreamDestroy(net->copy_stream);
    }
    delete net;
}

int pe_net_fp8_state(const pe_net* net, int* state) {
    PE_REQUIRE(net && state, "null pointer");
    const int opt = g_opt_forward_fp8.load();
    *state = opt >= 2 ? 1 : (opt == 0 ? 2 : net->fp8_state);
    return PE_OK;
}

int pe_net_info(const pe_net* net, int* batch, int* channels, int* height, int* width, int* desc_dim) {
    PE_REQUIRE(net != nullptr, "null net");
    if (batch) *batch = net->def.batch;
    if (channels) *channels = net->def.channels;
    if (height) *height = net->def.height;
    if (width) *width = net->def.width;
    if (desc_dim) *desc_dim = net->def.desc_dim;
    return PE_OK;
}
int pe_net_selected_channels(const pe_net* net, int* sel, int cap, int* n_sel) {
    PE_REQUIRE(net && n_sel, "null pointer");
    *n_sel = (int)net->def.sel.size();
    for (int i = 0; i < *n_sel && i < cap && sel; ++i) sel[i] = net->def.sel[i];
    return PE_OK;
}
int pe_net_describe(const pe_net* net, char* buf, size_t cap) {
    PE_REQUIRE(net && buf && cap > 0, "null pointer");
    static const char* act_names[] = {"none", "ReLU", "ELU", "PReLU"};
    std::ostringstream o;
    bool first = true;
    for (auto& l : net->def.layers) {
        if (!first) o << "\n";
        first = false;
        if (l.kind == L_CONV) o << "conv " << l.name << " " << l.a << " " << l.b;
        else if (l.kind == L_POOL) o << "pool " << l.name << " " << l.a << " " << l.b;
        else if (l.kind == L_ACT) o << "act " << l.name << " " << act_names[l.a & 3];
        else o << "fc " << l.name << " " << l.a;
    }
    const std::string s = o.str();
    PE_REQUIRE(s.size() + 1 <= cap, "buffer too small (%zu needed)", s.size() + 1);
    memcpy(buf, s.c_str(), s.size() + 1);
    return PE_OK;
}
int pe_net_num_blobs(const pe_net* net, int* n) {
    PE_REQUIRE(net && n, "null pointer");
    *n = (int)net->blobs.size();
    return PE_OK;
}
int pe_net_blob_info(const pe_net* net, int blob, char* name, size_t name_cap, int64_t* count) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_REQUIRE(blob >= 0 && blob < (int)net->blobs.size(), "blob %d out of range", blob);
    if (name && name_cap) {
        strncpy(name, net->def.blob_names[blob].c_str(), name_cap - 1);
        name[name_cap - 1] = 0;
    }
    if (count) *count = (int64_t)net->blobs[blob].size();
    return PE_OK;
}
int pe_net_set_blob(pe_net* net, int blob, const float* host, int64_t count) {
    PE_REQUIRE(net && host, "null pointer");
    PE_REQUIRE(blob >= 0 && blob < (int)net->blobs.size(), "blob %d out of range", blob);
    PE_REQUIRE(count == (int64_t)net->blobs[blob].size(), "blob %d holds %zu floats, got %lld", blob,
               net->blobs[blob].size(), (long long)count);
    memcpy(net->blobs[blob].data(), host, (size_t)count * sizeof(float));
    net->weights_dirty = true;
    net->fp8_state = 0;
    return PE_OK;
}
int pe_net_get_blob(const pe_net* net, int blob, float* host, int64_t count) {
    PE_REQUIRE(net && host, "null pointer");
    PE_REQUIRE(blob >= 0 && blob < (int)net->blobs.size(), "blob %d out of range", blob);
    PE_REQUIRE(count == (int64_t)net->blobs[blob].size(), "blob %d holds %zu floats, got %lld", blob,
               net->blobs[blob].size(), (long long)count);
    memcpy(host, net->blobs[blob].data(), (size_t)count * sizeof(float));
    return PE_OK;
}
int pe_net_fill(pe_net* net, uint64_t seed, int mode) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_REQUIRE(mode == 0 || mode == 1, "fill mode must be 0 or 1");
    for (size_t b = 0; b < net->blobs.size(); ++b) {
        const std::string& name = net->def.blob_names[b];
        std::vector<float>& v = net->blobs[b];
        const uint64_t key = stream_key(seed, 1000 + b);
        const char kind = name.back();
        if (kind == 'w') { /* Caffe xavier: U(-sqrt(3/fan_in), +sqrt(3/fan_in)), fan_in = count / num_output */
            const int64_t fan_in = (int64_t)v.size() / net->def.blob_shapes[b][0];
            const double s = std::sqrt(3.0 / (double)fan_in);
            for (size_t i = 0; i < v.size(); ++i) v[i] = uniform_f32(key, i, -s, s);
        } else if (kind == 'b') {
            for (size_t i = 0; i < v.size(); ++i) v[i] = mode == 0 ? 0.f : uniform_f32(key, i, -0.1, 0.1);
        } else {
            for (size_t i = 0; i < v.size(); ++i) v[i] = mode == 0 ? 0.25f : uniform_f32(key, i, -0.25, 0.5);
        }
    }
    net->weights_dirty = true;
    net->fp8_state = 0;
    return PE_OK;
}
int pe_net_save_weights(const pe_net* net, const char* path) {
    PE_REQUIRE(net && path, "null pointer");
    std::ofstream f(path, std::ios::binary);
    if (!f) { set_error("cannot write %s", path); return PE_ERR_IO; }
    const int32_t n = (int32_t)net->blobs.size();
    f.write("PEW1", 4);
    f.write(reinterpret_cast<const char*>(&n), 4);
    for (auto& b : net->blobs) {
        const int64_t c = (int64_t)b.size();
        f.write(reinterpret_cast<const char*>(&c), 8);
        f.write(reinterpret_cast<const char*>(b.data()), c * 4);
    }
    if (!f) { set_error("short write to %s", path); return PE_ERR_IO; }
    return PE_OK;
}
int pe_net_load_weights(pe_net* net, const char* path) {
    PE_REQUIRE(net && path, "null pointer");
    std::ifstream f(path, std::ios::binary);
    if (!f) { set_error("cannot open %s", path); return PE_ERR_IO; }
    char magic[4];
    int32_t n = 0;
    f.read(magic, 4);
    f.read(reinterpret_cast<char*>(&n), 4);
    if (!f || memcmp(magic, "PEW1", 4) != 0) { set_error("%s is not a PEW1 weight file", path); return PE_ERR_IO; }
    PE_REQUIRE(n == (int32_t)net->blobs.size(), "%s holds %d blobs, net has %zu", path, n, net->blobs.size());
    std::vector<std::vector<float>> tmp(net->blobs.size());
    for (size_t b = 0; b < tmp.size(); ++b) {
        int64_t c = 0;
        f.read(reinterpret_cast<char*>(&c), 8);
        if (!f || c != (int64_t)net->blobs[b].size()) { set_error("%s: blob %zu size mismatch", path, b); return PE_ERR_INVALID; }
        tmp[b].resize((size_t)c);
        f.read(reinterpret_cast<char*>(tmp[b].data()), c * 4);
        if (!f) { set_error("%s: truncated", path); return PE_ERR_IO; }
    }
    net->blobs.swap(tmp);
    net->weights_dirty = true;
    net->fp8_state = 0;
    return PE_OK;
}

int pe_compute_descriptors_dev(pe_net* net, const float* d_patches_hwc, int64_t n, int patch_channels,
                               float* d_out_desc) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_REQUIRE(n >= 0, "negative sample count");
    PE_REQUIRE(patch_channels == net->def.channels, "patches have %d channels, net input has %d",
               patch_channels, net->def.channels);
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    PE_REQUIRE(d_patches_hwc && d_out_desc, "null pointer");
    return net_forward_dev(net, d_patches_hwc, n, patch_channels, d_out_desc, net->def.desc_dim);
}

static int compute_descriptors_once(pe_net* net, const float* patches_hwc, int64_t n, int patch_channels, float* out_desc) {
    const int D = net->def.desc_dim;
    const size_t patch_floats = (size_t)4096 * patch_channels;
    const int64_t sub = 1024; /* patches per H2D stage: copy of stage i+1 overlaps compute of stage i */
    const int64_t stage_n = n < sub ? n : sub;
    PE_CHECK(net->in_stage[0].reserve(stage_n * patch_floats * sizeof(float)));
    if (n > sub) PE_CHECK(net->in_stage[1].reserve(stage_n * patch_floats * sizeof(float)));
    PE_CHECK(net->out_stage.reserve((size_t)n * D * sizeof(float)));
    int it = 0;
    for (int64_t first = 0; first < n; first += sub, ++it) {
        const int64_t m = n - first < sub ? n - first : sub;
        const int b = it & 1;
        if (it >= 2) PE_CUDA(cudaStreamWaitEvent(net->copy_stream, net->ev_done[b], 0));
        PE_CUDA(cudaMemcpyAsync(net->in_stage[b].p, patches_hwc + first * patch_floats, m * patch_floats * sizeof(float),
                                cudaMemcpyHostToDevice, net->copy_stream));
        PE_CUDA(cudaEventRecord(net->ev_in[b], net->copy_stream));
        PE_CUDA(cudaStreamWaitEvent(net->stream, net->ev_in[b], 0));
        PE_CHECK(net_forward_dev(net, net->in_stage[b].as<float>(), m, patch_channels,
                                 net->out_stage.as<float>() + (size_t)first * D, D));
        PE_CUDA(cudaEventRecord(net->ev_done[b], net->stream));
    }
    PE_CUDA(cudaMemcpyAsync(out_desc, net->out_stage.p, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToHost, net->stream));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    return forward_tc_check(net);
}

int pe_compute_descriptors(pe_net* net, const float* patches_hwc, int64_t n, int patch_channels, float* out_desc) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_REQUIRE(n >= 0, "negative sample count");
    PE_REQUIRE(patch_channels == net->def.channels, "patches have %d channels, net input has %d",
               patch_channels, net->def.channels);
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    PE_REQUIRE(patches_hwc && out_desc, "null pointer");
    /* PE_RETRY_FP16: opt-in e4m3 cross terms (forward_fp8 = 1) met an input beyond their range; the net has switched
     * to the all-fp16 split, the batch is forwarded again */
    int st = compute_descriptors_once(net, patches_hwc, n, patch_channels, out_desc);
    if (st == PE_RETRY_FP16) st = compute_descriptors_once(net, patches_hwc, n, patch_channels, out_desc);
    return st == PE_RETRY_FP16 ? PE_ERR_INVALID : st;
}

int pe_net_profile(pe_net* net, int enable) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    double ms[3];
    int64_t calls;
    PE_CHECK(prof_collect(&net->prof_events, 4, ms, &calls));
    net->profile = enable != 0;
    return PE_OK;
}
int pe_net_profile_read(pe_net* net, double ms[3], int64_t* calls) {
    PE_REQUIRE(net && ms && calls, "null pointer");
    PE_CHECK(net_ensure_device(net));
    return prof_collect(&net->prof_events, 4, ms, calls);
}
int pe_db_profile(pe_db* db, int enable) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    double ms[2];
    int64_t calls;
    PE_CHECK(prof_collect(&db->prof_events, 3, ms, &calls));
    db->profile = enable != 0;
    return PE_OK;
}
int pe_db_profile_read(pe_db* db, double ms[2], int64_t* calls) {
    PE_REQUIRE(db && ms && calls, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    return prof_collect(&db->prof_events, 3, ms, calls);
}
int pe_db_share_stream(pe_db* db, pe_net* net) {
    PE_REQUIRE(db && net, "null pointer");
    PE_REQUIRE(db->device == net->device, "DB and net live on different devices");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    if (db->own_stream && db->stream) cudaStreamDestroy(db->stream);
    db->stream = net->stream;
    db->own_stream = false;
    return PE_OK;
}

int pe_net_sync(pe_net* net) {
    PE_REQUIRE(net != nullptr, "null net");
    PE_CHECK(net_ensure_device(net));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    const int st = forward_tc_check(net);
    return st == PE_RETRY_FP16 ? PE_ERR_INVALID : st;   /* device entry points cannot replay the caller's batch */
}

int pe_synth_patches_dev(pe_net* net, uint64_t seed, int64_t first_index, int64_t n, int channels,
                         int poses_per_object, float* d_patches_hwc) {
    PE_REQUIRE(net && d_patches_hwc, "null pointer");
    PE_CHECK(net_ensure_device(net));
    return synth_patches_dev(net->stream, seed, first_index, n, channels, poses_per_object, d_patches_hwc);
}
int pe_synth_labels(uint64_t seed, int64_t first_index, int64_t n, int poses_per_object, float* labels) {
    PE_REQUIRE(labels && poses_per_object > 0 && n >= 0, "bad arguments");
    synth_labels_host(seed, first_index, n, poses_per_object, labels);
    return PE_OK;
}

}  // extern "C"

/* ---------------------------------------------------------------- binarize --- */
namespace {
__global__ void binarize_kernel(const float* __restrict__ desc, long long n, int dim, int words,
                                uint32_t* __restrict__ bits) {
    /* one warp per output word: bit j of word w is desc[32*w + j] > 0 */
    const long long gw = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gw >= n * words) return;
    const long long row = gw / words;
    const int w = (int)(gw - row * words);
    const int col = w * 32 + lane;
    const bool on = col < dim && desc[row * dim + col] > 0.f;
    const unsigned v = __ballot_sync(0xffffffffu, on);
    if (lane == 0) bits[gw] = v;
}
}  // namespace

extern "C" {

int pe_binarize_dev(pe_net* net, const float* d_desc, int64_t n, int dim, uint32_t* d_bits) {
    PE_REQUIRE(net != nullptr && dim > 0 && n >= 0, "bad arguments");
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    const int words = (dim + 31) / 32;
    const long long threads = n * words * 32;
    binarize_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, net->stream>>>(d_desc, n, dim, words, d_bits);
    PE_LAUNCHED();
    return PE_OK;
}

int pe_binarize(pe_net* net, const float* desc, int64_t n, int dim, uint32_t* bits) {
    PE_REQUIRE(net != nullptr && dim > 0 && n >= 0, "bad arguments");
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    PE_REQUIRE(desc && bits, "null pointer");
    const int words = (dim + 31) / 32;
    PE_CHECK(net->out_stage.reserve((size_t)n * dim * 4));
    PE_CHECK(net->fc_b.reserve((size_t)n * words * 4));
    PE_CUDA(cudaMemcpyAsync(net->out_stage.p, desc, (size_t)n * dim * 4, cudaMemcpyHostToDevice, net->stream));
    PE_CHECK(pe_binarize_dev(net, net->out_stage.as<float>(), n, dim, net->fc_b.as<uint32_t>()));
    PE_CUDA(cudaMemcpyAsync(bits, net->fc_b.p, (size_t)n * words * 4, cudaMemcpyDeviceToHost, net->stream));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    return PE_OK;
}

/* -------------------------------------------------------------------- db --- */
static int db_common(int64_t t_local, int64_t row_offset, const float* labels_all, int64_t t_total,
                     int poses_per_object, int device, pe_db** out, pe_db** made) {
    PE_REQUIRE(out != nullptr, "null output pointer");
    *out = nullptr;
    PE_REQUIRE(t_local >= 0 && row_offset >= 0 && t_total >= t_local + row_offset,
               "shard [%lld, %lld) outside 0..%lld", (long long)row_offset, (long long)(row_offset + t_local), (long long)t_total);
    PE_REQUIRE(t_total < (1LL << 31), "more than 2^31 templates");
    PE_REQUIRE(poses_per_object >= 0, "negative poses_per_object");
    PE_CHECK(set_device_checked(device));
    pe_db* db = new pe_db();
    db->device = device;
    db->t_local = t_local; db->t_total = t_total; db->row_offset = row_offset;
    db->poses_per_object = poses_per_object;
    cudaDeviceGetAttribute(&db->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete db;
        set_error("cudaStreamCreate failed");
        return PE_ERR_CUDA;
    }
    if (labels_all && t_total > 0) {
        if (cudaMalloc(&db->d_labels, (size_t)t_total * 8 * 4) != cudaSuccess ||
            cudaMemcpy(db->d_labels, labels_all, (size_t)t_total * 8 * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
            set_error("label upload failed: %s", cudaGetErrorString(cudaGetLastError()));
            pe_db_destroy(db);
            return PE_ERR_CUDA;
        }
    }
    *made = db;
    return PE_OK;
}

int pe_db_create(const float* desc, int64_t t_local, int dim, int64_t row_offset, const float* labels_all,
                 int64_t t_total, int poses_per_object, int device, pe_db** out) {
    PE_REQUIRE(dim >= 1 && (desc || t_local == 0), "bad descriptor matrix");
    pe_db* db = nullptr;
    PE_CHECK(db_common(t_local, row_offset, labels_all, t_total, poses_per_object, device, out, &db));
    db->dim = dim;
    const int st = db_set_descriptors(db, desc, false);
    if (st != PE_OK) { pe_db_destroy(db); return st; }
    *out = db;
    return PE_OK;
}
int pe_db_create_dev(const float* d_desc, int64_t t_local, int dim, int64_t row_offset, const float* labels_all_host,
                     int64_t t_total, int poses_per_object, int device, pe_db** out) {
    PE_REQUIRE(dim >= 1 && (d_desc || t_local == 0), "bad descriptor matrix");
    pe_db* db = nullptr;
    PE_CHECK(db_common(t_local, row_offset, labels_all_host, t_total, poses_per_object, device, out, &db));
    db->dim = dim;
    const int st = db_set_descriptors(db, d_desc, true);
    if (st != PE_OK) { pe_db_destroy(db); return st; }
    *out = db;
    return PE_OK;
}
int pe_db_create_hamming(const uint32_t* codes, int64_t t_local, int words_per_code, int64_t row_offset,
                         const float* labels_all, int64_t t_total, int poses_per_object, int device, pe_db** out) {
    PE_REQUIRE(words_per_code == 1 || words_per_code == 2 || words_per_code == 4 || words_per_code == 8,
               "words_per_code must be 1, 2, 4 or 8");
    PE_REQUIRE(codes || t_local == 0, "null codes");
    pe_db* db = nullptr;
    PE_CHECK(db_common(t_local, row_offset, labels_all, t_total, poses_per_object, device, out, &db));
    db->hamming = true;
    db->words = words_per_code;
    db->dim = words_per_code * 32;
    const size_t bytes = (size_t)(t_local > 0 ? t_local : 1) * words_per_code * 4;
    cudaPointerAttributes attr;
    const bool on_dev = codes && cudaPointerGetAttributes(&attr, codes) == cudaSuccess && attr.type == cudaMemoryTypeDevice;
    cudaGetLastError();
    if (cudaMalloc(&db->d_codes, bytes) != cudaSuccess ||
        (t_local > 0 && cudaMemcpy(db->d_codes, codes, (size_t)t_local * words_per_code * 4,
                                   on_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice) != cudaSuccess)) {
        set_error("code upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        pe_db_destroy(db);
        return PE_ERR_CUDA;
    }
    const int st = knn_tc_prepare(db);      /* +-1 tensor tiles of the codes (32 / 64 / 128 bit) */
    if (st != PE_OK) { pe_db_destroy(db); return st; }
    *out = db;
    return PE_OK;
}
void pe_db_destroy(pe_db* db) {
    if (!db) return;
    cudaSetDevice(db->device);
    if (db->stream && db->own_stream) cudaStreamSynchronize(db->stream);
    else cudaDeviceSynchronize();
    for (cudaEvent_t e : db->prof_events) cudaEventDestroy(e);
    if (db->d_desc) cudaFree(db->d_desc);
    if (db->d_norm) cudaFree(db->d_norm);
    if (db->d_codes) cudaFree(db->d_codes);
    if (db->d_labels) cudaFree(db->d_labels);
    if (db->d_tc) cudaFree(db->d_tc);
    if (db->d_center) cudaFree(db->d_center);
    if (db->d_norm_c) cudaFree(db->d_norm_c);
    for (DevBuf* b : {&db->key_scale, &db->flag, &db->fb_d, &db->fb_i, &db->fb_done}) b->release();
    if (db->d_tc_err) cudaFreeHost(db->d_tc_err);
    db->tau.release();
    for (DevBuf* b : {&db->q, &db->q_in, &db->cand_d, &db->cand_i, &db->out_i, &db->out_d, &db->qlab, &db->knn_idx,
                      &db->scratch, &db->scratch2, &db->gather, &db->boot, &db->m_q, &db->m_ci, &db->m_cd, &db->m_ri, &db->m_rd})
        b->release();
    if (db->ev_match) cudaEventDestroy(db->ev_match);
    for (int i = 0; i < 2; ++i) {
        db->m_desc[i].release();
        if (db->ev_free[i]) cudaEventDestroy(db->ev_free[i]);
    }
    if (db->stream && db->own_stream) cudaStreamDestroy(db->stream);
    delete db;
}
int pe_db_info(const pe_db* db, int64_t* t_local, int64_t* t_total, int* dim, int* is_hamming) {
    PE_REQUIRE(db != nullptr, "null db");
    if (t_local) *t_local = db->t_local;
    if (t_total) *t_total = db->t_total;
    if (dim) *dim = db->hamming ? db->words : db->dim;
    if (is_hamming) *is_hamming = db->hamming ? 1 : 0;
    return PE_OK;
}
int pe_db_sync(pe_db* db) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    PE_CHECK(comm_check(db->last_comm));
    return knn_tc_check(db);
}

int pe_db_knn_flagged(pe_db* db, int64_t* n) {
    PE_REQUIRE(db && n, "null pointer");
    *n = 0;
    if (!db->flag.p) return PE_OK;
    PE_CUDA(cudaSetDevice(db->device));
    int h = 0;
    PE_CUDA(cudaMemcpyAsync(&h, db->flag.p, sizeof(int), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    *n = h;
    return PE_OK;
}

int pe_knn_dev(pe_db* db, const float* d_queries, int64_t nq, int k, int32_t* d_out_idx, float* d_out_dist) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_REQUIRE(nq == 0 || (d_queries && d_out_idx && d_out_dist), "null pointer");
    return knn_l2_local(db, d_queries, nq, k, d_out_idx, d_out_dist);
}
int pe_knn(pe_db* db, const float* queries, int64_t nq, int k, int32_t* out_idx, float* out_dist) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_REQUIRE(!db->hamming, "pe_knn on a Hamming DB");
    PE_REQUIRE(k >= 1 && k <= 16 && nq >= 0, "bad k / nq");
    if (nq == 0) return PE_OK;
    PE_REQUIRE(queries && out_idx && out_dist, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->q_in.reserve((size_t)nq * db->dim * 4));
    PE_CHECK(db->out_i.reserve((size_t)nq * k * 4));
    PE_CHECK(db->out_d.reserve((size_t)nq * k * 4));
    PE_CUDA(cudaMemcpyAsync(db->q_in.p, queries, (size_t)nq * db->dim * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CHECK(knn_l2_local(db, db->q_in.as<float>(), nq, k, db->out_i.as<int32_t>(), db->out_d.as<float>()));
    PE_CUDA(cudaMemcpyAsync(out_idx, db->out_i.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_dist, db->out_d.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    return knn_tc_check(db);
}
int pe_knn_hamming_dev(pe_db* db, const uint32_t* d_qcodes, int64_t nq, int k, int32_t* d_out_idx, int32_t* d_out_dist) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_REQUIRE(nq == 0 || (d_qcodes && d_out_idx && d_out_dist), "null pointer");
    return knn_hamming_local(db, d_qcodes, nq, k, d_out_idx, d_out_dist);
}
int pe_knn_hamming(pe_db* db, const uint32_t* qcodes, int64_t nq, int k, int32_t* out_idx, int32_t* out_dist) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_REQUIRE(db->hamming, "pe_knn_hamming on an L2 DB");
    PE_REQUIRE(k >= 1 && k <= 16 && nq >= 0, "bad k / nq");
    if (nq == 0) return PE_OK;
    PE_REQUIRE(qcodes && out_idx && out_dist, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->q_in.reserve((size_t)nq * db->words * 4));
    PE_CHECK(db->out_i.reserve((size_t)nq * k * 4));
    PE_CHECK(db->out_d.reserve((size_t)nq * k * 4));
    PE_CUDA(cudaMemcpyAsync(db->q_in.p, qcodes, (size_t)nq * db->words * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CHECK(knn_hamming_local(db, db->q_in.as<uint32_t>(), nq, k, db->out_i.as<int32_t>(), db->out_d.as<int32_t>()));
    PE_CUDA(cudaMemcpyAsync(out_idx, db->out_i.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_dist, db->out_d.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    return PE_OK;
}

/* ------------------------------------------------------------ reductions --- */
static int upload_idx_labels(pe_db* db, const int32_t* knn_idx, const float* q_labels, int64_t nq, int k) {
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->knn_idx.reserve((size_t)(nq > 0 ? nq : 1) * k * 4));
    PE_CHECK(db->qlab.reserve((size_t)(nq > 0 ? nq : 1) * 8 * 4));
    if (nq > 0) {
        PE_CUDA(cudaMemcpyAsync(db->knn_idx.p, knn_idx, (size_t)nq * k * 4, cudaMemcpyHostToDevice, db->stream));
        if (q_labels) PE_CUDA(cudaMemcpyAsync(db->qlab.p, q_labels, (size_t)nq * 8 * 4, cudaMemcpyHostToDevice, db->stream));
    }
    return PE_OK;
}

int pe_histogram_dev(pe_db* db, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq, int k,
                     const int* rot_inv, int n_rot, const float* bins, int n_bins, float* out) {
    PE_REQUIRE(db && bins && out && (rot_inv || n_rot == 0), "null pointer");
    PE_REQUIRE(nq >= 0 && (nq == 0 || (d_knn_idx && d_q_labels)), "null pointer");
    return pose_histogram_dev(db, d_knn_idx, d_q_labels, nq, k, rot_inv, n_rot, bins, n_bins, out);
}
int pe_histogram(pe_db* db, const int32_t* knn_idx, const float* q_labels, int64_t nq, int k, const int* rot_inv,
                 int n_rot, const float* bins, int n_bins, float* out) {
    PE_REQUIRE(db && bins && out && (rot_inv || n_rot == 0), "null pointer");
    PE_REQUIRE(nq >= 0 && (nq == 0 || (knn_idx && q_labels)), "null pointer");
    PE_REQUIRE(k >= 1 && k <= 32, "k=%d outside 1..32", k);
    for (int64_t i = 0; i < nq; ++i) {
        const float c = q_labels[i * 8];
        PE_REQUIRE(c >= 0.f && c < (float)n_rot, "query %lld: class %g has no rotInv entry", (long long)i, c);
    }
    PE_CHECK(upload_idx_labels(db, knn_idx, q_labels, nq, k));
    return pose_histogram_dev(db, db->knn_idx.as<int>(), db->qlab.as<float>(), nq, k, rot_inv, n_rot, bins, n_bins, out);
}
int pe_confusion(pe_db* db, const int32_t* knn_idx, const float* q_labels, int64_t nq, int k,
                 const int* class_to_local, int n_classes, int n_local, int nr_test_poses, float* out) {
    PE_REQUIRE(db && class_to_local && out, "null pointer");
    PE_REQUIRE(nq >= 0 && (nq == 0 || (knn_idx && q_labels)), "null pointer");
    PE_REQUIRE(k >= 1 && k <= 32, "k=%d outside 1..32", k);
    PE_CHECK(upload_idx_labels(db, knn_idx, q_labels, nq, k));
    return pose_confusion_dev(db, db->knn_idx.as<int>(), db->qlab.as<float>(), nq, k, class_to_local, n_classes,
                              n_local, nr_test_poses, out);
}
int pe_knn_bootstrap(pe_db* db, const int32_t* knn3_idx, int64_t nq, int nr_query_poses, int32_t* out, int32_t* out_len) {
    PE_REQUIRE(db != nullptr && nq >= 0, "bad arguments");
    if (nq == 0) return PE_OK;
    PE_REQUIRE(knn3_idx && out && out_len, "null pointer");
    PE_CHECK(upload_idx_labels(db, knn3_idx, nullptr, nq, 3));
    PE_CHECK(db->boot.reserve((size_t)nq * 5 * 4));
    int* d_out = db->boot.as<int>();
    int* d_len = d_out + nq * 4;
    PE_CHECK(pose_bootstrap_dev(db, db->knn_idx.as<int>(), nq, nr_query_poses, d_out, d_len));
    PE_CUDA(cudaMemcpyAsync(out, d_out, (size_t)nq * 16, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_len, d_len, (size_t)nq * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    return PE_OK;
}
int pe_knn_accuracy(pe_db* db, const int32_t* max_sim_tmpl0, const int32_t* bootstrap, int64_t nq, int nr_query_poses,
                    float* intra_pct, float* inter_pct) {
    PE_REQUIRE(db && intra_pct && inter_pct && nq >= 0, "bad arguments");
    PE_REQUIRE(nq == 0 || (max_sim_tmpl0 && bootstrap), "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->boot.reserve((size_t)(nq > 0 ? nq : 1) * 5 * 4));
    int* d_boot = db->boot.as<int>();
    int* d_ms = d_boot + nq * 4;
    if (nq > 0) {
        PE_CUDA(cudaMemcpyAsync(d_boot, bootstrap, (size_t)nq * 16, cudaMemcpyHostToDevice, db->stream));
        PE_CUDA(cudaMemcpyAsync(d_ms, max_sim_tmpl0, (size_t)nq * 4, cudaMemcpyHostToDevice, db->stream));
    }
    return pose_accuracy_dev(db, d_ms, d_boot, nq, nr_query_poses, intra_pct, inter_pct);
}

}  // extern "C"
