/*
 * prototxt.cpp - TEST-phase reader for the Caffe text format used by the reference's
 * network/manifold*.prototxt, plus a built-in table of the seven variants.
 *
 * Replaces what `caffe::Net<float>(path, caffe::TEST)` does at reference
 * src/networksolver.cpp:286: keep layers without `include {phase: TRAIN}`, wire the
 * Slice/Concat/Silence channel plumbing (network/manifold_rgb_16.prototxt:6-33) into a channel
 * list, and record Convolution / Pooling / ReLU / ELU / PReLU / InnerProduct in order with the
 * learnable blobs in Caffe order (weight, bias; PReLU slope).
 */
#include <cmath>
#include <cstdlib>
#include <map>
#include <memory>
#include <sstream>

#include "common.h"

namespace pe {
namespace {

struct Node;
using NodePtr = std::unique_ptr<Node>;
struct Field {
    std::string key;
    std::string scalar;  /* valid when !block */
    NodePtr block;
};
struct Node {
    std::vector<Field> fields;
    std::vector<const Field*> all(const std::string& k) const {
        std::vector<const Field*> r;
        for (auto& f : fields)
            if (f.key == k) r.push_back(&f);
        return r;
    }
    const Field* first(const std::string& k) const {
        for (auto& f : fields)
            if (f.key == k) return &f;
        return nullptr;
    }
    std::string str(const std::string& k, const std::string& dflt = "") const {
        auto f = first(k);
        return (f && !f->block) ? f->scalar : dflt;
    }
    bool get_int(const std::string& k, int dflt, int* out) const {
        auto f = first(k);
        if (!f) { *out = dflt; return true; }
        if (f->block) return false;
        char* end = nullptr;
        long v = std::strtol(f->scalar.c_str(), &end, 10);
        if (end == f->scalar.c_str() || *end) return false;
        *out = (int)v;
        return true;
    }
    const Node* block(const std::string& k) const {
        auto f = first(k);
        return (f && f->block) ? f->block.get() : nullptr;
    }
};

bool tokenize(const std::string& text, std::vector<std::string>* toks) {
    size_t i = 0, n = text.size();
    while (i < n) {
        char c = text[i];
        if (c == '#') {
            while (i < n && text[i] != '\n') ++i;
        } else if (isspace((unsigned char)c)) {
            ++i;
        } else if (c == '{' || c == '}' || c == ':') {
            toks->push_back(std::string(1, c));
            ++i;
        } else if (c == '"' || c == '\'') {
            size_t j = text.find(c, i + 1);
            if (j == std::string::npos) return false;
            toks->push_back(std::string("\"") + text.substr(i + 1, j - i - 1)); /* leading " marks a string */
            i = j + 1;
        } else {
            size_t j = i;
            while (j < n && !isspace((unsigned char)text[j]) && text[j] != '{' && text[j] != '}' &&
                   text[j] != ':' && text[j] != '#')
                ++j;
            toks->push_back(text.substr(i, j - i));
            i = j;
        }
    }
    return true;
}

bool parse_block(const std::vector<std::string>& t, size_t* pos, Node* out, int depth) {
    if (depth > 32) return false;
    while (*pos < t.size() && t[*pos] != "}") {
        Field f;
        f.key = t[(*pos)++];
        if (f.key == "{" || f.key == ":" || f.key.empty() || f.key[0] == '"') return false;
        if (*pos >= t.size()) return false;
        if (t[*pos] == ":") {
            ++*pos;
            if (*pos >= t.size()) return false;
        }
        if (t[*pos] == "{") {
            ++*pos;
            f.block.reset(new Node());
            if (!parse_block(t, pos, f.block.get(), depth + 1)) return false;
            if (*pos >= t.size() || t[*pos] != "}") return false;
            ++*pos;
        } else {
            const std::string& v = t[(*pos)++];
            if (v == "}" || v == ":") return false;
            f.scalar = (!v.empty() && v[0] == '"') ? v.substr(1) : v;
        }
        out->fields.push_back(std::move(f));
    }
    return true;
}

int pooled_size(int in, int k, int s) {
    /* Caffe PoolingLayer::Reshape: static_cast<int>(ceil(static_cast<float>(in - k) / s)) + 1 */
    return (int)std::ceil((float)(in - k) / (float)s) + 1;
}

int act_kind(const std::string& t) {
    if (t == "ReLU") return PE_ACT_RELU;
    if (t == "ELU") return PE_ACT_ELU;
    if (t == "PReLU") return PE_ACT_PRELU;
    return 0;
}

#define PARSE_FAIL(...)              \
    do {                             \
        set_error(__VA_ARGS__);      \
        return PE_ERR_PARSE;         \
    } while (0)

}  // namespace

int parse_prototxt(const std::string& text, NetDef* net) {
    std::vector<std::string> toks;
    if (!tokenize(text, &toks)) PARSE_FAIL("prototxt: unterminated string");
    Node root;
    size_t pos = 0;
    if (!parse_block(toks, &pos, &root, 0) || pos != toks.size())
        PARSE_FAIL("prototxt: syntax error near token %zu", pos);

    *net = NetDef();
    const Node* ishape = root.block("input_shape");
    std::vector<int> dims;
    if (ishape) {
        for (auto f : ishape->all("dim")) dims.push_back(std::atoi(f->scalar.c_str()));
    } else {
        for (auto f : root.all("input_dim")) dims.push_back(std::atoi(f->scalar.c_str()));
    }
    if (dims.size() != 4 || dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 || dims[3] <= 0)
        PARSE_FAIL("prototxt: need a 4-d input_shape");
    net->batch = dims[0]; net->channels = dims[1]; net->height = dims[2]; net->width = dims[3];
    const std::string data_name = root.str("input");
    if (data_name.empty()) PARSE_FAIL("prototxt: missing `input:`");

    std::map<std::string, std::vector<int>> chan; /* pre-conv blobs: name -> channel list */
    std::vector<int> all(net->channels);
    for (int c = 0; c < net->channels; ++c) all[c] = c;
    chan[data_name] = all;

    std::string cur_blob;
    bool have_cur = false;
    int sc = 0, sh = 0, sw = 0; /* current activation shape */

    auto layers = root.all("layer");
    if (layers.empty()) layers = root.all("layers");
    for (auto lf : layers) {
        if (!lf->block) PARSE_FAIL("prototxt: `layer` must be a block");
        const Node& L = *lf->block;
        bool train_only = false;
        for (auto inc : L.all("include"))
            if (inc->block && inc->block->str("phase") == "TRAIN") train_only = true;
        if (train_only) continue;
        const std::string type = L.str("type"), name = L.str("name");
        std::vector<std::string> bottoms, tops;
        for (auto f : L.all("bottom")) bottoms.push_back(f->scalar);
        for (auto f : L.all("top")) tops.push_back(f->scalar);

        if (type == "Slice") {
            const Node* sp = L.block("slice_param");
            int axis = 1;
            if (sp && !sp->get_int("axis", 1, &axis)) PARSE_FAIL("%s: bad axis", name.c_str());
            if (axis != 1 || bottoms.size() != 1 || !chan.count(bottoms[0]))
                PARSE_FAIL("%s: only channel slices of input blobs are supported", name.c_str());
            const std::vector<int>& src = chan[bottoms[0]];
            std::vector<int> pts{0};
            if (sp)
                for (auto f : sp->all("slice_point")) pts.push_back(std::atoi(f->scalar.c_str()));
            pts.push_back((int)src.size());
            if (pts.size() - 1 != tops.size()) PARSE_FAIL("%s: slice_point/top mismatch", name.c_str());
            for (size_t i = 0; i + 1 < pts.size(); ++i) {
                if (pts[i] < 0 || pts[i + 1] < pts[i] || pts[i + 1] > (int)src.size())
                    PARSE_FAIL("%s: bad slice_point", name.c_str());
                chan[tops[i]] = std::vector<int>(src.begin() + pts[i], src.begin() + pts[i + 1]);
            }
        } else if (type == "Concat") {
            const Node* cp = L.block("concat_param");
            int axis = 1;
            if (cp && !cp->get_int("axis", 1, &axis)) PARSE_FAIL("%s: bad axis", name.c_str());
            if (axis != 1 || tops.size() != 1) PARSE_FAIL("%s: only channel concat is supported", name.c_str());
            std::vector<int> out;
            for (auto& b : bottoms) {
                if (!chan.count(b)) PARSE_FAIL("%s: unknown bottom %s", name.c_str(), b.c_str());
                out.insert(out.end(), chan[b].begin(), chan[b].end());
            }
            chan[tops[0]] = out;
        } else if (type == "Silence") {
            /* consumes unused Slice tops; nothing to compute */
        } else if (type == "Convolution") {
            const Node* p = L.block("convolution_param");
            int nf = 0, k = 0, stride = 1, pad = 0;
            if (!p || !p->get_int("num_output", 0, &nf) || !p->get_int("kernel_size", 0, &k) ||
                !p->get_int("stride", 1, &stride) || !p->get_int("pad", 0, &pad) || nf <= 0 || k <= 0)
                PARSE_FAIL("%s: bad convolution_param", name.c_str());
            if (stride != 1 || pad != 0) PARSE_FAIL("%s: only stride 1 / pad 0 is supported", name.c_str());
            if (bottoms.size() != 1 || tops.size() != 1) PARSE_FAIL("%s: one bottom/top expected", name.c_str());
            if (!have_cur) {
                if (!chan.count(bottoms[0])) PARSE_FAIL("%s: unknown bottom %s", name.c_str(), bottoms[0].c_str());
                net->sel = chan[bottoms[0]];
                sc = (int)net->sel.size(); sh = net->height; sw = net->width;
                have_cur = true;
            } else if (bottoms[0] != cur_blob) {
                PARSE_FAIL("%s: bottom %s is not the running blob %s", name.c_str(), bottoms[0].c_str(), cur_blob.c_str());
            }
            if (sh < k || sw < k) PARSE_FAIL("%s: kernel larger than input", name.c_str());
            Layer ly;
            ly.kind = L_CONV; ly.name = name; ly.a = nf; ly.b = k;
            ly.blob0 = (int)net->blob_names.size(); ly.blob1 = ly.blob0 + 1;
            net->blob_names.push_back(name + ".w");
            net->blob_shapes.push_back({nf, sc, k, k});
            net->blob_names.push_back(name + ".b");
            net->blob_shapes.push_back({nf});
            net->layers.push_back(ly);
            sc = nf; sh = sh - k + 1; sw = sw - k + 1;
            cur_blob = tops[0];
        } else if (type == "Pooling") {
            const Node* p = L.block("pooling_param");
            int k = 0, stride = 1;
            if (!have_cur || !p || !p->get_int("kernel_size", 0, &k) || !p->get_int("stride", 1, &stride) || k <= 0 || stride <= 0)
                PARSE_FAIL("%s: bad pooling_param", name.c_str());
            if (p->str("pool", "MAX") != "MAX") PARSE_FAIL("%s: only MAX pooling is supported", name.c_str());
            if (bottoms.size() != 1 || bottoms[0] != cur_blob || tops.size() != 1)
                PARSE_FAIL("%s: bottom is not the running blob", name.c_str());
            if (sh < k || sw < k) PARSE_FAIL("%s: kernel larger than input", name.c_str());
            Layer ly;
            ly.kind = L_POOL; ly.name = name; ly.a = k; ly.b = stride;
            net->layers.push_back(ly);
            sh = pooled_size(sh, k, stride); sw = pooled_size(sw, k, stride);
            cur_blob = tops[0];
        } else if (act_kind(type)) {
            if (!have_cur || bottoms.size() != 1 || bottoms[0] != cur_blob || tops.size() != 1)
                PARSE_FAIL("%s: bottom is not the running blob", name.c_str());
            Layer ly;
            ly.kind = L_ACT; ly.name = name; ly.a = act_kind(type);
            if (ly.a == PE_ACT_PRELU) {
                ly.blob0 = (int)net->blob_names.size();
                net->blob_names.push_back(name + ".a");
                net->blob_shapes.push_back({sc});
            }
            net->layers.push_back(ly);
            cur_blob = tops[0];
        } else if (type == "InnerProduct") {
            const Node* p = L.block("inner_product_param");
            int nout = 0;
            if (!have_cur || !p || !p->get_int("num_output", 0, &nout) || nout <= 0)
                PARSE_FAIL("%s: bad inner_product_param", name.c_str());
            if (bottoms.size() != 1 || bottoms[0] != cur_blob || tops.size() != 1)
                PARSE_FAIL("%s: bottom is not the running blob", name.c_str());
            Layer ly;
            ly.kind = L_FC; ly.name = name; ly.a = nout;
            ly.blob0 = (int)net->blob_names.size(); ly.blob1 = ly.blob0 + 1;
            net->blob_names.push_back(name + ".w");
            net->blob_shapes.push_back({nout, (int64_t)sc * sh * sw});
            net->blob_names.push_back(name + ".b");
            net->blob_shapes.push_back({nout});
            net->layers.push_back(ly);
            sc = nout; sh = 1; sw = 1;
            cur_blob = tops[0];
        } else {
            PARSE_FAIL("%s: unsupported TEST-phase layer type `%s`", name.c_str(), type.c_str());
        }
    }
    if (!have_cur) PARSE_FAIL("prototxt: no compute layers");
    net->desc_dim = sc * sh * sw;
    return PE_OK;
}

/* ---- built-in variants: facts from network/manifold*.prototxt (input dims :4; channel
 * plumbing manifold_rgb_16:6-33; activations manifold:33-38, rgbdnor_16:35, rgbd_16:66;
 * outputs manifold:104 (3), manifold_wang:104 (64), *_16:104/135 (16), _bin:113-154) ---- */
namespace {
struct Variant {
    const char* name;
    int batch, channels;
    const char* groups; /* comma list of slice tops concatenated for conv1, or "" */
    const char* act;
    int n_fc;
    int fc[4];
    int tail; /* 0 none, 1 triplet (manifold:111-166), 2 binarization loss (_bin:156-163) */
};
const Variant kVariants[] = {
    {"manifold", 330, 4, "", "ReLU", 2, {256, 3, 0, 0}, 1},
    {"manifold_wang", 300, 4, "", "ReLU", 2, {256, 64, 0, 0}, 0},
    {"manifold_rgb_16", 300, 7, "rgb", "ELU", 2, {256, 16, 0, 0}, 0},
    {"manifold_rgbd_16", 300, 7, "rgb,dep", "PReLU", 2, {256, 16, 0, 0}, 0},
    {"manifold_rgbnor_16", 300, 7, "rgb,nor", "ELU", 2, {256, 16, 0, 0}, 0},
    {"manifold_rgbdnor_16", 300, 7, "", "ELU", 2, {256, 16, 0, 0}, 0},
    {"manifold_rgbdnor_16_bin", 300, 7, "", "PReLU", 4, {256, 16, 512, 128}, 2},
};

void emit_layer(std::ostringstream& o, const std::string& name, const std::string& type,
                const std::vector<std::string>& bottoms, const std::vector<std::string>& tops,
                const std::string& body, bool train_only = false) {
    o << "layer {\n  name: \"" << name << "\"\n  type: \"" << type << "\"\n";
    for (auto& b : bottoms) o << "  bottom: \"" << b << "\"\n";
    for (auto& t : tops) o << "  top: \"" << t << "\"\n";
    o << body;
    if (train_only) o << "  include {phase: TRAIN}\n";
    o << "}\n";
}
}  // namespace

bool builtin_prototxt(const std::string& variant, std::string* text) {
    const Variant* v = nullptr;
    for (auto& k : kVariants)
        if (variant == k.name) v = &k;
    if (!v) return false;
    std::ostringstream o;
    o << "name: \"MANIFOLD\"\ninput: \"data\"\ninput_shape{dim: " << v->batch << " dim: " << v->channels
      << " dim: 64 dim: 64}\n";
    std::string src = "data";
    const std::string groups = v->groups;
    if (!groups.empty()) {
        emit_layer(o, "slicer", "Slice", {"data"}, {"rgb", "dep", "nor"},
                   "  slice_param { axis: 1 slice_point: 3 slice_point: 4 }\n");
        std::vector<std::string> used, unused;
        for (const char* g : {"rgb", "dep", "nor"}) {
            bool in = (("," + groups + ",").find(std::string(",") + g + ",") != std::string::npos);
            (in ? used : unused).push_back(g);
        }
        emit_layer(o, "concat", "Concat", used, {"input"}, "  concat_param { axis: 1 }\n");
        emit_layer(o, "silencer", "Silence", unused, {}, "");
        src = "input";
    }
    const std::string filler = "    weight_filler { type: \"xavier\" }\n    bias_filler { type: \"constant\" }\n";
    emit_layer(o, "conv1", "Convolution", {src}, {"conv1"},
               "  convolution_param {\n    num_output: 16\n    kernel_size: 8\n" + filler + "  }\n");
    emit_layer(o, "pool1", "Pooling", {"conv1"}, {"pool1"},
               "  pooling_param { pool: MAX kernel_size: 2 stride: 2 }\n");
    emit_layer(o, "relu1", v->act, {"pool1"}, {"pool1"}, "");
    emit_layer(o, "conv2", "Convolution", {"pool1"}, {"conv2"},
               "  convolution_param {\n    num_output: 7\n    kernel_size: 5\n" + filler + "  }\n");
    emit_layer(o, "pool2", "Pooling", {"conv2"}, {"pool2"},
               "  pooling_param { pool: MAX kernel_size: 2 stride: 2 }\n");
    emit_layer(o, "relu2", v->act, {"pool2"}, {"pool2"}, "");
    const char* fc_names[4] = {"ip1", "ip2", "bin_ip1", "bin_ip2"};
    std::string prev = "pool2";
    for (int i = 0; i < v->n_fc; ++i) {
        std::ostringstream body;
        body << "  inner_product_param {\n    num_output: " << v->fc[i] << "\n" << filler << "  }\n";
        emit_layer(o, fc_names[i], "InnerProduct", {prev}, {fc_names[i]}, body.str());
        if (i + 1 < v->n_fc)
            emit_layer(o, std::string("relu_") + fc_names[i], v->act, {fc_names[i]}, {fc_names[i]}, "");
        prev = fc_names[i];
    }
    if (v->tail == 1) {
        emit_layer(o, "slicer", "Slice", {prev}, {"feat_triplets", "feat_pairs"},
                   "  slice_param { axis: 0 slice_point: 198 }\n", true);
        emit_layer(o, "tripletLoss", "TripletLoss", {"feat_triplets", "label_triplets"}, {"loss1"}, "", true);
    } else if (v->tail == 2) {
        emit_layer(o, "loss", "BinarizationLoss", {prev, prev}, {"loss"}, "", true);
    }
    *text = o.str();
    return true;
}

}  // namespace pe
