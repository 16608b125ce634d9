/*
 * io.cpp - on-disk sample sets without HDF5.
 *
 * The reference stores sample sets through hdf5Handler (reference src/hdf5handler.cpp:11-132): dataset "data"
 * [N][C][H][W] f32 (Caffe layout; `write` transposes Sample.data from OpenCV's interleaved HWC, `read` transposes
 * back) and dataset "label" [N][label_dim] f32.  HDF5 is absent on this box, so the same two arrays go into one flat
 * little-endian file:
 *
 *     "PESS0001" | int64 n | int32 C | int32 H | int32 W | int32 label_dim | data [n][C][H][W] f32 | label [n][label_dim] f32
 *
 * i.e. byte for byte the two HDF5 datasets behind a 32-byte header (tools/h5_to_pess.py dumps one into the other where
 * h5py exists).  pe_samples_write / pe_samples_read take and return the HWC patches every other entry point of the
 * library uses, with hdf5Handler's transposes inside.  Host code only.
 */
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <cstring>
#include <vector>

#include "common.h"

using namespace pe;

namespace {
struct Header {
    char magic[8];
    int64_t n;
    int32_t c, h, w, label_dim;
};
static_assert(sizeof(Header) == 32, "PESS header is 32 bytes");

int read_header(FILE* f, const char* path, Header* hd) {
    if (fread(hd, sizeof(Header), 1, f) != 1 || memcmp(hd->magic, "PESS0001", 8) != 0) {
        set_error("%s is not a PESS0001 sample file", path);
        return PE_ERR_IO;
    }
    if (hd->n < 0 || hd->c <= 0 || hd->h <= 0 || hd->w <= 0 || hd->label_dim < 0 || hd->c > 4096 || hd->h > 65536 || hd->w > 65536) {
        set_error("%s: implausible header", path);
        return PE_ERR_INVALID;
    }
    return PE_OK;
}
}  // namespace

extern "C" {

int pe_samples_write(const char* path, const float* data_hwc, const float* labels, int64_t n, int channels, int height, int width,
                     int label_dim) {
    PE_REQUIRE(path && n >= 0 && channels > 0 && height > 0 && width > 0 && label_dim >= 0, "bad arguments");
    PE_REQUIRE(n == 0 || (data_hwc && (labels || label_dim == 0)), "null pointer");
    FILE* f = fopen(path, "wb");
    if (!f) { set_error("cannot write %s", path); return PE_ERR_IO; }
    Header hd;
    memcpy(hd.magic, "PESS0001", 8);
    hd.n = n; hd.c = channels; hd.h = height; hd.w = width; hd.label_dim = label_dim;
    bool ok = fwrite(&hd, sizeof(hd), 1, f) == 1;
    const size_t img = (size_t)channels * height * width;
    std::vector<float> tmp(img);
    for (int64_t i = 0; i < n && ok; ++i) {          /* hdf5Handler::write: HWC -> CHW, one sample at a time */
        const float* s = data_hwc + (size_t)i * img;
        for (int ch = 0; ch < channels; ++ch)
            for (int y = 0; y < height; ++y)
                for (int x = 0; x < width; ++x) tmp[((size_t)ch * height + y) * width + x] = s[((size_t)y * width + x) * channels + ch];
        ok = fwrite(tmp.data(), 4, img, f) == img;
    }
    if (ok && n > 0 && label_dim > 0) ok = fwrite(labels, 4, (size_t)n * label_dim, f) == (size_t)n * label_dim;
    ok = fclose(f) == 0 && ok;
    if (!ok) { set_error("short write to %s", path); return PE_ERR_IO; }
    return PE_OK;
}

int pe_samples_info(const char* path, int64_t* n, int* channels, int* height, int* width, int* label_dim) {
    PE_REQUIRE(path != nullptr, "null path");
    FILE* f = fopen(path, "rb");
    if (!f) { set_error("cannot open %s", path); return PE_ERR_IO; }
    Header hd;
    const int st = read_header(f, path, &hd);
    fclose(f);
    if (st != PE_OK) return st;
    if (n) *n = hd.n;
    if (channels) *channels = hd.c;
    if (height) *height = hd.h;
    if (width) *width = hd.w;
    if (label_dim) *label_dim = hd.label_dim;
    return PE_OK;
}

int pe_samples_read(const char* path, int64_t first, int64_t count, float* out_data_hwc, float* out_labels) {
    PE_REQUIRE(path && first >= 0 && count >= 0, "bad arguments");
    FILE* f = fopen(path, "rb");
    if (!f) { set_error("cannot open %s", path); return PE_ERR_IO; }
    Header hd;
    int st = read_header(f, path, &hd);
    if (st == PE_OK && first + count > hd.n) { set_error("%s holds %lld samples, asked for [%lld, %lld)", path, (long long)hd.n, (long long)first, (long long)(first + count)); st = PE_ERR_INVALID; }
    const size_t img = (size_t)hd.c * hd.h * hd.w;
    std::vector<float> tmp(st == PE_OK ? img : 0);
    bool ok = st == PE_OK;
    if (ok && out_data_hwc && count > 0) {
        ok = fseek(f, (long)(sizeof(Header) + (size_t)first * img * 4), SEEK_SET) == 0;
        for (int64_t i = 0; i < count && ok; ++i) {   /* hdf5Handler::read: CHW -> HWC */
            ok = fread(tmp.data(), 4, img, f) == img;
            float* d = out_data_hwc + (size_t)i * img;
            for (int ch = 0; ch < hd.c && ok; ++ch)
                for (int y = 0; y < hd.h; ++y)
                    for (int x = 0; x < hd.w; ++x) d[((size_t)y * hd.w + x) * hd.c + ch] = tmp[((size_t)ch * hd.h + y) * hd.w + x];
        }
    }
    if (ok && out_labels && count > 0 && hd.label_dim > 0) {
        ok = fseek(f, (long)(sizeof(Header) + (size_t)hd.n * img * 4 + (size_t)first * hd.label_dim * 4), SEEK_SET) == 0 &&
             fread(out_labels, 4, (size_t)count * hd.label_dim, f) == (size_t)count * hd.label_dim;
    }
    fclose(f);
    if (st != PE_OK) return st;
    if (!ok) { set_error("%s: truncated", path); return PE_ERR_IO; }
    return PE_OK;
}

}  // extern "C"

/* ------------------------------------------------------------------ PLY meshes ---- */
/* Model::loadModel (reference src/model.cpp:296-352) without cv::viz / VTK: a minimal PLY reader (ascii and
 * binary_little_endian; triangles only, as the reference asserts at :318) that yields what the reference keeps of a mesh -
 * points, per-vertex colours (127 grey when the file has none, /255 as at :359), per-vertex texture coordinates, faces,
 * centroid.  Host code; the arrays feed pe_render_views[_tex]. */
struct pe_mesh {
    std::vector<float> v, c, uv;
    std::vector<int32_t> f;
    bool has_color = false, has_uv = false;
    float centroid[3] = {0, 0, 0};
};

namespace {

struct PlyProp { std::string name, type, count_type; bool list = false; };
struct PlyElem { std::string name; long long count = 0; std::vector<PlyProp> props; };

int ply_type_size(const std::string& t) {
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}

/* one scalar of `type` from a binary stream, as double */
bool ply_read_bin(std::istream& in, const std::string& t, double* out) {
    unsigned char b[8];
    const int n = ply_type_size(t);
    if (n == 0 || !in.read(reinterpret_cast<char*>(b), n)) return false;
    if (t == "char" || t == "int8") { *out = (signed char)b[0]; return true; }
    if (t == "uchar" || t == "uint8") { *out = b[0]; return true; }
    if (t == "short" || t == "int16") { int16_t x; memcpy(&x, b, 2); *out = x; return true; }
    if (t == "ushort" || t == "uint16") { uint16_t x; memcpy(&x, b, 2); *out = x; return true; }
    if (t == "int" || t == "int32") { int32_t x; memcpy(&x, b, 4); *out = x; return true; }
    if (t == "uint" || t == "uint32") { uint32_t x; memcpy(&x, b, 4); *out = x; return true; }
    if (t == "float" || t == "float32") { float x; memcpy(&x, b, 4); *out = x; return true; }
    double x; memcpy(&x, b, 8); *out = x;
    return true;
}

}  // namespace

extern "C" {

int pe_mesh_load_ply(const char* path, pe_mesh** out) {
    PE_REQUIRE(path && out, "null pointer");
    *out = nullptr;
    std::ifstream in(path, std::ios::binary);
    if (!in) { pe::set_error("cannot open %s", path); return PE_ERR_IO; }
    std::string line;
    if (!std::getline(in, line) || line.substr(0, 3) != "ply") { pe::set_error("%s is not a PLY file", path); return PE_ERR_IO; }
    bool ascii = false, have_format = false;
    std::vector<PlyElem> elems;
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ls(line);
        std::string kw;
        ls >> kw;
        if (kw == "format") {
            std::string fmt;
            ls >> fmt;
            if (fmt == "ascii") ascii = true;
            else if (fmt != "binary_little_endian") { pe::set_error("%s: PLY format '%s' unsupported", path, fmt.c_str()); return PE_ERR_UNSUPPORTED; }
            have_format = true;
        } else if (kw == "element") {
            PlyElem e;
            ls >> e.name >> e.count;
            elems.push_back(e);
        } else if (kw == "property") {
            if (elems.empty()) { pe::set_error("%s: property before element", path); return PE_ERR_IO; }
            PlyProp p;
            std::string t;
            ls >> t;
            if (t == "list") { p.list = true; ls >> p.count_type >> p.type >> p.name; }
            else { p.type = t; ls >> p.name; }
            if (ply_type_size(p.type) == 0 || (p.list && ply_type_size(p.count_type) == 0)) {
                pe::set_error("%s: unknown PLY type in '%s'", path, line.c_str());
                return PE_ERR_IO;
            }
            elems.back().props.push_back(p);
        } else if (kw == "end_header") break;
    }
    if (!have_format) { pe::set_error("%s: no format line", path); return PE_ERR_IO; }
    pe_mesh* m = new pe_mesh();
    auto fail = [&](const char* why) { pe::set_error("%s: %s", path, why); delete m; return PE_ERR_IO; };
    for (const PlyElem& e : elems) {
        const bool is_v = e.name == "vertex", is_f = e.name == "face";
        int ix = -1, iy = -1, iz = -1, ir = -1, ig = -1, ib = -1, iu = -1, iv = -1;
        if (is_v)
            for (size_t p = 0; p < e.props.size(); ++p) {
                const std::string& n = e.props[p].name;
                if (n == "x") ix = (int)p; else if (n == "y") iy = (int)p; else if (n == "z") iz = (int)p;
                else if (n == "red" || n == "r" || n == "diffuse_red") ir = (int)p;
                else if (n == "green" || n == "g" || n == "diffuse_green") ig = (int)p;
                else if (n == "blue" || n == "b" || n == "diffuse_blue") ib = (int)p;
                else if (n == "u" || n == "s" || n == "texture_u" || n == "texture_s") iu = (int)p;
                else if (n == "v" || n == "t" || n == "texture_v" || n == "texture_t") iv = (int)p;
            }
        if (is_v && (ix < 0 || iy < 0 || iz < 0)) return fail("vertex element without x / y / z");
        if (is_v) {
            m->has_color = ir >= 0 && ig >= 0 && ib >= 0;
            m->has_uv = iu >= 0 && iv >= 0;
            m->v.reserve((size_t)e.count * 3);
        }
        std::vector<double> vals(e.props.size());
        for (long long r = 0; r < e.count; ++r) {
            for (size_t p = 0; p < e.props.size(); ++p) {
                const PlyProp& pr = e.props[p];
                if (!pr.list) {
                    if (ascii) { if (!(in >> vals[p])) return fail("truncated (ascii scalar)"); }
                    else if (!ply_read_bin(in, pr.type, &vals[p])) return fail("truncated (binary scalar)");
                    continue;
                }
                double cnt = 0;
                if (ascii) { if (!(in >> cnt)) return fail("truncated (list count)"); }
                else if (!ply_read_bin(in, pr.count_type, &cnt)) return fail("truncated (list count)");
                const bool idx_list = is_f && (pr.name == "vertex_indices" || pr.name == "vertex_index");
                if (idx_list && (int)cnt != 3) return fail("a face is not a triangle (the reference asserts triangles, src/model.cpp:318)");
                for (int j = 0; j < (int)cnt; ++j) {
                    double x = 0;
                    if (ascii) { if (!(in >> x)) return fail("truncated (list)"); }
                    else if (!ply_read_bin(in, pr.type, &x)) return fail("truncated (list)");
                    if (idx_list) m->f.push_back((int32_t)x);
                }
            }
            if (is_v) {
                m->v.push_back((float)vals[ix]); m->v.push_back((float)vals[iy]); m->v.push_back((float)vals[iz]);
                if (m->has_color) {
                    /* uchar colours /255 (src/model.cpp:359); float colour properties are taken as they are */
                    const bool bytes = ply_type_size(e.props[ir].type) == 1;
                    for (int ci : {ir, ig, ib}) m->c.push_back(bytes ? (float)vals[ci] / 255.0f : (float)vals[ci]);
                }
                if (m->has_uv) { m->uv.push_back((float)vals[iu]); m->uv.push_back((float)vals[iv]); }
            }
        }
    }
    const size_t nv = m->v.size() / 3;
    if (nv == 0) return fail("no vertices");
    if (!m->has_color) m->c.assign(nv * 3, 127.0f / 255.0f);       /* src/model.cpp:325 */
    for (int32_t i : m->f)
        if (i < 0 || (size_t)i >= nv) return fail("face index out of range");
    double cs[3] = {0, 0, 0};
    for (size_t i = 0; i < nv; ++i) for (int d = 0; d < 3; ++d) cs[d] += m->v[3 * i + d];
    for (int d = 0; d < 3; ++d) m->centroid[d] = (float)(cs[d] / (double)nv);
    *out = m;
    return PE_OK;
}

int pe_mesh_info(const pe_mesh* m, int* n_vertices, int* n_faces, int* has_colors, int* has_uv, float* centroid3) {
    PE_REQUIRE(m != nullptr, "null mesh");
    if (n_vertices) *n_vertices = (int)(m->v.size() / 3);
    if (n_faces) *n_faces = (int)(m->f.size() / 3);
    if (has_colors) *has_colors = m->has_color ? 1 : 0;
    if (has_uv) *has_uv = m->has_uv ? 1 : 0;
    if (centroid3) for (int d = 0; d < 3; ++d) centroid3[d] = m->centroid[d];
    return PE_OK;
}

int pe_mesh_data(const pe_mesh* m, float* vertices, float* colors, float* uv, int32_t* faces) {
    PE_REQUIRE(m != nullptr, "null mesh");
    if (vertices) memcpy(vertices, m->v.data(), m->v.size() * 4);
    if (colors) memcpy(colors, m->c.data(), m->c.size() * 4);
    if (uv) { PE_REQUIRE(m->has_uv, "mesh has no texture coordinates"); memcpy(uv, m->uv.data(), m->uv.size() * 4); }
    if (faces) memcpy(faces, m->f.data(), m->f.size() * 4);
    return PE_OK;
}

void pe_mesh_destroy(pe_mesh* m) { delete m; }

}  // extern "C"
