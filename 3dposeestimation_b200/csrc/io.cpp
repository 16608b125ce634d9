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
