"""ctypes front-end of the CPU oracle (oracle/libpose_oracle.so) + a Caffe-prototxt reader.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (3dposeestimation_b200/) never imports it.

The prototxt reader here is an independent Python implementation of the TEST-phase subset
(reference network/manifold*.prototxt); the product has its own C++ parser and the tests
compare the two.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpose_oracle.so")

CONV, POOL, ACT, FC = 1, 2, 3, 4
ACT_KINDS = {"ReLU": 1, "ELU": 2, "PReLU": 3}


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


class _Layer(C.Structure):
    _fields_ = [("kind", C.c_int), ("a", C.c_int), ("b", C.c_int), ("blob0", C.c_int), ("blob1", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp, ip, u32p = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint32)
        L.orc_forward.restype = C.c_int
        L.orc_forward.argtypes = [C.POINTER(_Layer), C.c_int, C.POINTER(fp), ip, C.c_int, C.c_int, C.c_int,
                                  fp, C.c_int64, C.c_int, fp, C.c_int]
        L.orc_forward_stage.restype = C.c_int
        L.orc_forward_im2col.argtypes = [C.POINTER(_Layer), C.c_int, C.POINTER(fp), ip, C.c_int, C.c_int, C.c_int,
                                         fp, C.c_int64, C.c_int, C.c_int, fp]
        L.orc_forward_im2col.restype = C.c_int
        L.orc_forward_im2col_mt.argtypes = [C.POINTER(_Layer), C.c_int, C.POINTER(fp), ip, C.c_int, C.c_int, C.c_int,
                                            fp, C.c_int64, C.c_int, C.c_int, fp, C.c_int]
        L.orc_forward_im2col_mt.restype = C.c_int
        L.orc_forward_stage.argtypes = [C.POINTER(_Layer), C.c_int, C.POINTER(fp), ip, C.c_int, C.c_int, C.c_int,
                                        fp, C.c_int, C.c_int, fp, C.c_int]
        L.orc_knn_l2.restype = None
        L.orc_knn_l2.argtypes = [fp, C.c_int64, fp, C.c_int64, C.c_int, C.c_int, ip, fp, fp, C.c_int]
        L.orc_l2_distance.restype = C.c_float
        L.orc_l2_distance.argtypes = [fp, fp, C.c_int]
        L.orc_binarize.restype = None
        L.orc_binarize.argtypes = [fp, C.c_int64, C.c_int, u32p]
        L.orc_knn_hamming.restype = None
        L.orc_knn_hamming.argtypes = [u32p, C.c_int64, u32p, C.c_int64, C.c_int, C.c_int, ip, ip, C.c_int]
        L.orc_merge_topk_f32.restype = None
        L.orc_merge_topk_f32.argtypes = [ip, fp, C.c_int, C.c_int64, C.c_int, ip, fp]
        L.orc_merge_topk_i32.restype = None
        L.orc_merge_topk_i32.argtypes = [ip, ip, C.c_int, C.c_int64, C.c_int, ip, ip]
        L.orc_angular_distance.restype = C.c_float
        L.orc_angular_distance_atan2.argtypes = [fp, fp]
        L.orc_angular_distance_atan2.restype = C.c_float
        L.orc_set_angular_form.argtypes = [C.c_int]
        L.orc_set_angular_form.restype = None
        L.orc_angular_distance.argtypes = [fp, fp]
        L.orc_rot22.restype = C.c_float
        L.orc_rot22.argtypes = [fp]
        L.orc_histogram.restype = None
        L.orc_histogram.argtypes = [ip, C.c_int64, C.c_int, fp, fp, C.c_int, ip, fp, C.c_int, fp, fp]
        L.orc_confusion.restype = None
        L.orc_confusion.argtypes = [ip, C.c_int64, C.c_int, fp, fp, ip, C.c_int, C.c_int, fp]
        L.orc_knn_bootstrap.restype = None
        L.orc_knn_bootstrap.argtypes = [ip, C.c_int64, C.c_int, C.c_int, ip, ip]
        L.orc_knn_accuracy.restype = None
        L.orc_knn_accuracy.argtypes = [ip, ip, C.c_int64, C.c_int, fp, fp]
        L.orc_depth2normals.restype = None
        L.orc_depth2normals.argtypes = [fp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, fp]
        L.orc_patch_rect.restype = C.c_int
        L.orc_patch_rect.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, ip, ip, ip, ip]
        L.orc_sample_patch.restype = C.c_int
        L.orc_sample_patch.argtypes = [C.POINTER(C.c_uint8), fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                       C.c_float, C.c_float, fp]
        L.orc_init_sphere.restype = C.c_int
        L.orc_init_sphere.argtypes = [C.c_int, fp, C.c_int]
        L.orc_view_pose.restype = None
        L.orc_view_pose.argtypes = [fp, C.c_float, C.c_float, fp]
        L.orc_render.restype = None
        L.orc_project_centroid.argtypes = [fp, fp, fp, ip]
        L.orc_project_centroid.restype = C.c_float
        L.orc_pose_label.argtypes = [fp, C.c_int, fp]
        L.orc_pose_label.restype = None
        L.orc_render_tex.argtypes = [fp, fp, fp, C.c_int, ip, C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_int, fp, fp, C.c_int, C.c_int,
                                     C.c_float, C.c_float, C.POINTER(C.c_uint8), fp]
        L.orc_render_tex.restype = None
        L.orc_render.argtypes = [fp, fp, C.c_int, ip, C.c_int, fp, fp, C.c_int, C.c_int, C.c_float, C.c_float, C.POINTER(C.c_uint8), fp]
        i64p = C.POINTER(C.c_int64)
        L.orc_max_sim_tmpl.restype = None
        L.orc_max_sim_tmpl.argtypes = [fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, ip]
        L.orc_triplet_indices.restype = None
        L.orc_triplet_indices.argtypes = [fp, ip, ip, ip, ip, i64p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_uint64, ip, ip]
        L.orc_gather_batch.restype = None
        L.orc_gather_batch.argtypes = [fp, fp, fp, ip, C.c_int, C.c_int, C.c_int64, C.c_uint64, C.c_int, fp]
        L.orc_elevation.restype = C.c_float
        L.orc_elevation.argtypes = [fp]
        L.orc_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


# --------------------------------------------------------------------------- prototxt ---
_TOKEN = re.compile(r'"[^"]*"|\'[^\']*\'|[{}:]|[^\s{}:]+')


def _parse_block(tokens, pos):
    """Caffe text format: `key: value` and `key { ... }`; repeated keys accumulate."""
    out: dict[str, list] = {}
    while pos < len(tokens) and tokens[pos] != "}":
        key = tokens[pos]
        pos += 1
        if tokens[pos] == ":":
            pos += 1
        if tokens[pos] == "{":
            val, pos = _parse_block(tokens, pos + 1)
            assert tokens[pos] == "}", "unbalanced braces"
            pos += 1
        else:
            val = tokens[pos].strip("\"'")
            pos += 1
        out.setdefault(key, []).append(val)
    return out, pos


def parse_text(text: str) -> dict:
    text = re.sub(r"#[^\n]*", "", text)
    tokens = _TOKEN.findall(text)
    out, pos = _parse_block(tokens, 0)
    assert pos == len(tokens), "trailing tokens"
    return out


@dataclass
class NetDef:
    batch: int
    channels: int
    height: int
    width: int
    sel: list                       # input-blob channels fed to the first conv
    layers: list = field(default_factory=list)      # (kind, name, a, b, blob0, blob1)
    blob_names: list = field(default_factory=list)  # e.g. "conv1.w"
    blob_shapes: list = field(default_factory=list)
    desc_dim: int = 0

    def describe(self) -> str:
        inv = {v: k for k, v in ACT_KINDS.items()}
        lines = []
        for kind, name, a, b, _, _ in self.layers:
            if kind == CONV:
                lines.append(f"conv {name} {a} {b}")
            elif kind == POOL:
                lines.append(f"pool {name} {a} {b}")
            elif kind == ACT:
                lines.append(f"act {name} {inv[a]}")
            else:
                lines.append(f"fc {name} {a}")
        return "\n".join(lines)


def _pooled(n, k, s):
    return int(np.ceil(np.float32(n - k) / np.float32(s))) + 1


def parse_prototxt(text: str) -> NetDef:
    """TEST-phase view of a manifold*.prototxt (layers with include{phase:TRAIN} dropped)."""
    root = parse_text(text)
    dims = [int(d) for d in root["input_shape"][0]["dim"]]
    n, c, h, w = dims
    data_name = root["input"][0]
    chan = {data_name: list(range(c))}   # blob name -> channel list (pre-conv plumbing)
    net = NetDef(n, c, h, w, sel=[])
    cur_blob, shape = None, None
    for layer in root.get("layer", []):
        inc = layer.get("include", [])
        if any(i.get("phase", [""])[0] == "TRAIN" for i in inc):
            continue
        ltype, name = layer["type"][0], layer["name"][0]
        bottoms, tops = layer.get("bottom", []), layer.get("top", [])
        if ltype == "Slice":
            sp = layer["slice_param"][0]
            assert int(sp.get("axis", ["1"])[0]) == 1
            src = chan[bottoms[0]]
            pts = [0] + [int(p) for p in sp.get("slice_point", [])] + [len(src)]
            assert len(pts) - 1 == len(tops)
            for i, t in enumerate(tops):
                chan[t] = src[pts[i]:pts[i + 1]]
        elif ltype == "Concat":
            assert int(layer.get("concat_param", [{}])[0].get("axis", ["1"])[0]) == 1
            chan[tops[0]] = sum((chan[b] for b in bottoms), [])
        elif ltype == "Silence":
            pass
        elif ltype == "Convolution":
            p = layer["convolution_param"][0]
            nf, k = int(p["num_output"][0]), int(p["kernel_size"][0])
            assert int(p.get("stride", ["1"])[0]) == 1 and int(p.get("pad", ["0"])[0]) == 0
            if cur_blob is None:
                net.sel = list(chan[bottoms[0]])
                shape = (len(net.sel), h, w)
            else:
                assert bottoms[0] == cur_blob, f"{name}: bottom {bottoms[0]} != {cur_blob}"
            b0 = len(net.blob_names)
            net.blob_names += [f"{name}.w", f"{name}.b"]
            net.blob_shapes += [(nf, shape[0], k, k), (nf,)]
            net.layers.append((CONV, name, nf, k, b0, b0 + 1))
            shape = (nf, shape[1] - k + 1, shape[2] - k + 1)
            cur_blob = tops[0]
        elif ltype == "Pooling":
            p = layer["pooling_param"][0]
            assert p.get("pool", ["MAX"])[0] == "MAX" and bottoms[0] == cur_blob
            k, s = int(p["kernel_size"][0]), int(p.get("stride", ["1"])[0])
            net.layers.append((POOL, name, k, s, -1, -1))
            shape = (shape[0], _pooled(shape[1], k, s), _pooled(shape[2], k, s))
            cur_blob = tops[0]
        elif ltype in ACT_KINDS:
            assert bottoms[0] == cur_blob
            b0 = -1
            if ltype == "PReLU":
                b0 = len(net.blob_names)
                net.blob_names.append(f"{name}.a")
                net.blob_shapes.append((shape[0],))
            net.layers.append((ACT, name, ACT_KINDS[ltype], 0, b0, -1))
            cur_blob = tops[0]
        elif ltype == "InnerProduct":
            assert bottoms[0] == cur_blob
            nout = int(layer["inner_product_param"][0]["num_output"][0])
            fan_in = int(np.prod(shape))
            b0 = len(net.blob_names)
            net.blob_names += [f"{name}.w", f"{name}.b"]
            net.blob_shapes += [(nout, fan_in), (nout,)]
            net.layers.append((FC, name, nout, 0, b0, b0 + 1))
            shape = (nout, 1, 1)
            cur_blob = tops[0]
        else:
            raise ValueError(f"unsupported TEST-phase layer type {ltype}")
    net.desc_dim = int(np.prod(shape))
    return net


# ----------------------------------------------------------------------------- forward ---
class OracleNet:
    def __init__(self, net: NetDef, blobs: list):
        assert len(blobs) == len(net.blob_shapes)
        self.net = net
        self.blobs = [_f32(b).reshape(-1) for b in blobs]
        for b, shp in zip(self.blobs, net.blob_shapes):
            assert b.size == int(np.prod(shp)), (b.size, shp)
        self._layers = (_Layer * len(net.layers))(*[_Layer(k, a, b, b0, b1) for k, _, a, b, b0, b1 in net.layers])
        self._blob_ptrs = (C.POINTER(C.c_float) * len(self.blobs))(*[_p(b, C.c_float) for b in self.blobs])
        self._sel = _i32(net.sel)

    def forward(self, patches_hwc: np.ndarray, threads: int = 0) -> np.ndarray:
        """computeDescriptors: patches n x H x W x C (HWC) -> n x D."""
        x = _f32(patches_hwc)
        n, h, w, c = x.shape
        assert (h, w, c) == (self.net.height, self.net.width, self.net.channels)
        out = np.empty((n, self.net.desc_dim), np.float32)
        d = lib().orc_forward(self._layers, len(self.net.layers), self._blob_ptrs, _p(self._sel, C.c_int32),
                              len(self.net.sel), h, w, _p(x, C.c_float), n, c, _p(out, C.c_float), threads)
        if d != self.net.desc_dim:
            raise RuntimeError(f"oracle forward failed ({d})")
        return out

    def forward_im2col(self, patches_hwc: np.ndarray, batch: int | None = None, threads: int = 1) -> np.ndarray:
        """the same forward the way Caffe-CPU runs it (im2col + SGEMM, `batch` samples per pass); threads = 1: a build without
        a threaded BLAS, threads = 0 / n: the samples of a batch spread over all / n OpenMP threads"""
        x = _f32(patches_hwc)
        n, h, w, c = x.shape
        assert (h, w, c) == (self.net.height, self.net.width, self.net.channels)
        out = np.empty((n, self.net.desc_dim), np.float32)
        d = lib().orc_forward_im2col_mt(self._layers, len(self.net.layers), self._blob_ptrs, _p(self._sel, C.c_int32),
                                        len(self.net.sel), h, w, _p(x, C.c_float), n, c, int(batch or self.net.batch), _p(out, C.c_float),
                                        int(threads))
        if d != self.net.desc_dim:
            raise RuntimeError(f"oracle im2col forward failed ({d})")
        return out

    def stage(self, patch_hwc: np.ndarray, stage: int) -> np.ndarray:
        x = _f32(patch_hwc)
        h, w, c = x.shape
        buf = np.empty(1 << 20, np.float32)
        cnt = lib().orc_forward_stage(self._layers, len(self.net.layers), self._blob_ptrs, _p(self._sel, C.c_int32),
                                      len(self.net.sel), h, w, _p(x, C.c_float), c, stage, _p(buf, C.c_float), buf.size)
        if cnt < 0:
            raise RuntimeError("oracle stage failed")
        return buf[:cnt].copy()


# ---------------------------------------------------------------------------- matching ---
def knn_l2(q, t, k, threads=0, want_gap=False):
    q, t = _f32(q), _f32(t)
    nq, dim = q.shape
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.float32)
    gap = np.empty(nq, np.float32)
    lib().orc_knn_l2(_p(q, C.c_float), nq, _p(t, C.c_float), t.shape[0], dim, k, _p(idx, C.c_int32),
                     _p(dist, C.c_float), _p(gap, C.c_float), threads)
    return (idx, dist, gap) if want_gap else (idx, dist)


def l2_distance(a, b):
    a, b = _f32(a), _f32(b)
    return float(lib().orc_l2_distance(_p(a, C.c_float), _p(b, C.c_float), a.size))


def binarize(desc):
    d = _f32(desc)
    n, dim = d.shape
    bits = np.empty((n, (dim + 31) // 32), np.uint32)
    lib().orc_binarize(_p(d, C.c_float), n, dim, _p(bits, C.c_uint32))
    return bits


def knn_hamming(q, t, k, threads=0):
    q = np.ascontiguousarray(q, np.uint32)
    t = np.ascontiguousarray(t, np.uint32)
    nq, words = q.shape
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.int32)
    lib().orc_knn_hamming(_p(q, C.c_uint32), nq, _p(t, C.c_uint32), t.shape[0], words, k, _p(idx, C.c_int32),
                          _p(dist, C.c_int32), threads)
    return idx, dist


def merge_topk(idx, dist):
    """idx/dist: n_shards x nq x k candidate lists holding GLOBAL indices."""
    idx = _i32(idx)
    s, nq, k = idx.shape
    oi = np.empty((nq, k), np.int32)
    if np.issubdtype(np.asarray(dist).dtype, np.integer):
        dist = _i32(dist)
        od = np.empty((nq, k), np.int32)
        lib().orc_merge_topk_i32(_p(idx, C.c_int32), _p(dist, C.c_int32), s, nq, k, _p(oi, C.c_int32), _p(od, C.c_int32))
    else:
        dist = _f32(dist)
        od = np.empty((nq, k), np.float32)
        lib().orc_merge_topk_f32(_p(idx, C.c_int32), _p(dist, C.c_float), s, nq, k, _p(oi, C.c_int32), _p(od, C.c_float))
    return oi, od


def angular_distance(qa, qb):
    qa, qb = _f32(qa), _f32(qb)
    return float(lib().orc_angular_distance(_p(qa, C.c_float), _p(qb, C.c_float)))


def angular_distance_atan2(qa, qb):
    """Eigen 3.3+ form of angularDistance (radians)"""
    a, b = _f32(qa), _f32(qb)
    return float(lib().orc_angular_distance_atan2(_p(a, C.c_float), _p(b, C.c_float)))


def set_angular_form(form: int):
    """0: Eigen 3.2 acos form (default, what the CUDA path implements), 1: Eigen 3.3+ atan2 form"""
    lib().orc_set_angular_form(int(form))


def rot22(q):
    q = _f32(q)
    return float(lib().orc_rot22(_p(q, C.c_float)))


def histogram(knn_idx, q_labels, t_labels, nr_tmpl_poses, rot_inv, bins, want_err=False):
    knn_idx, q_labels, t_labels = _i32(knn_idx), _f32(q_labels), _f32(t_labels)
    rot_inv, bins = _i32(rot_inv), _f32(bins)
    nq, k = knn_idx.shape
    out = np.empty(bins.size + 2, np.float32)
    err = np.empty(max(nq, 1), np.float32)
    lib().orc_histogram(_p(knn_idx, C.c_int32), nq, k, _p(q_labels, C.c_float), _p(t_labels, C.c_float),
                        nr_tmpl_poses, _p(rot_inv, C.c_int32), _p(bins, C.c_float), bins.size,
                        _p(out, C.c_float), _p(err, C.c_float))
    return (out, err[:nq]) if want_err else out


def confusion(knn_idx, q_labels, t_labels, class_to_local, n_local, nr_test_poses):
    knn_idx, q_labels, t_labels = _i32(knn_idx), _f32(q_labels), _f32(t_labels)
    c2l = _i32(class_to_local)
    nq, k = knn_idx.shape
    out = np.empty((n_local, n_local), np.float32)
    lib().orc_confusion(_p(knn_idx, C.c_int32), nq, k, _p(q_labels, C.c_float), _p(t_labels, C.c_float),
                        _p(c2l, C.c_int32), n_local, nr_test_poses, _p(out, C.c_float))
    return out


def knn_bootstrap(knn3_idx, nr_query_poses, nr_tmpl_poses):
    knn3_idx = _i32(knn3_idx)
    nq = knn3_idx.shape[0]
    assert knn3_idx.shape[1] == 3
    out = np.empty((nq, 4), np.int32)
    ln = np.empty(nq, np.int32)
    lib().orc_knn_bootstrap(_p(knn3_idx, C.c_int32), nq, nr_query_poses, nr_tmpl_poses, _p(out, C.c_int32), _p(ln, C.c_int32))
    return out, ln


def knn_accuracy(max_sim_tmpl0, bootstrap, nr_query_poses):
    m, b = _i32(max_sim_tmpl0), _i32(bootstrap)
    intra, inter = C.c_float(), C.c_float()
    lib().orc_knn_accuracy(_p(m, C.c_int32), _p(b, C.c_int32), m.size, nr_query_poses, C.byref(intra), C.byref(inter))
    return intra.value, inter.value


# ------------------------------------------------------------------ patch extraction ---
def depth2normals(depth, fx, fy, ox=0.0, oy=0.0):
    """src/helper.cpp:53-103 on one rows x cols f32 depth map -> rows x cols x 3 normals"""
    d = _f32(depth)
    out = np.empty(d.shape + (3,), np.float32)
    lib().orc_depth2normals(_p(d, C.c_float), d.shape[0], d.shape[1], fx, fy, ox, oy, _p(out, C.c_float))
    return out


def patch_rect(rows, cols, cx, cy, z, fx, fy, m=0.2):
    v = [C.c_int32() for _ in range(4)]
    ok = lib().orc_patch_rect(rows, cols, int(cx), int(cy), z, fx, fy, m, *[C.byref(x) for x in v])
    return bool(ok), tuple(x.value for x in v)


def sample_patch(color_u8, depth, normals, cx, cy, z, fx, fy, m=0.2):
    """samplePatchWithScale (src/datasetmanager.cpp:299-346) -> 64 x 64 x 7 f32, or None where its asserts fail"""
    c = np.ascontiguousarray(color_u8, np.uint8)
    d, n = _f32(depth), _f32(normals)
    out = np.empty((64, 64, 7), np.float32)
    st = lib().orc_sample_patch(_p(c, C.c_uint8), _p(d, C.c_float), _p(n, C.c_float), d.shape[0], d.shape[1], int(cx), int(cy),
                                z, fx, fy, m, _p(out, C.c_float))
    return out if st == 0 else None


# ------------------------------------------------------------------ template renderer ---
def init_sphere(depth):
    """SphereRenderer::initSphere (src/sphere.cpp:186-226) -> [n, 3] unit vectors in the reference's order"""
    cap = 10 * 4 ** depth + 2
    out = np.empty((cap, 3), np.float32)
    n = lib().orc_init_sphere(depth, _p(out, C.c_float), cap)
    return out[:n].copy()


def view_pose(vertex, scale, angle_deg):
    """SphereRenderer::createTransformation -> row-major 3x4 [R|t]"""
    v = _f32(vertex)
    out = np.empty((3, 4), np.float32)
    lib().orc_view_pose(_p(v, C.c_float), scale, angle_deg, _p(out, C.c_float))
    return out


def sphere_views(depth, scale, rotation, skip_lower_hemi=True):
    """the pose list of SphereRenderer::createViews (src/sphere.cpp:69-94): scale / rotation = (first, step, last)"""
    f32 = np.float32
    def frange(a):
        out, i = [], f32(a[0])
        while i <= f32(a[2]):
            out.append(float(i)); i = f32(i + f32(a[1]))
        return out
    poses = []
    for sc in frange(scale):
        for pos in init_sphere(depth):
            if pos[2] < 0 and skip_lower_hemi:
                continue
            for rot in frange(rotation):
                poses.append(view_pose(pos, sc, rot))
    return np.stack(poses) if poses else np.zeros((0, 3, 4), np.float32)


def render(vertices, colors, faces, cam, rt, rows=480, cols=640, z_near=0.001, z_far=4.0, uv=None, tex=None):
    """Painter::paint with RealWorldCamera + Model -> (BGR u8 [rows, cols, 3], depth f32 [rows, cols], 0 = background);
    uv [nv, 2] + tex [th, tw, 3] BGR u8: textured mesh (GL_MODULATE, GL_REPEAT, GL_LINEAR)"""
    v, c, f = _f32(vertices), _f32(colors), _i32(faces)
    k, p = _f32(cam), _f32(rt)
    col = np.empty((rows, cols, 3), np.uint8)
    dep = np.empty((rows, cols), np.float32)
    if uv is not None and tex is not None:
        t, w = _f32(uv), np.ascontiguousarray(tex, np.uint8)
        lib().orc_render_tex(_p(v, C.c_float), _p(c, C.c_float), _p(t, C.c_float), v.shape[0], _p(f, C.c_int32), f.shape[0],
                             _p(w, C.c_uint8), w.shape[0], w.shape[1], _p(k, C.c_float), _p(p, C.c_float),
                             rows, cols, z_near, z_far, _p(col, C.c_uint8), _p(dep, C.c_float))
    else:
        lib().orc_render(_p(v, C.c_float), _p(c, C.c_float), v.shape[0], _p(f, C.c_int32), f.shape[0], _p(k, C.c_float), _p(p, C.c_float),
                         rows, cols, z_near, z_far, _p(col, C.c_uint8), _p(dep, C.c_float))
    return col, dep


def project_centroid(cam, rt, centroid=None):
    """(center_x, center_y, z) of extractRealSamplesPaul: cam * centroid / z truncated to int, z = centroid z"""
    k, p = _f32(cam), _f32(rt).reshape(12)
    c = _f32(centroid) if centroid is not None else None
    xy = np.empty(2, np.int32)
    z = lib().orc_project_centroid(_p(k, C.c_float), _p(p, C.c_float), _p(c, C.c_float) if c is not None else None, _p(xy, C.c_int32))
    return int(xy[0]), int(xy[1]), float(z)


def pose_label(rt, index):
    """label of a rendered template view: {index, Quaternionf(R).coeffs() xyzw, pose.inverse().translation()}"""
    p = _f32(rt).reshape(12)
    out = np.empty(8, np.float32)
    lib().orc_pose_label(_p(p, C.c_float), int(index), _p(out, C.c_float))
    return out


# ----------------------------------------------------------------------- triplet miner ---
def max_sim_tmpl(tmpl_labels, train_labels, n_objects, inplane):
    """datasetManager::computeMaxSimTmpl[Inplane] (src/datasetmanager.cpp:687-783) -> [n_objects * n_train, 2]"""
    tl, ql = _f32(tmpl_labels).reshape(-1, 8), _f32(train_labels).reshape(-1, 8)
    n_t, n_q = tl.shape[0] // n_objects, ql.shape[0] // n_objects
    out = np.empty((n_objects * n_q, 2), np.int32)
    lib().orc_max_sim_tmpl(_p(tl, C.c_float), _p(ql, C.c_float), n_objects, n_t, n_q, int(bool(inplane)), _p(out, C.c_int32))
    return out


def elevation(label):
    l = _f32(label)
    return lib().orc_elevation(_p(l, C.c_float))


def triplet_indices(tmpl_labels, max_sim, knn_table, knn_len, rot_inv, neg_offsets, n_objects, nr_training_poses, batch_size,
                    triplet_size, it, seed):
    """networkSolver::buildBatch's index selection (src/networksolver.cpp:52-186) -> ([n_triplets, 5], flags)"""
    tl = _f32(tmpl_labels).reshape(-1, 8)
    n_t = tl.shape[0] // n_objects
    ms = _i32(max_sim)
    kt = _i32(knn_table) if knn_table is not None else None
    kl = _i32(knn_len) if knn_len is not None else None
    ri = _i32(rot_inv)
    no = np.ascontiguousarray(neg_offsets, np.int64) if neg_offsets is not None else None
    n = batch_size // triplet_size
    out, flag = np.empty((n, 5), np.int32), np.empty(n, np.int32)
    lib().orc_triplet_indices(_p(tl, C.c_float), _p(ms, C.c_int32), _p(kt, C.c_int32) if kt is not None else None,
                              _p(kl, C.c_int32) if kl is not None else None, _p(ri, C.c_int32),
                              _p(no, C.c_int64) if no is not None else None, n_objects, n_t, nr_training_poses, batch_size,
                              triplet_size, it, seed, _p(out, C.c_int32), _p(flag, C.c_int32))
    return out, flag


def gather_batch(tmpl_pool, train_pool, neg_pool, idx, first_linear_id, seed, random_background=0):
    """the batch_caffe fill of buildBatchQueue (src/networksolver.cpp:32-40) -> [n_triplets * 5, C, 64, 64]"""
    tp, qp = _f32(tmpl_pool), _f32(train_pool)
    npool = _f32(neg_pool) if neg_pool is not None else None
    ix = _i32(idx)
    ch = tp.shape[-1]
    out = np.empty((ix.shape[0] * 5, ch, 64, 64), np.float32)
    lib().orc_gather_batch(_p(tp, C.c_float), _p(qp, C.c_float), _p(npool, C.c_float) if npool is not None else None,
                           _p(ix, C.c_int32), ix.shape[0], ch, first_linear_id, seed, random_background, _p(out, C.c_float))
    return out


def num_threads():
    return lib().orc_num_threads()
