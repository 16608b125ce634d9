"""On-disk formats, CPU only: the PESS0001 sample file (hdf5Handler's two datasets behind a header)
and the .caffemodel -> PEW1 converter (own protobuf wire reader, exercised with a model encoded here)."""
import importlib.util
import os
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "tools", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_sample_file_round_trip_and_layout(capi, tmp_path):
    rng = np.random.default_rng(1)
    data = rng.random((5, 6, 4, 7), dtype=np.float32)          # n, H, W, C
    labels = rng.random((5, 8), dtype=np.float32)
    path = str(tmp_path / "s.pess")
    capi.samples_write(path, data, labels)
    assert capi.samples_info(path) == {"n": 5, "channels": 7, "height": 6, "width": 4, "label_dim": 8}
    d, l = capi.samples_read(path)
    assert np.array_equal(d, data) and np.array_equal(l, labels)
    d2, l2 = capi.samples_read(path, 2, 2)
    assert np.array_equal(d2, data[2:4]) and np.array_equal(l2, labels[2:4])
    # on disk: hdf5Handler::write's layout, i.e. the "data" dataset is [N][C][H][W], then "label"
    raw = open(path, "rb").read()
    assert raw[:8] == b"PESS0001" and struct.unpack("<qiiii", raw[8:32]) == (5, 7, 6, 4, 8)
    disk = np.frombuffer(raw, "<f4", 5 * 7 * 6 * 4, 32).reshape(5, 7, 6, 4)
    assert np.array_equal(disk, data.transpose(0, 3, 1, 2))
    assert np.array_equal(np.frombuffer(raw, "<f4", 40, 32 + disk.nbytes).reshape(5, 8), labels)
    # the h5 dump tool writes the same bytes from the dataset arrays
    _tool("h5_to_pess").write_pess(str(tmp_path / "t.pess"), data.transpose(0, 3, 1, 2), labels)
    assert open(str(tmp_path / "t.pess"), "rb").read() == raw
    with pytest.raises(capi.PoseError):
        capi.samples_read(path, 4, 3)
    with pytest.raises(capi.PoseError):
        capi.samples_info(str(tmp_path / "missing.pess"))
    open(str(tmp_path / "bad.pess"), "wb").write(b"not a sample file at all, really not")
    with pytest.raises(capi.PoseError):
        capi.samples_info(str(tmp_path / "bad.pess"))


# ---- a tiny protobuf encoder, only to fabricate a .caffemodel for the converter -------------------------------------
def _vi(v):
    out = b""
    while True:
        b = v & 0x7F
        v >>= 7
        out += bytes([b | (0x80 if v else 0)])
        if not v:
            return out


def _ld(num, payload):
    return _vi(num << 3 | 2) + _vi(len(payload)) + payload


def _blob(arr, packed=True, legacy_dims=False):
    a = np.asarray(arr, "<f4")
    out = b""
    if legacy_dims:
        dims = ([1] * (4 - a.ndim) + list(a.shape))
        for i, d in enumerate(dims):
            out += _vi((i + 1) << 3 | 0) + _vi(d)
    else:
        out += _ld(7, _ld(1, b"".join(_vi(d) for d in a.shape)))
    if packed:
        out += _ld(5, a.tobytes())
    else:
        out += b"".join(_vi(5 << 3 | 5) + struct.pack("<f", float(x)) for x in a.ravel())
    return out


def test_caffemodel_converter(capi, po, synth, tmp_path):
    from tests import netdefs
    tool = _tool("caffemodel_to_pew")
    ref = po.parse_prototxt(netdefs.prototxt("manifold_rgbd_16"))          # PReLU net: weight, bias and slope blobs
    blobs = synth.make_blobs(ref.blob_names, ref.blob_shapes, 11, mode=1)
    # group the blobs per layer as Caffe stores them: conv / ip layers hold (weight, bias), PReLU layers one slope blob
    layers, i = [], 0
    while i < len(blobs):
        name = ref.blob_names[i]
        take = 2 if (i + 1 < len(blobs) and len(ref.blob_shapes[i]) >= 2 and len(ref.blob_shapes[i + 1]) == 1 and
                     ref.blob_shapes[i + 1][0] == ref.blob_shapes[i][0]) else 1
        layers.append((name.split(".")[0], [np.asarray(b, np.float32).reshape(s) for b, s in zip(blobs[i:i + take], ref.blob_shapes[i:i + take])]))
        i += take
    model = _ld(1, b"manifold") + _ld(100, _ld(1, b"data") + _ld(2, b"Input"))          # a layer without blobs is skipped
    for k, (name, arrs) in enumerate(layers):
        body = _ld(1, name.encode()) + _ld(2, b"Convolution")
        for j, a in enumerate(arrs):
            body += _ld(7, _blob(a, packed=(k + j) % 2 == 0, legacy_dims=(k % 3 == 2)))
        model += _ld(100, body)
    src, dst = tmp_path / "m.caffemodel", tmp_path / "m.pew"
    src.write_bytes(model)
    parsed = tool.parse_caffemodel(model)
    assert [n for n, _ in parsed] == [n for n, _ in layers]
    assert tool.main(["x", str(src), str(dst)]) == 0
    net = capi.Net(text=netdefs.prototxt("manifold_rgbd_16"))
    net.load_weights(str(dst))
    got = net.get_blobs()
    assert len(got) == len(blobs)
    for g, b in zip(got, blobs):
        assert np.array_equal(np.asarray(g).ravel(), np.asarray(b, np.float32).ravel())
    # legacy V1 container (NetParameter.layers = 2, name 4, blobs 6) parses the same
    v1 = b"".join(_ld(2, _ld(4, n.encode()) + b"".join(_ld(6, _blob(a)) for a in arrs)) for n, arrs in layers)
    flat = [a for _, bl in tool.parse_caffemodel(v1) for a, _ in bl]
    assert all(np.array_equal(x, np.asarray(b, np.float32).ravel()) for x, b in zip(flat, blobs)) and len(flat) == len(blobs)


def test_ply_reader_round_trip(capi, tmp_path):
    """pe_mesh_load_ply = what Model::loadModel keeps of cv::viz::Mesh::load (reference src/model.cpp:296-352): points,
    colours / 255 (127 grey when absent), per-vertex texture coordinates, triangle faces, centroid; ascii and binary"""
    from tests import scenes
    rng = np.random.default_rng(1)
    v, f = scenes.icosphere(2, 0.07)
    col = rng.integers(0, 256, (len(v), 3))
    uv = rng.random((len(v), 2)).astype(np.float32)
    for binary in (False, True):
        p = tmp_path / f"m{int(binary)}.ply"
        scenes.write_ply(p, v, f, col, uv, binary, uv_names=("texture_u", "texture_v") if binary else ("u", "v"),
                   index_name="vertex_index" if binary else "vertex_indices")
        m = capi.load_ply(p)
        assert np.array_equal(m["vertices"], v) and np.array_equal(m["faces"], f)
        assert np.array_equal(m["colors"], (col / 255.0).astype(np.float32)) and m["has_colors"]
        assert np.array_equal(m["uv"], uv)
        np.testing.assert_allclose(m["centroid"], v.astype(np.float64).mean(0), atol=1e-7)
    p = tmp_path / "plain.ply"
    scenes.write_ply(p, v, f)
    m = capi.load_ply(p)
    assert m["uv"] is None and not m["has_colors"] and np.all(m["colors"] == np.float32(127 / 255))
    # errors: missing file, not a PLY, a quad face, an index out of range, a truncated body
    with pytest.raises(capi.PoseError) as e:
        capi.load_ply(tmp_path / "nope.ply")
    assert e.value.code == capi.PE_ERR_IO
    (tmp_path / "bad.ply").write_text("solid x\n")
    with pytest.raises(capi.PoseError):
        capi.load_ply(tmp_path / "bad.ply")
    (tmp_path / "quad.ply").write_text("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
                                        "element face 1\nproperty list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n1 1 0\n0 1 0\n4 0 1 2 3\n")
    with pytest.raises(capi.PoseError) as e:
        capi.load_ply(tmp_path / "quad.ply")
    assert "triangle" in str(e.value)
    scenes.write_ply(tmp_path / "oor.ply", v, np.array([[0, 1, len(v)]], np.int32))
    with pytest.raises(capi.PoseError):
        capi.load_ply(tmp_path / "oor.ply")
    data = (tmp_path / "m1.ply").read_bytes()
    (tmp_path / "cut.ply").write_bytes(data[:len(data) // 2])
    with pytest.raises(capi.PoseError):
        capi.load_ply(tmp_path / "cut.ply")


def test_pose_labels_equal_oracle(capi, po):
    """pe_pose_labels (host arithmetic of the template producer) == the oracle's restatement of the reference's label
    (src/datasetmanager.cpp:484-491, 523-530), bit for bit, on view-sphere poses and random rotations (all four branches)"""
    from scipy.spatial.transform import Rotation
    views = capi.sphere_views(2, (0.4, 0.35, 1.1), (-45, 15, 45), False)
    extra = np.zeros((300, 3, 4), np.float32)
    extra[:, :, :3] = Rotation.random(300, random_state=8).as_matrix().astype(np.float32)
    extra[:, :, 3] = np.random.default_rng(2).standard_normal((300, 3)).astype(np.float32)
    poses = np.concatenate([views, extra])
    got = capi.pose_labels(poses, 11)
    want = np.stack([po.pose_label(p, 11) for p in poses])
    assert np.array_equal(got, want)
    assert capi.pose_labels(poses[:0], 0).shape == (0, 8)
