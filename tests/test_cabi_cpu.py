"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/pose_b200.h
declares, and its host logic (prototxt reader, blob bookkeeping, fillers, weight files, label
generator, error returns) behaves.  No compute entry point is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests import netdefs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "pose_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.findall(r"PE_API\s+[\w\s\*]+?\b(pe_\w+)\s*\(", text)


def test_library_exports_every_declared_symbol(capi):
    names = header_symbols()
    assert len(names) >= 55 and len(set(names)) == len(names)
    lib = capi.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/pose_b200.h but not exported"
    assert set(names) == set(capi.SIGNATURES), set(names) ^ set(capi.SIGNATURES)
    assert lib.pe_version() == 100


def test_host_mirror_library_loads_and_exports_its_doors(capi):
    """libpose_b200_host.so (the compiled C++ mirror of networkEvaluator / networkSolver) links and loads without a GPU"""
    src = open(os.path.join(ROOT, "3dposeestimation_b200", "csrc", "host_cabi.cpp")).read()
    doors = re.findall(r"HOST_API\s+[\w\s\*]+?\b(pe_host_\w+)\s*\(", src)
    assert set(doors) == set(capi.HOST_SIGNATURES) and len(doors) >= 7
    H = capi.host_lib()
    for n in doors:
        assert hasattr(H, n)
    # no device here: the mirror reports the library's error instead of computing anything on the host
    if capi.device_count() == 0:
        z = np.zeros((2, 64, 64, 7), np.float32)
        lab = np.zeros((2, 8), np.float32)
        rot = np.zeros(2, np.int32)
        out = np.zeros((10, 7, 64, 64), np.float32)
        idx, ms = np.zeros((2, 5), np.int32), np.zeros((2, 2), np.int32)
        vp = lambda a: a.ctypes.data_as(capi.vp)
        code = H.pe_host_build_batch(vp(z), vp(lab), 2, 1, vp(z), vp(lab), 1, 7, vp(rot), 1, None, None, 10, 0, 1, 0, vp(out), vp(idx),
                                     vp(ms), None)
        assert code != 0 and b"pose_b200" in H.pe_host_last_error()


def test_producer_mirror_host_side(capi, tmp_path):
    """include/datasetmanager_b200.hpp without a GPU: the view count of SphereRenderer::createViews' pose list, Model::loadModel
    reading a PLY, and - with no device - the producers reporting the library's error instead of producing anything"""
    from tests import scenes
    H = capi.host_lib()
    vp = lambda a: a.ctypes.data_as(capi.vp)
    cam9 = np.array([572.4, 0, 325.3, 0, 573.6, 242.0, 0, 0, 1], np.float32)
    sc, ro = np.float32([0.4, 1.1, 1.0]), np.float32([-45, 15, 45])
    n = C.c_int64()
    assert H.pe_host_create_views(None, vp(cam9), 1, vp(sc), vp(ro), 1, None, None, None, 0, C.byref(n)) == 0
    assert n.value == len(capi.sphere_views(1, sc, ro)) and n.value % 7 == 0
    patches, labels = np.zeros((4, 64, 64, 7), np.float32), np.zeros((4, 8), np.float32)
    code = H.pe_host_create_synth_samples(b"/nonexistent.ply", vp(cam9), 0, 0, 0, 0, None, None, None, 0, 0, vp(patches), vp(labels), 4,
                                          C.byref(n))
    assert code != 0 and b"nonexistent.ply" in H.pe_host_last_error()
    if capi.device_count() == 0:
        v, f = scenes.icosphere(1, 0.05)
        scenes.write_ply(tmp_path / "m.ply", v, f)
        code = H.pe_host_create_synth_samples(str(tmp_path / "m.ply").encode(), vp(cam9), 0, 0, 0, 0, None, None, None, 0, 0, vp(patches),
                                              vp(labels), 4, C.byref(n))
        assert code != 0 and b"pose_b200" in H.pe_host_last_error() and not patches.any()


def test_every_entry_point_survives_null_arguments():
    """SURVEY 8b "Errors": the C ABI returns status codes, never faults - every entry point called with null pointers and
    zero sizes, in a process of its own (tests/nullfuzz_cabi.py); the GPU suite repeats this with a device present"""
    import subprocess
    import sys
    r = subprocess.run([sys.executable, "-m", "tests.nullfuzz_cabi"], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "all returned" in r.stdout, r.stdout[-600:] + r.stderr[-600:]


def test_example_program_builds_and_reports_errors():
    """examples/evaluate (the evaluation half of the reference's main.cpp on the mirror headers) is built with the library
    and answers a missing config / no argument the way main.cpp does, without a GPU"""
    import subprocess
    exe = os.path.join(ROOT, "examples", "evaluate")
    assert os.path.exists(exe)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "Specify config file as argument" in r.stderr          # main.cpp:16-20
    r = subprocess.run([exe, "/nonexistent/config.ini"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "cannot open config" in r.stderr


def test_header_cites_reference_interfaces():
    text = open(os.path.join(ROOT, "include", "pose_b200.h")).read()
    for cite in ("src/networkevaluator.cpp:9-54", "src/networkevaluator.cpp:269-355", "src/networkevaluator.cpp:199-245",
                 "src/networksolver.cpp:474-504", "src/networkevaluator.cpp:153-175", "src/networksolver.cpp:286"):
        assert cite in text


@pytest.mark.parametrize("name", list(netdefs.VARIANTS))
def test_prototxt_reader_matches_python_reader(capi, po, name):
    ref = po.parse_prototxt(netdefs.prototxt(name))
    for net in (capi.Net(text=netdefs.prototxt(name)), capi.Net(named=name)):
        assert (net.batch, net.channels, net.height, net.width, net.desc_dim) == \
            (ref.batch, ref.channels, ref.height, ref.width, ref.desc_dim)
        assert net.sel == ref.sel == netdefs.EXPECTED[name][0]
        assert net.describe() == ref.describe()
        info = net.blob_info()
        assert [n for n, _ in info] == ref.blob_names == netdefs.expected_blob_names(name)
        assert [c for _, c in info] == [int(np.prod(s)) for s in ref.blob_shapes]
        net.close()


@pytest.mark.skipif(not os.path.isdir("/root/reference/network"), reason="reference checkout not present")
@pytest.mark.parametrize("name", list(netdefs.VARIANTS))
def test_reads_the_reference_prototxt_files(capi, po, name):
    path = f"/root/reference/network/{name}.prototxt"
    ref = po.parse_prototxt(open(path).read())
    net = capi.Net(path=path)
    assert net.describe() == ref.describe() and net.sel == ref.sel and net.desc_dim == ref.desc_dim
    assert [n for n, _ in net.blob_info()] == ref.blob_names


def test_fillers_match_python_generator(capi, po, synth):
    for vi, name in enumerate(("manifold_rgbd_16", "manifold_rgbdnor_16_bin", "manifold")):
        ref = po.parse_prototxt(netdefs.prototxt(name))
        net = capi.Net(named=name)
        for mode in (0, 1):
            net.fill(1234 + vi, mode)
            want = synth.make_blobs(ref.blob_names, ref.blob_shapes, 1234 + vi, mode=mode)
            got = net.get_blobs()
            for w, g, nm in zip(want, got, ref.blob_names):
                assert np.array_equal(w.reshape(-1), g), nm
        w1 = net.get_blobs()[0]
        bound = np.sqrt(3.0 / (ref.blob_shapes[0][1] * 64))      # Caffe xavier: fan_in = C*8*8
        assert np.abs(w1).max() <= bound and np.abs(w1).max() > 0.9 * bound


def test_set_get_save_load(capi, tmp_path):
    net = capi.Net(named="manifold_rgbd_16")
    net.fill(7, 1)
    blobs = net.get_blobs()
    path = tmp_path / "w.pew"
    net.save_weights(path)
    other = capi.Net(named="manifold_rgbd_16")
    other.load_weights(path)
    assert all(np.array_equal(a, b) for a, b in zip(blobs, other.get_blobs()))
    blobs[3][:] = 0.5
    other.set_blobs(blobs)
    assert np.all(other.get_blobs()[3] == 0.5)
    wrong = capi.Net(named="manifold")                         # different blob list
    with pytest.raises(capi.PoseError) as e:
        wrong.load_weights(path)
    assert e.value.code == capi.PE_ERR_INVALID
    with pytest.raises(capi.PoseError) as e:
        other.load_weights(tmp_path / "missing.pew")
    assert e.value.code == capi.PE_ERR_IO
    with pytest.raises(capi.PoseError):
        other.set_blobs([np.zeros(3, np.float32)])


def test_error_returns(capi):
    with pytest.raises(capi.PoseError) as e:
        capi.Net(text="input: \"data\" layer { name: \"x\" ")
    assert e.value.code == capi.PE_ERR_PARSE
    with pytest.raises(capi.PoseError) as e:
        capi.Net(named="no_such_net")
    assert e.value.code == capi.PE_ERR_INVALID
    # a valid Caffe net outside the manifold family (conv1 with 8 filters)
    text = netdefs.prototxt("manifold_rgbdnor_16").replace("num_output: 16\n    kernel_size: 8", "num_output: 8\n    kernel_size: 8")
    with pytest.raises(capi.PoseError) as e:
        capi.Net(text=text)
    assert e.value.code == capi.PE_ERR_UNSUPPORTED
    with pytest.raises(capi.PoseError) as e:
        capi.Net(text=netdefs.prototxt("manifold").replace('type: "ReLU"', 'type: "Sigmoid"'))
    assert e.value.code == capi.PE_ERR_PARSE
    with pytest.raises(capi.PoseError) as e:
        capi.Net(path="/nonexistent/net.prototxt")
    assert e.value.code == capi.PE_ERR_IO
    assert b"cannot open" in capi.lib().pe_last_error()


def test_synth_labels_match_python_generator(capi, synth):
    want = synth.make_labels(synth.BASE_SEED, 5, 33)
    got = capi.synth_labels(synth.BASE_SEED, 0, 5 * 33, 33)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-7)
    assert np.array_equal(got[:, 0], want[:, 0])
    part = capi.synth_labels(synth.BASE_SEED, 40, 20, 33)
    assert np.array_equal(part, got[40:60])
    np.testing.assert_allclose(np.linalg.norm(got[:, 1:5], axis=1), 1.0, atol=1e-6)


def test_no_cpu_fallback(capi):
    """without a device every compute entry point must fail loudly, never compute on the host"""
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    net = capi.Net(named="manifold_rgbdnor_16")
    net.fill(1, 1)
    with pytest.raises(capi.PoseError) as e:
        net.compute_descriptors(np.zeros((1, 64, 64, 7), np.float32))
    assert e.value.code == capi.PE_ERR_CUDA
    with pytest.raises(capi.PoseError) as e:
        capi.TemplateDB(desc=np.zeros((4, 16), np.float32))
    assert e.value.code == capi.PE_ERR_CUDA
