"""A fixed-seed slice of the randomised shape sweep (tests/fuzz_gpu.py; the full sweep is `python -m tests.fuzz_gpu 500 <seed>`):
dims 1..128, k 1..16, template counts around the block / tile / chunk boundaries, both matcher paths, Hamming codes of 32 /
64 / 128 bits, odd batch sizes of every net, random neighbour tables through the evaluation kernels - all against the oracle."""
import numpy as np
import pytest

from tests import fuzz_gpu, netdefs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tensor", [1, 0], ids=["tensor", "simt"])
def test_knn_shape_sweep(gpu, po, tensor):
    rng = np.random.default_rng(20240607 + tensor)
    gpu.set_option("knn_tensor", tensor)
    try:
        for c in fuzz_gpu.knn_cases(rng, 60 if tensor else 25):
            try:
                fuzz_gpu.run_knn(gpu, po, c)
            except AssertionError as e:
                raise AssertionError(f"{c}: {e}") from e
        for _ in range(20):
            fuzz_gpu.run_hamming(gpu, po, rng)
    finally:
        gpu.set_option("knn_tensor", 1)


def test_forward_batch_sizes_and_eval_tables(gpu, po, synth):
    rng = np.random.default_rng(77)
    for name in netdefs.VARIANTS:
        fuzz_gpu.run_forward(gpu, po, synth, rng, name)
    for _ in range(40):
        fuzz_gpu.run_eval(gpu, po, synth, rng)


def test_render_and_patch_random_scenes(gpu, po):
    """triangle soups (slivers, degenerate and huge faces, faces behind the camera / through the near plane), odd frame sizes,
    textures; patch cuts at random centres and depths incl. impossible ones: bit-exact against the oracle"""
    rng = np.random.default_rng(5150)
    covered = [fuzz_gpu.run_render(gpu, po, rng) for _ in range(24)]
    assert np.mean(covered) > 0.05                   # the scenes do put triangles on screen
    accepted = [fuzz_gpu.run_patch(gpu, po, rng) for _ in range(12)]
    assert 0.2 < np.mean(accepted) < 0.95            # both accepted and rejected cuts occur


def test_every_entry_point_survives_null_arguments_with_a_device(gpu):
    """tests/nullfuzz_cabi.py on a box with a GPU: with a device present the calls get past the "no CUDA device" answer,
    so the argument checks themselves are what returns"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "tests.nullfuzz_cabi"], cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "all returned" in r.stdout, r.stdout[-800:] + r.stderr[-800:]
