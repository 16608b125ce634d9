"""Seeded synthetic inputs for the evaluation path: patches, labels, weights.

Everything derives from a counter-based splitmix64 hash, so the arrays do not depend on numpy's
Generator implementation and the committed golden outputs (tests/golden/) stay valid.
Pure numpy; no CUDA.  The device-side generator (pe_synth_patches_dev) follows the same recipe
for bench-scale data but is not required to match bit for bit.
"""
from __future__ import annotations

import numpy as np

BASE_SEED = 20240607
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _key(seed: int, stream: int) -> np.uint64:
    return splitmix64(np.uint64(seed) ^ (np.uint64(stream) * np.uint64(0xD1342543DE82EF95) & _M64))


def uniform01(seed: int, stream: int, n: int, first: int = 0) -> np.ndarray:
    """n floats in [0,1): 24 high bits of splitmix64(key + counter)."""
    with np.errstate(over="ignore"):
        ctr = np.arange(first, first + n, dtype=np.uint64) + _key(seed, stream)
    return (splitmix64(ctr) >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)


def uniform(seed: int, stream: int, shape, lo: float, hi: float) -> np.ndarray:
    n = int(np.prod(shape))
    u = uniform01(seed, stream, n)
    return (np.float32(lo) + u * np.float32(hi - lo)).astype(np.float32).reshape(shape)


# ------------------------------------------------------------------------------ weights ---
def make_blobs(blob_names, blob_shapes, seed: int, mode: int = 1):
    """Caffe xavier weights U(-sqrt(3/fan_in), sqrt(3/fan_in)); mode 0: bias 0 / PReLU 0.25 (the
    prototxts' fillers); mode 1 (parity): bias U(-0.1,0.1), PReLU slopes U(-0.25,0.5)."""
    blobs = []
    for i, (name, shp) in enumerate(zip(blob_names, blob_shapes)):
        if name.endswith(".w"):
            s = float(np.sqrt(3.0 / int(np.prod(shp[1:]))))
            blobs.append(uniform(seed, 1000 + i, shp, -s, s))
        elif name.endswith(".b"):
            blobs.append(np.zeros(shp, np.float32) if mode == 0 else uniform(seed, 1000 + i, shp, -0.1, 0.1))
        else:
            blobs.append(np.full(shp, 0.25, np.float32) if mode == 0 else uniform(seed, 1000 + i, shp, -0.25, 0.5))
    return blobs


# ------------------------------------------------------------------------------ patches ---
def _tri(x):
    """triangle wave with period 1 and range [0,1]"""
    f = x - np.floor(x)
    return (np.float32(1.0) - np.abs(np.float32(2.0) * f - np.float32(1.0))).astype(np.float32)


def make_patches(seed: int, first: int, n: int, channels: int, poses_per_object: int,
                 noise_seed: int | None = None, noise: float = 0.0) -> np.ndarray:
    """n x 64 x 64 x channels float32 HWC patches for samples [first, first+n).

    Sample i is object i // poses_per_object, pose i % poses_per_object.  Channel order as the
    reference builds it (src/datasetmanager.cpp:327-343): 3 colour, depth, 3 normal.
      colour: 0.5*U[0,1) + 0.5*class pattern (triangle wave, class-dependent direction/frequency,
              pose-dependent phase)
      depth : 0 outside a centred disc of radius 16..28 px, U[0.3,0.7) inside
      normal: random unit vectors inside the disc, 0 outside
    noise > 0 adds U(-noise, noise) keyed by noise_seed (to derive "test" patches from templates).
    """
    H = W = 64
    out = np.zeros((n, H, W, channels), np.float32)
    yy, xx = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    for j in range(n):
        i = first + j
        obj, pose = divmod(i, poses_per_object)
        par = uniform01(seed, 7, 8, first=obj * 8)           # class parameters
        pp = uniform01(seed, 8, 4, first=i * 4)              # pose parameters
        r = np.float32(16.0) + np.float32(12.0) * pp[0]
        inside = ((yy - 31.5) ** 2 + (xx - 31.5) ** 2) <= r * r
        u = uniform01(seed, 9, H * W * 7, first=i * H * W * 7).reshape(H, W, 7)
        for c in range(min(channels, 3)):
            fx = np.float32(0.5) + np.float32(2.5) * par[2 * c]
            fy = np.float32(0.5) + np.float32(2.5) * par[2 * c + 1]
            pat = _tri((xx * fx + yy * fy) / np.float32(64.0) + pp[1] + np.float32(0.37 * c))
            out[j, :, :, c] = np.float32(0.5) * u[:, :, c] + np.float32(0.5) * pat
        if channels > 3:
            out[j, :, :, 3] = np.where(inside, np.float32(0.3) + np.float32(0.4) * u[:, :, 3], np.float32(0))
        if channels > 4:
            v = u[:, :, 4:7] * np.float32(2.0) - np.float32(1.0)
            nrm = np.sqrt((v * v).sum(-1, keepdims=True)) + np.float32(1e-6)
            out[j, :, :, 4:7] = np.where(inside[..., None], v / nrm, np.float32(0))[..., : channels - 4]
        if noise > 0:
            ns = uniform01(seed if noise_seed is None else noise_seed, 10, H * W * channels,
                           first=i * H * W * channels).reshape(H, W, channels)
            out[j] += (ns * np.float32(2.0) - np.float32(1.0)) * np.float32(noise)
    return out


# ------------------------------------------------------------------------------- labels ---
def random_quaternions(seed: int, stream: int, n: int, first: int = 0) -> np.ndarray:
    """uniform unit quaternions (Shoemake), Eigen coeffs() order x,y,z,w"""
    u = uniform01(seed, stream, 3 * n, first=3 * first).reshape(n, 3).astype(np.float64)
    a, b = np.sqrt(1.0 - u[:, 0]), np.sqrt(u[:, 0])
    t1, t2 = 2 * np.pi * u[:, 1], 2 * np.pi * u[:, 2]
    q = np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=1)
    return q.astype(np.float32)


def quat_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], axis=-1)


def make_labels(seed: int, n_objects: int, poses_per_object: int) -> np.ndarray:
    """(n_objects*poses) x 8 labels {class, qx,qy,qz,qw, 0,0,0} (include/datatypes.h:34-48)."""
    n = n_objects * poses_per_object
    lab = np.zeros((n, 8), np.float32)
    lab[:, 0] = np.repeat(np.arange(n_objects, dtype=np.float32), poses_per_object)
    lab[:, 1:5] = random_quaternions(seed, 20, n)
    return lab


def perturb_labels(labels: np.ndarray, seed: int, max_deg: float = 15.0) -> np.ndarray:
    """query labels: each quaternion rotated by a random rotation of at most max_deg degrees"""
    n = labels.shape[0]
    axis = uniform(seed, 21, (n, 3), -1.0, 1.0).astype(np.float64)
    axis /= np.linalg.norm(axis, axis=1, keepdims=True) + 1e-12
    ang = np.deg2rad(max_deg) * uniform01(seed, 22, n).astype(np.float64)
    dq = np.concatenate([axis * np.sin(ang / 2)[:, None], np.cos(ang / 2)[:, None]], axis=1)
    out = labels.copy()
    q = quat_mul(dq, labels[:, 1:5].astype(np.float64))
    out[:, 1:5] = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    return out


def rot_inv_pattern(n_classes: int) -> np.ndarray:
    """rotInv flags: 0 for ~80 % of classes, 1 for ~10 %, 2 for ~10 % (pattern of config.ini:10)."""
    r = np.zeros(n_classes, np.int32)
    r[2::10] = 1
    r[6::10] = 2
    return r
