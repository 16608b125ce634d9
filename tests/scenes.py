"""Seeded synthetic RGB-D frames for the patch-extraction tests (numpy only; shared by tests/golden/make_golden_patch.py,
the CPU oracle tests and the GPU parity tests).  A frame is a tilted background plane with a sphere and a box in front
of it, a rectangle of missing depth (zeros, as a real sensor gives), and smooth random colour."""
import struct

import numpy as np

CAM = dict(fx=572.4114, fy=573.57043, ox=325.2611, oy=242.04899)     # LineMOD intrinsics (640 x 480)


def make_frame(seed, rows=480, cols=640):
    rng = np.random.default_rng(seed)
    sx, sy = cols / 640.0, rows / 480.0
    yy, xx = np.mgrid[0:rows, 0:cols].astype(np.float32)
    depth = (1.2 + 0.0005 / sx * (xx - cols / 2) + 0.0003 / sy * (yy - rows / 2)).astype(np.float32)
    cx, cy, rad = cols * (0.4 + 0.2 * rng.random()), rows * (0.4 + 0.2 * rng.random()), 0.17 * rows
    r2 = (xx - cx) ** 2 + (yy - cy) ** 2
    sphere = 0.85 - 0.25 * np.sqrt(np.maximum(0, rad * rad - r2)) / rad
    depth = np.where(r2 < rad * rad, sphere, depth).astype(np.float32)
    bx, by = int(cols * 0.72), int(rows * 0.2)
    depth[by:by + rows // 5, bx:bx + cols // 6] = 0.95
    depth[rows // 8:rows // 8 + rows // 16, cols // 10:cols // 10 + cols // 12] = 0.0       # sensor holes
    depth += (rng.random((rows, cols)).astype(np.float32) - 0.5) * 0.002                      # sensor noise
    depth[depth < 0.01] = 0.0
    color = rng.integers(0, 256, (rows // 8 + 2, cols // 8 + 2, 3)).astype(np.float32)
    color = np.kron(color, np.ones((8, 8, 1), np.float32))[:rows, :cols]                      # blocky texture
    color = np.clip(color + 20 * np.sin(xx / 7.0)[..., None], 0, 255).astype(np.uint8)
    return color, depth.astype(np.float32), (int(cx), int(cy))


def cam_for(rows, cols):
    """intrinsics scaled to a smaller frame"""
    return dict(fx=CAM["fx"] * cols / 640.0, fy=CAM["fy"] * rows / 480.0, ox=CAM["ox"] * cols / 640.0, oy=CAM["oy"] * rows / 480.0)


# ------------------------------------------------------------------ meshes for the renderer tests ---
def icosphere(level, radius=1.0):
    """(vertices [n, 3], faces [m, 3]) of a subdivided icosahedron (its own subdivision: shared midpoints by dictionary)"""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    v = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
         (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(level):
        cache, nf = {}, []
        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return (np.array(v) * radius).astype(np.float32), np.array(f, np.int32)


def cube(half=0.05):
    v = np.array([[x, y, z] for x in (-half, half) for y in (-half, half) for z in (-half, half)], np.float32)
    f = np.array([[0, 1, 3], [0, 3, 2], [4, 6, 7], [4, 7, 5], [0, 4, 5], [0, 5, 1], [2, 3, 7], [2, 7, 6], [0, 2, 6], [0, 6, 4], [1, 5, 7], [1, 7, 3]], np.int32)
    c = (v / (2 * half) + 0.5).astype(np.float32)            # colour = position in the unit cube
    return v, c, f


def ground_plane():
    """a ground plane through the camera's near plane (camera at the origin, identity pose): Y = 0.1, X in [-2, 2],
    Z from -1 (behind the camera) to 3; vertex colour (Z + 1) / 4 in all channels; two triangles sharing the diagonal"""
    v = np.array([[-2, 0.1, -1], [2, 0.1, -1], [2, 0.1, 3], [-2, 0.1, 3]], np.float32)
    c = np.repeat(((v[:, 2:3] + 1) / 4), 3, axis=1).astype(np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    return v, c, f


def textured_quad(seed=5, th=37, tw=53):
    """fronto-parallel quad at Z = 1 covering the frame, uv running past [0, 1] (GL_REPEAT), random BGR texture"""
    rng = np.random.default_rng(seed)
    v = np.array([[-0.7, -0.5, 1], [0.7, -0.5, 1], [0.7, 0.5, 1], [-0.7, 0.5, 1]], np.float32)
    uv = np.array([[-0.3, -0.2], [1.6, -0.2], [1.6, 1.4], [-0.3, 1.4]], np.float32)
    c = np.array([[1, 1, 1], [1, 0.5, 1], [0.25, 1, 1], [1, 1, 0.5]], np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    tex = rng.integers(0, 256, (th, tw, 3)).astype(np.uint8)
    return v, c, uv, f, tex


def write_ply(path, v, f, colors=None, uv=None, binary=False, uv_names=("u", "v"), index_name="vertex_indices"):
    hdr = ["ply", "format binary_little_endian 1.0" if binary else "format ascii 1.0", "comment written by the test",
           f"element vertex {len(v)}", "property float x", "property float y", "property float z"]
    if colors is not None:
        hdr += ["property uchar red", "property uchar green", "property uchar blue"]
    if uv is not None:
        hdr += [f"property float {uv_names[0]}", f"property float {uv_names[1]}"]
    hdr += [f"element face {len(f)}", f"property list uchar int {index_name}", "end_header"]
    with open(path, "wb") as fh:
        fh.write(("\n".join(hdr) + "\n").encode())
        for i in range(len(v)):
            if binary:
                fh.write(struct.pack("<3f", *v[i]))
                if colors is not None:
                    fh.write(struct.pack("<3B", *colors[i]))
                if uv is not None:
                    fh.write(struct.pack("<2f", *uv[i]))
            else:
                row = [repr(float(x)) for x in v[i]]
                if colors is not None:
                    row += [str(int(x)) for x in colors[i]]
                if uv is not None:
                    row += [repr(float(x)) for x in uv[i]]
                fh.write((" ".join(row) + "\n").encode())
        for t in f:
            if binary:
                fh.write(struct.pack("<B3i", 3, *t))
            else:
                fh.write(("3 " + " ".join(str(int(x)) for x in t) + "\n").encode())
