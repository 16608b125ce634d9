"""Template renderer oracle (oracle/oracle_render.c) against analytic scenes.  The reference renders through OpenGL,
which is unavailable on this box and has no CPU twin, so parity for this row is UNPINNED: these tests pin the restated
pipeline maths (projection, coverage, 1/Z interpolation, perspective-correct colour, depth test, view sphere) to
closed-form answers instead.  CPU only."""
import numpy as np

from tests import scenes

CAM = np.array([572.4114, 573.57043, 325.2611, 242.04899], np.float32)


def eye_pose(tz):
    rt = np.zeros((3, 4), np.float32)
    rt[:, :3] = np.eye(3)
    rt[2, 3] = tz
    return rt


def test_view_sphere_counts_and_poses(po):
    for depth, n, upper in ((0, 12, 8), (1, 42, 25), (2, 162, 89)):
        s = po.init_sphere(depth)
        assert s.shape == (n, 3) and int((s[:, 2] >= 0).sum()) == upper
        assert np.abs(np.linalg.norm(s, axis=1) - 1).max() < 2e-7
        assert len({tuple(np.round(p, 5)) for p in s}) == n                    # no duplicates survive the linear search
    views = po.sphere_views(2, (0.4, 0.2, 0.8), (-45, 15, 45))
    assert views.shape == (3 * 89 * 7, 3, 4)
    for rt in views[::37]:
        r = rt[:, :3].astype(np.float64)
        assert np.abs(r @ r.T - np.eye(3)).max() < 1e-6 and abs(np.linalg.det(r) - 1) < 1e-6
        assert rt[0, 3] == 0 and rt[1, 3] == 0 and rt[2, 3] in (np.float32(0.4), np.float32(0.6), np.float32(0.8))
    # the camera looks at the object centre along +z: the view direction of pose (vertex, angle 0) is -vertex
    s = po.init_sphere(1)
    rt = po.view_pose(s[7], 0.5, 0.0)
    np.testing.assert_allclose(rt[2, :3], -s[7], atol=1e-6)


def test_single_triangle_depth_colour_and_coverage(po):
    z0 = 0.8
    v = np.array([[-0.2, -0.15, 0.0], [0.25, -0.1, 0.1], [0.0, 0.2, -0.1]], np.float32)       # slanted: depths 0.8, 0.9, 0.7
    c = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
    f = np.array([[0, 1, 2]], np.int32)
    col, dep = po.render(v, c, f, CAM, eye_pose(z0))
    # analytic: ray through the pixel centre against the triangle's plane, barycentrics in 3-D
    vc = v.astype(np.float64) + [0, 0, z0]
    n = np.cross(vc[1] - vc[0], vc[2] - vc[0])
    ys, xs = np.mgrid[0:480, 0:640]
    d = np.stack([(xs + 0.5 - CAM[2]) / CAM[0], (ys + 0.5 - CAM[3]) / CAM[1], np.ones_like(xs, np.float64)], -1)
    t = (vc[0] @ n) / (d @ n)
    p = d * t[..., None]
    def bary(p):
        m = np.stack([vc[1] - vc[0], vc[2] - vc[0]], 1)
        uv = np.linalg.lstsq(m, (p - vc[0]).reshape(-1, 3).T, rcond=None)[0].T.reshape(p.shape[:2] + (2,))
        return 1 - uv[..., 0] - uv[..., 1], uv[..., 0], uv[..., 1]
    b0, b1, b2 = bary(p)
    inside = (b0 > 2e-3) & (b1 > 2e-3) & (b2 > 2e-3)
    outside = (b0 < -2e-3) | (b1 < -2e-3) | (b2 < -2e-3)
    assert inside.sum() > 20000
    assert np.all(dep[inside] > 0) and np.all(dep[outside] == 0) and np.all(col[outside] == 0)
    # vertices are snapped to 1/256 px: depth agrees with the exact plane to ~1e-5 relative
    np.testing.assert_allclose(dep[inside], p[..., 2][inside], rtol=3e-5)
    want = np.stack([b2, b1, b0], -1) * 255          # BGR of (r, g, b) = (b0, b1, b2) weights
    assert np.abs(col[inside].astype(np.float64) - want[inside]).max() <= 1.0


def test_watertight_quad_no_culling_and_depth_test(po):
    # two triangles sharing a diagonal: every pixel centre of the quad is covered exactly once (no cracks, no double hits)
    v = np.array([[-0.1, -0.1, 0], [0.1, -0.1, 0], [0.1, 0.1, 0], [-0.1, 0.1, 0]], np.float32)
    white = np.ones((4, 3), np.float32)
    quad = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    col, dep = po.render(v, white, quad, CAM, eye_pose(0.6))
    hit = dep > 0
    ys, xs = np.nonzero(hit)
    assert hit[ys.min():ys.max() + 1, xs.min():xs.max() + 1].all()                  # a solid rectangle: no crack on the diagonal
    assert np.all(col[hit] == 255) and np.allclose(dep[hit], 0.6, rtol=1e-6)
    # reversed winding renders the same (the reference never enables culling)
    col2, dep2 = po.render(v, white, quad[:, ::-1], CAM, eye_pose(0.6))
    assert np.array_equal(dep > 0, dep2 > 0) and np.allclose(dep, dep2, rtol=1e-6) and np.array_equal(col, col2)   # same coverage
    # a nearer red triangle in front of the quad wins where it covers, whatever the draw order
    v2 = np.concatenate([v, np.array([[-0.05, -0.05, -0.1], [0.05, -0.05, -0.1], [0.0, 0.05, -0.1]], np.float32)])
    c2 = np.concatenate([white, np.array([[1, 0, 0]] * 3, np.float32)])
    for faces in (np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6]], np.int32), np.array([[4, 5, 6], [0, 1, 2], [0, 2, 3]], np.int32)):
        col3, dep3 = po.render(v2, c2, faces, CAM, eye_pose(0.6))
        front = np.isclose(dep3, 0.5, rtol=1e-6)
        assert front.sum() > 1000 and np.all(col3[front] == [0, 0, 255])            # BGR red
        assert np.array_equal(dep3 > 0, hit)
    # beyond the far plane / behind the near plane nothing is drawn
    assert not po.render(v, white, quad, CAM, eye_pose(4.5))[1].any()
    assert not po.render(v, white, quad, CAM, eye_pose(-0.5))[1].any()


def test_sphere_depth_matches_ray_casting(po):
    radius, dist = 0.1, 0.7
    v, f = scenes.icosphere(4, radius)
    c = np.full(v.shape, 0.5, np.float32)
    col, dep = po.render(v, c, f, CAM, eye_pose(dist))
    ys, xs = np.mgrid[0:480, 0:640]
    d = np.stack([(xs + 0.5 - CAM[2]) / CAM[0], (ys + 0.5 - CAM[3]) / CAM[1], np.ones_like(xs, np.float64)], -1)
    # |t d - c|^2 = r^2 with c = (0, 0, dist)
    a = (d * d).sum(-1); b = -2 * d[..., 2] * dist; cc = dist * dist - radius * radius
    disc = b * b - 4 * a * cc
    hit = disc > 0
    z = np.where(hit, (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a), 0)
    core = disc > 0.3 * disc.max()                      # away from the silhouette, where the facets hug the sphere
    assert core.sum() > 5000 and np.all(dep[core] > 0)
    assert np.abs(dep[core] - z[core]).max() < 4e-4      # facet sag of a level-4 icosphere of radius 0.1 (~0.2 mm)
    assert np.all(dep[~hit & (disc < -0.02 * disc.max())] == 0)
    assert np.all(col[core] == 128)                     # 0.5 * 255 + 0.5 -> 128
