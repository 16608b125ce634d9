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


def test_near_plane_clipping_of_a_ground_plane(po):
    """triangles with vertices behind the camera are clipped at Zc = near, not dropped (GL clips every primitive;
    reference near plane src/painter.cpp:9): depth and perspective-correct colour of the visible part are the analytic
    plane's, rows whose plane depth is below near stay empty, the shared (clipped) diagonal leaves no crack"""
    v, c, f = scenes.ground_plane()
    near, far = 0.3, 4.0
    col, dep = po.render(v, c, f, CAM, eye_pose(0.0), z_near=near, z_far=far)
    fy, oy = float(CAM[1]), float(CAM[3])
    py = np.arange(480) + 0.5
    with np.errstate(divide="ignore"):
        z_row = np.where(py > oy, 0.1 * fy / (py - oy), np.inf)          # plane depth along image row py
    visible = (z_row > near * 1.001) & (z_row < 3.0 * 0.999)
    hidden = (z_row < near * 0.999) | (z_row > 3.0 * 1.001)
    assert visible.sum() > 150 and (hidden & (py > oy)).sum() > 40
    assert np.all(dep[hidden] == 0)
    band = dep[visible]
    assert np.all(band > 0), "a crack along the clipped diagonal or a dropped triangle"
    np.testing.assert_allclose(band, np.repeat(z_row[visible][:, None], 640, 1), rtol=2e-4)
    want_c = np.round(255 * (z_row[visible] + 1) / 4)
    assert np.abs(col[visible][:, :, 0].astype(np.float64) - want_c[:, None]).max() <= 1.0
    # a triangle entirely in front of the near plane is untouched by the clipper
    v2 = v.copy(); v2[:, 2] = np.array([0.5, 0.5, 3, 3], np.float32)
    a = po.render(v2, c, f, CAM, eye_pose(0.0), z_near=near, z_far=far)
    b = po.render(v2, c, f, CAM, eye_pose(0.0), z_near=0.001, z_far=far)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    # entirely behind: nothing
    v3 = v.copy(); v3[:, 2] = -np.abs(v3[:, 2]) - 0.1
    assert not po.render(v3, c, f, CAM, eye_pose(0.0), z_near=near, z_far=far)[1].any()


def test_textured_quad_matches_bilinear_repeat_sampling(po):
    """GL_TEXTURE_2D as Model::loadModel / bindVBOs configure it (src/model.cpp:330-350, src/painter.cpp:140-160):
    GL_MODULATE with the vertex colour, GL_REPEAT, GL_LINEAR on level 0"""
    v, c, uv, f, tex = scenes.textured_quad()
    col, dep = po.render(v, c, f, CAM, eye_pose(0.0), uv=uv, tex=tex)
    fx, fy, ox, oy = [float(x) for x in CAM]
    yy, xx = np.mgrid[0:480, 0:640].astype(np.float64)
    X, Y = (xx + 0.5 - ox) / fx, (yy + 0.5 - oy) / fy               # Z = 1
    inside = (np.abs(X) < 0.69) & (np.abs(Y) < 0.49)
    assert np.abs(dep[inside] - 1.0).max() < 1e-6
    # the quad's attributes are affine in (X, Y) on the plane Z = 1 (both triangles agree: bilinear with no xy term)
    s = -0.3 + (X + 0.7) / 1.4 * 1.9
    t = -0.2 + (Y + 0.5) / 1.0 * 1.6
    th, tw = tex.shape[:2]
    fu, fv = s * tw - 0.5, t * th - 0.5
    x0, y0 = np.floor(fu), np.floor(fv)
    ax, ay = fu - x0, fv - y0
    x0i, y0i = x0.astype(int) % tw, y0.astype(int) % th
    x1i, y1i = (x0i + 1) % tw, (y0i + 1) % th
    texf = tex.astype(np.float64)
    samp = ((1 - ax) * (1 - ay))[..., None] * texf[y0i, x0i] + (ax * (1 - ay))[..., None] * texf[y0i, x1i] + \
           ((1 - ax) * ay)[..., None] * texf[y1i, x0i] + (ax * ay)[..., None] * texf[y1i, x1i]      # BGR
    # vertex colours: corners (1,1,1) (1,.5,1) (.25,1,1) (1,1,.5) - each triangle interpolates its own three corners
    def tri_colour(i0, i1, i2):
        a = np.stack([v[[i0, i1, i2], 0], v[[i0, i1, i2], 1], np.ones(3)], 1).astype(np.float64)
        w = np.linalg.solve(a.T, np.stack([X.ravel(), Y.ravel(), np.ones(X.size)]))
        return (w.T @ c[[i0, i1, i2]].astype(np.float64)).reshape(480, 640, 3), w.T.reshape(480, 640, 3)
    c1, w1 = tri_colour(0, 1, 2)
    c2, w2 = tri_colour(0, 2, 3)
    in1 = inside & (w1.min(-1) > 0.02)
    in2 = inside & (w2.min(-1) > 0.02)
    assert in1.sum() > 50000 and in2.sum() > 50000
    for m, cc in ((in1, c1), (in2, c2)):
        want_rgb = cc[m] * samp[m][:, ::-1] / 255.0
        got_rgb = col[m][:, ::-1].astype(np.float64) / 255.0
        assert np.abs(got_rgb - want_rgb).max() < 1.5 / 255
    # without texture coordinates the same mesh renders its vertex colours only
    plain = po.render(v, c, f, CAM, eye_pose(0.0))[0]
    assert np.abs(plain[in1][:, ::-1].astype(np.float64) / 255 - c1[in1]).max() < 1.5 / 255


def test_pose_label_matches_scipy_and_the_rotation(po):
    """the label of a template view (src/datasetmanager.cpp:484-491, 523-530): Eigen's Quaternionf(R) (Shoemake's branches)
    reproduces R, agrees with scipy's conversion up to sign, and the translation part is that of the inverse pose"""
    from scipy.spatial.transform import Rotation
    views = po.sphere_views(2, (0.4, 0.35, 1.1), (-45, 15, 45), skip_lower_hemi=False)
    rng = np.random.default_rng(4)
    extra = np.zeros((200, 3, 4), np.float32)
    extra[:, :, :3] = Rotation.random(200, random_state=5).as_matrix().astype(np.float32)       # all four branches
    extra[:, :, 3] = rng.standard_normal((200, 3)).astype(np.float32)
    branches = set()
    for rt in list(views[::9]) + list(extra):
        lab = po.pose_label(rt, 7)
        r = rt[:, :3].astype(np.float64)
        q = lab[1:5].astype(np.float64)
        assert lab[0] == 7 and abs(np.linalg.norm(q) - 1) < 2e-6
        assert np.abs(Rotation.from_quat(q).as_matrix() - r).max() < 3e-6
        ref = Rotation.from_matrix(r).as_quat()
        assert min(np.abs(q - ref).max(), np.abs(q + ref).max()) < 3e-6
        np.testing.assert_allclose(lab[5:], -(r.T @ rt[:, 3].astype(np.float64)), atol=3e-6)
        tr = np.trace(r)
        branches.add(3 if tr > 0 else int(np.argmax(np.diag(r))))
        if tr > 0:
            assert q[3] > 0                       # Eigen's trace branch gives w > 0
    assert branches == {0, 1, 2, 3}
