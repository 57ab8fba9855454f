"""not gpu: the CPU restatement of the depth-mesh rendering path (oracle/render_cpu.py) against the facts
the reference states about it, and the host-side mirror (depth_renderer.py) against the oracle.  PARITY UNPINNED: the
reference renders through OpenGL, unavailable on this box, and holds no golden frames."""
import numpy as np
import pytest

from efficientdepthestimation_b200 import depth_renderer as DR
from oracle import render_cpu as R


@pytest.mark.parametrize("density", [0, 1, 3, 5])
def test_mesh_counts_and_layout(density):
    """render.py:465-546: (2^d + 1)^2 vertices, 2 * 4^d triangles (a, b, c), (c, b, d), z = 1 - depth / 255."""
    depth = np.random.default_rng(density).integers(0, 256, (48, 64), dtype=np.uint8)
    v, uv, idx = R.mesh_from_depth(depth, density, displacement=4.0)
    n = 2 ** density + 1
    assert v.shape == (n * n, 3) and uv.shape == (n * n, 2) and idx.shape == (6 * (n - 1) ** 2,)
    assert v.dtype == np.float32 and idx.dtype == np.uint32
    assert v[0, 0] == -1 and v[n - 1, 0] == 1 and np.isclose(v[0, 1], 0.75 - 0.125) and np.isclose(v[-1, 1], -0.625)
    assert (uv[0] == (0, 1)).all() and (uv[-1] == (1, 0)).all()
    assert list(idx[:6]) == [0, n, 1, 1, n, n + 1]
    assert v[:, 2].min() >= 0 and v[:, 2].max() <= 4.0
    assert np.isclose(v[0, 2], (1 - depth[47, 0] / 255.0) * 4.0)            # vertex (row 0, col 0) -> pixel (height - 1, 0)
    # every triangle is counter-clockwise in the y-up plane: the reference culls back faces (render.py:631)
    t = idx.reshape(-1, 3)
    a, b, c = v[t[:, 0], :2], v[t[:, 1], :2], v[t[:, 2], :2]
    assert ((b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0]) > 0).all()


def test_host_matrices_and_animation_match_the_oracle():
    assert np.array_equal(DR.get_perspective_matrix(18, 4 / 3), R.perspective_matrix(18, 4 / 3))
    for ax, name in ((DR.Axis.X, "x"), (DR.Axis.Y, "y"), (DR.Axis.Z, "z")):
        assert np.array_equal(DR.get_rotation_matrix(0.3, ax), R.rotation_matrix(0.3, name))
    assert np.array_equal(DR.get_translation_matrix(1, 2, 3), R.translation_matrix(1, 2, 3))
    cam = DR.Camera(window_size=(64, 48), fov_y=18)
    anim = DR.benchmark_camera_animation()
    want = R.benchmark_camera_mvps(64, 48, 5)
    anim.reset()
    for i in range(5):
        anim.update(1.0 / 60)
        cam.view = DR.get_translation_matrix(dz=-10) @ anim.transform
        assert np.allclose(cam.view_projection_matrix, want[i], rtol=0, atol=0), i
    with pytest.raises(RuntimeError):
        anim.transform = np.eye(4)


def test_flat_mesh_facing_the_camera_reproduces_the_texture():
    """A flat mesh (depth 255 -> z = 0) seen through an orthographic-like mvp that maps the quad onto the whole window:
    every pixel centre samples the texture at its own texel centre, so the frame IS the texture (GL_LINEAR at texel
    centres returns the texel) -- checks fill rule (no holes / double hits on shared edges), uv interpolation, fetch."""
    h, w = 24, 32
    rng = np.random.default_rng(1)
    tex = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    v, uv, idx = R.mesh_from_depth(np.full((h, w), 255, np.uint8), 3)
    mvp = np.diag([1.0, 1.0 / v[0, 1], 1.0, 1.0]).astype(np.float32)           # stretch the quad's y extent to [-1, 1]
    frame = R.render(v, uv, idx, mvp, tex, w, h)
    assert np.array_equal(frame, tex)


def test_depth_test_and_culling():
    """Two overlapping quads at different depths: the nearer one wins everywhere it covers (GL_LESS); a clockwise
    triangle is culled; nothing drawn -> opaque black (render.py:634)."""
    tex = np.zeros((2, 2, 4), np.uint8)
    tex[..., 0], tex[..., 3] = 200, 255
    verts = np.array([[-1, -1, 0.5], [1, -1, 0.5], [-1, 1, 0.5], [1, 1, 0.5],            # far quad (z_ndc 0.5), full window
                      [-0.5, -0.5, -0.5], [0.5, -0.5, -0.5], [-0.5, 0.5, -0.5], [0.5, 0.5, -0.5]], np.float32)
    uvs = np.array([[0, 0]] * 4 + [[1, 1]] * 4, np.float32)
    tex[1, 1, 0] = 50                                                                     # the near quad samples texel (1, 1)
    idx = np.array([0, 1, 2, 2, 1, 3, 4, 5, 6, 6, 5, 7], np.uint32)
    f = R.render(verts, uvs, idx, np.eye(4, dtype=np.float32), tex, 16, 16)
    assert (f[..., 3] == 255).all()
    assert (f[4:12, 4:12, 0] == 50).all() and f[0, 0, 0] == 200 and f[15, 15, 0] == 200
    f2 = R.render(verts, uvs, idx[[0, 2, 1]], np.eye(4, dtype=np.float32), tex, 16, 16)  # clockwise: culled
    assert (f2[..., :3] == 0).all() and (f2[..., 3] == 255).all()
