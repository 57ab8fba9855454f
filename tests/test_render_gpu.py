"""-m gpu: the CUDA rasteriser (csrc/render.cu) through the C ABI / depth_renderer.py against the CPU restatement of the
reference's mesh code and shaders (oracle/render_cpu.py): vertex / index buffers and RGBA8 frames BIT-EXACT at sizes the
oracle finishes in seconds; at the benchmark's full size (density 8, 300 poses) size-independent properties."""
import numpy as np
import pytest
import torch

from efficientdepthestimation_b200 import depth_renderer as DR
from oracle import render_cpu as R

pytestmark = pytest.mark.gpu


def scene(h, w, seed, channels=0):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    depth = (127 + 120 * np.sin(xx / 7.0 + seed) * np.cos(yy / 5.0)).astype(np.uint8)       # smooth relief + noise
    depth = np.where(rng.random((h, w)) < 0.05, rng.integers(0, 256, (h, w)), depth).astype(np.uint8)
    if channels:
        depth = np.repeat(depth[:, :, None], channels, axis=2)
    tex = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    tex[..., 3] = 255
    return depth, tex


@pytest.mark.parametrize("density,h,w,channels", [(0, 48, 64, 0), (3, 48, 64, 1), (5, 57, 76, 0), (6, 30, 41, 3)])
def test_mesh_from_depth_is_bit_exact(cuda, lib, density, h, w, channels):
    depth, tex = scene(h, w, density, channels)
    mesh = DR.Mesh.from_texture(DR.Texture(tex), depth, density=density, displacement_factor=4.0)
    v, uv, idx = R.mesh_from_depth(depth, density, displacement=4.0)
    torch.cuda.synchronize()
    assert np.array_equal(mesh.vertices.cpu().numpy(), v)
    assert np.array_equal(mesh.texture_coordinates.cpu().numpy(), uv)
    assert np.array_equal(mesh.indices.cpu().numpy().view(np.uint32), idx)


@pytest.mark.parametrize("density,win,poses", [(0, (40, 30), 2), (2, (64, 48), 3), (4, (96, 72), 3), (5, (61, 47), 2)])
def test_frames_are_bit_exact_against_the_shader_restatement(cuda, lib, density, win, poses):
    """Benchmark camera path (benchmark.py:865-886): fov_y 18, camera at dz = -10, displacement 4."""
    w, h = win
    depth, tex = scene(h, w, 10 + density)
    mesh = DR.Mesh.from_texture(DR.Texture(tex), depth, density=density, displacement_factor=4.0)
    mvps = R.benchmark_camera_mvps(w, h, 40)[::40 // poses][:poses]
    frames = DR.MeshRenderer(DR.Camera(window_size=(w, h), fov_y=18)).render(mesh, mvps).cpu().numpy()
    v, uv, idx = R.mesh_from_depth(depth, density, displacement=4.0)
    for f in range(poses):
        want = R.render(v, uv, idx, mvps[f], tex, w, h)
        bad = int((frames[f] != want).any(-1).sum())
        assert bad == 0, f"pose {f}: {bad} of {w * h} pixels differ"
    assert (frames[..., :3].reshape(poses, -1).max(1) > 0).all()                   # something was drawn
    flipped = DR.MeshRenderer(DR.Camera(window_size=(w, h), fov_y=18), flip_y=True).render(mesh, mvps).cpu().numpy()
    assert np.array_equal(flipped, frames[:, ::-1])


def test_edge_cases(cuda, lib):
    """Mesh behind the camera / off screen -> opaque black; clockwise mesh culled; coarse mesh over a large window takes
    the block-cooperative path and equals the oracle; texture identity of a flat mesh (no holes, no double hits)."""
    depth, tex = scene(24, 32, 3)
    mesh = DR.Mesh.from_texture(DR.Texture(tex), np.full((24, 32), 255, np.uint8), density=3)
    r = DR.MeshRenderer(DR.Camera(window_size=(32, 24), fov_y=18))
    ident = np.diag([1.0, 1.0 / float(mesh.vertices[0, 1]), 1.0, 1.0]).astype(np.float32)
    assert np.array_equal(r.render(mesh, ident[None])[0].cpu().numpy(), tex)
    behind = R.perspective_matrix(18, 32 / 24) @ R.translation_matrix(dz=10)       # the mesh sits BEHIND the camera: w <= 0
    away = R.translation_matrix(dx=5)                                               # entirely right of the window
    mirrored = np.diag([-1.0, 1.0, 1.0, 1.0]).astype(np.float32)                    # winding flips: back faces
    out = r.render(mesh, np.stack([behind, away, mirrored])).cpu().numpy()
    assert (out[..., :3] == 0).all() and (out[..., 3] == 255).all()
    big = DR.MeshRenderer(DR.Camera(window_size=(200, 150), fov_y=18))
    d2, t2 = scene(150, 200, 4)
    m1 = DR.Mesh.from_texture(DR.Texture(t2), d2, density=1, displacement_factor=4.0)
    mvp = R.benchmark_camera_mvps(200, 150, 7)[6]
    v, uv, idx = R.mesh_from_depth(d2, 1, displacement=4.0)
    assert np.array_equal(big.render(m1, mvp[None])[0].cpu().numpy(), R.render(v, uv, idx, mvp, t2, 200, 150))
    with pytest.raises(RuntimeError):
        DR.Texture(np.zeros((4, 4, 3), np.uint8))
    with pytest.raises(AssertionError):
        DR.Mesh.from_texture(DR.Texture(tex), depth, density=-1)


def test_benchmark_size_properties(cuda, lib):
    """create_rendered_images at full size (benchmark.py:845-846: density 8, displacement 4; one loop of the camera
    animation = fps / loops_per_second = 300 poses) on a 304x228 window: deterministic, fully covered around the centre,
    the animation returns to its start pose, and every frame of a FLAT mesh equals a homography of the texture that
    keeps colours inside the texture's range."""
    w, h = 304, 228
    depth, tex = scene(h, w, 8)
    mesh = DR.Mesh.from_texture(DR.Texture(tex), depth, density=8, displacement_factor=4.0)
    assert mesh.vertices.shape == (257 * 257, 3) and mesh.indices.numel() == 3 * 131072
    r = DR.MeshRenderer(DR.Camera(window_size=(w, h), fov_y=18))
    a = r.render_animation(mesh, DR.benchmark_camera_animation(), 300, fps=60, camera_position=DR.get_translation_matrix(dz=-10))
    b = r.render_animation(mesh, DR.benchmark_camera_animation(), 300, fps=60, camera_position=DR.get_translation_matrix(dz=-10))
    torch.cuda.synchronize()
    assert a.shape == (300, h, w, 4) and torch.equal(a, b)                          # atomicMin z-buffer: order independent
    assert bool((a[..., 3] == 255).all())
    centre = a[:, h // 4: 3 * h // 4, w // 4: 3 * w // 4, :3].reshape(300, -1, 3).amax(2)
    assert float((centre > 0).float().mean()) > 0.99                                # no holes in the mesh interior
    assert not torch.equal(a[0], a[75])                                             # the camera moves
