"""Depth-mesh rendering on the GPU without a graphics API (host side of ede_mesh_from_depth / ede_render_mesh).

Mirrors the part of ``DepthRenderer/DepthRenderer`` that ``Benchmark/benchmark.py:845-1039`` (create_rendered_images) drives:
``Texture`` / ``Mesh.from_texture`` (render.py:333-372, 465-546), ``Camera`` (render.py:14-75), the camera animations
(animation.py:6-118) and the matrix helpers (utils.py:20-100) keep their names and argument meaning; ``MeshRenderer``
replaces the GLFW window loop (render.py:568-859) by a headless call that renders MANY camera poses of one mesh into a
device buffer of RGBA8 frames (written to files by the caller where an image is wanted).  The reference's own GL path is
unavailable on this box (no graphics stack); parity is against the CPU restatement of the shader math (oracle/render_cpu.py,
tests only).  There is no CPU fallback: without CUDA every entry point raises.
"""
import ctypes as C
import enum
from typing import Optional, Sequence

import numpy as np
import torch

from . import _cabi as A


class Axis(enum.Enum):
    X = enum.auto()
    Y = enum.auto()
    Z = enum.auto()


# ------------------------------------------------------------------------------------------------ utils.py:20-100
def get_perspective_matrix(fov_y, aspect_ratio, near=0.01, far=1000.0):
    return np.array([[fov_y / aspect_ratio, 0, 0, 0], [0, fov_y, 0, 0],
                     [0, 0, (far + near) / (near - far), (2 * near * far) / (near - far)], [0, 0, -1, 0]], dtype=np.float32)


def get_rotation_matrix(angle, axis=Axis.X, dtype=np.float32, degrees=False):
    m = np.eye(4, dtype=dtype)
    if degrees:
        angle = np.deg2rad(angle)
    i, j = {Axis.X: (1, 2), Axis.Y: (2, 0), Axis.Z: (0, 1)}[axis]
    m[i, i] = m[j, j] = np.cos(angle)
    m[i, j], m[j, i] = -np.sin(angle), np.sin(angle)
    return m


def get_translation_matrix(dx: float = 0, dy: float = 0, dz: float = 0, dtype=np.float32):
    m = np.eye(4, dtype=dtype)
    m[0, 3], m[1, 3], m[2, 3] = dx, dy, dz
    return m


# ------------------------------------------------------------------------------------------------ animation.py:6-118
class Animation:
    def __init__(self):
        self.elapsed = 0.0
        self._transform = np.eye(4, dtype=np.float32)

    def update(self, delta):
        self.elapsed += delta

    def reset(self):
        self._transform = np.eye(4, dtype=np.float32)
        self.elapsed = 0.0

    def apply(self, other):
        return other @ self.transform

    @property
    def transform(self):
        return self._transform

    @transform.setter
    def transform(self, transform):
        self._transform = transform


class RotateAxisBounce(Animation):
    def __init__(self, angle=np.pi / 2, axis=Axis.Y, speed=1.0, offset=0):
        super().__init__()
        self.angle, self.axis, self.speed, self.offset = angle, axis, speed, offset

    def update(self, delta):
        super().update(delta)
        new_angle = np.sin(2.0 * np.pi * (self.speed * self.elapsed + self.offset)) * self.angle
        self.transform = get_rotation_matrix(new_angle, axis=self.axis)


class RotateXYBounce(Animation):
    def __init__(self, angle=np.pi / 2, speed=1.0, offset=0):
        super().__init__()
        self.angle, self.speed, self.offset = angle, speed, offset

    def update(self, delta):
        super().update(delta)
        phase = 2.0 * np.pi * (self.speed * self.elapsed + self.offset)
        self.transform = (get_rotation_matrix(np.sin(phase) * self.angle, axis=Axis.Y)
                          @ get_rotation_matrix(np.cos(phase) * self.angle, axis=Axis.X))


class Translate(Animation):
    def __init__(self, distance=1.0, axis=Axis.X, speed=1.0, offset=0.0):
        super().__init__()
        self.distance, self.speed, self.axis, self.offset = distance, speed, axis, offset

    def update(self, delta):
        super().update(delta)
        t = np.sin(self.speed * self.elapsed * 2.0 * np.pi + self.offset * 2.0 * np.pi) * self.distance
        d = {Axis.X: (t, 0.0, 0.0), Axis.Y: (0.0, t, 0.0), Axis.Z: (0.0, 0.0, t)}[self.axis]
        self.transform = get_translation_matrix(*d)


class Compose(Animation):
    def __init__(self, animations):
        super().__init__()
        self.animations = animations

    def update(self, delta):
        super().update(delta)
        for animation in self.animations:
            animation.update(delta)

    def reset(self):
        for animation in self.animations:
            animation.reset()

    @property
    def transform(self):
        transform = np.eye(4, dtype=np.float32)
        for animation in self.animations:
            transform = transform @ animation.transform
        return transform

    @transform.setter
    def transform(self, transform):
        raise RuntimeError(f"{self.__class__.__name__} does not support setter for transform.")


# ------------------------------------------------------------------------------------------------ render.py
class Camera:
    """render.py:14-75: `view` is set by the caller, `projection` from get_perspective_matrix(fov_y, aspect)."""

    def __init__(self, window_size, fov_y=60, near=0.01, far=1000.0):
        self.window_size = tuple(int(v) for v in window_size)          # (width, height)
        self.fov_y, self.near, self.far = fov_y, near, far
        self.view = np.eye(4, dtype=np.float32)
        self.projection = get_perspective_matrix(fov_y, self.aspect_ratio, near=near, far=far)

    window_width = property(lambda self: self.window_size[0])
    window_height = property(lambda self: self.window_size[1])
    aspect_ratio = property(lambda self: self.window_size[0] / self.window_size[1])

    @property
    def view_projection_matrix(self):
        return self.projection @ self.view


def _cuda_u8(x, what):
    t = torch.as_tensor(x)
    if t.dtype != torch.uint8:
        raise RuntimeError(f"{what} must be uint8, got {t.dtype}")
    if not torch.cuda.is_available():
        raise RuntimeError("efficientdepthestimation_b200.depth_renderer: no CUDA device (there is no CPU fallback)")
    return t.cuda().contiguous() if t.device.type != "cuda" else t.contiguous()


class Texture:
    """RGBA texture (render.py:333-372): uint8 [height][width][4], row 0 = v 0 (the caller flips, benchmark.py:911)."""

    def __init__(self, image):
        t = _cuda_u8(image, "texture image")
        if t.dim() != 3 or t.size(2) != 4:
            raise RuntimeError(f"Texture expects an RGBA image [height][width][4], got {tuple(t.shape)}")
        self.image = t

    def copy(self):
        return Texture(self.image.clone())


class Mesh:
    """Vertex / uv / index buffers on the device (render.py:375-463); `transform` = model matrix."""

    def __init__(self, texture: Texture, vertices, texture_coordinates, indices):
        self.texture, self.vertices, self.texture_coordinates, self.indices = texture, vertices, texture_coordinates, indices
        self.transform = np.eye(4, dtype=np.float32)

    @staticmethod
    def from_texture(texture: Texture, depth_map=None, density=0, displacement_factor: float = 1.0) -> "Mesh":
        """render.py:465-546 on the GPU: (2^density + 1)^2 vertices, z = (1 - depth / 255) * displacement_factor."""
        if density % 1 != 0 or density < 0:
            raise AssertionError(f"Density must be a non-negative whole number, got {density}.")
        if depth_map is None:
            depth_map = torch.zeros(texture.image.shape[:2], dtype=torch.uint8)     # z = 1 everywhere (render.py:512)
        d = _cuda_u8(depth_map, "depth map")
        if d.dim() not in (2, 3):
            raise RuntimeError(f"depth map must be [height][width] or [height][width][channels], got {tuple(d.shape)}")
        n = 2 ** int(density) + 1
        dev = d.device
        vertices = torch.empty(n * n, 3, dtype=torch.float32, device=dev)
        uv = torch.empty(n * n, 2, dtype=torch.float32, device=dev)
        indices = torch.empty(max(1, 6 * (n - 1) * (n - 1)), dtype=torch.int32, device=dev)
        a = A.EdeMeshFromDepth(depth=d.data_ptr(), vertices=vertices.data_ptr(), uv=uv.data_ptr(), indices=indices.data_ptr(),
                               h=d.size(0), w=d.size(1), depth_stride=d.size(2) if d.dim() == 3 else 1, density=int(density),
                               displacement=float(displacement_factor))
        with torch.cuda.device(dev):
            A.check(A.lib().ede_mesh_from_depth(C.byref(a), torch.cuda.current_stream(dev).cuda_stream), "ede_mesh_from_depth")
        return Mesh(texture, vertices, uv, indices)


class MeshRenderer:
    """Headless replacement of the window loop (render.py:568-859): `render(mesh, mvps)` draws every camera pose of
    `mvps` ([frames][4][4], row-major, = camera.view_projection_matrix @ mesh.transform as in render.py:811) and returns
    the frames as a uint8 CUDA tensor [frames][height][width][4] in glReadPixels order (row 0 = bottom; flip_y=True
    gives image order)."""

    def __init__(self, camera: Camera, flip_y: bool = False):
        self.camera, self.flip_y = camera, bool(flip_y)
        self._ws = None

    @property
    def frame_buffer_shape(self):
        return self.camera.window_size

    def render(self, mesh: Mesh, mvps) -> torch.Tensor:
        dev = mesh.vertices.device
        m = torch.as_tensor(np.ascontiguousarray(np.asarray(mvps, dtype=np.float32).reshape(-1, 16))).to(dev)
        w, h = self.camera.window_size
        frames = torch.empty(m.size(0), h, w, 4, dtype=torch.uint8, device=dev)
        tex = mesh.texture.image
        a = A.EdeRender(vertices=mesh.vertices.data_ptr(), uv=mesh.texture_coordinates.data_ptr(),
                        indices=mesh.indices.data_ptr(), mvp=m.data_ptr(), texture=tex.data_ptr(), frames=frames.data_ptr(),
                        n_vertices=mesh.vertices.size(0), n_triangles=mesh.indices.numel() // 3, n_frames=m.size(0),
                        tex_h=tex.size(0), tex_w=tex.size(1), height=h, width=w, flip_y=1 if self.flip_y else 0)
        need = int(A.lib().ede_render_workspace_bytes(C.byref(a)))
        if need <= 0:
            raise RuntimeError("ede_render_workspace_bytes: bad render description")
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        a.workspace = self._ws.data_ptr()
        with torch.cuda.device(dev):
            A.check(A.lib().ede_render_mesh(C.byref(a), torch.cuda.current_stream(dev).cuda_stream), "ede_render_mesh")
        return frames

    def render_animation(self, mesh: Mesh, animation: Animation, n_frames: int, fps: float = 60.0,
                         camera_position: Optional[np.ndarray] = None) -> torch.Tensor:
        """The per-frame update of benchmark.py:1016-1018: animation.update(1 / fps); view = camera_position @
        animation.transform; one mvp per frame."""
        cam_pos = np.eye(4, dtype=np.float32) if camera_position is None else camera_position
        animation.reset()
        mvps = []
        for _ in range(int(n_frames)):
            animation.update(1.0 / fps)
            self.camera.view = cam_pos @ animation.transform
            mvps.append(self.camera.view_projection_matrix @ mesh.transform)
        return self.render(mesh, np.stack(mvps))


def benchmark_camera_animation() -> Compose:
    """The camera path of Benchmark/benchmark.py:875-886."""
    rotation_angle = 2.5
    lps = 0.5 / rotation_angle
    return Compose([RotateAxisBounce(np.deg2rad(rotation_angle), axis=Axis.Y, offset=0.5, speed=-lps),
                    RotateAxisBounce(np.deg2rad(rotation_angle / 5.0), axis=Axis.X, offset=0.5, speed=-lps),
                    Translate(distance=0.30, speed=lps),
                    Translate(distance=0.15, axis=Axis.Y, offset=0.25, speed=lps)])
