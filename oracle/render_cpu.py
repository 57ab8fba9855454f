"""CPU restatement of the reference's depth-mesh rendering path.  TEST INFRASTRUCTURE (see
``oracle/__init__.py``): only tests/ and the CPU legs of bench tools may import it.

What it restates (paths relative to the reference repository root):
  * mesh from depth            DepthRenderer/DepthRenderer/render.py:465-546 (``Mesh.from_texture``) and the
                               ``vertices[:, 2] *= displacement_factor`` of Benchmark/benchmark.py:949
  * vertex shader              DepthRenderer/DepthRenderer/shaders/shader.vert:1-14   gl_Position = mvp * vec4(p, 1)
  * fragment shader            DepthRenderer/DepthRenderer/shaders/shader.frag:1-8    texture2D(colourSampler, uv)
  * matrices / camera          DepthRenderer/DepthRenderer/utils.py:20-100 (perspective "fov_y" matrix, rotations,
                               translations), render.py:14-75 (Camera), :811 (mvp = projection @ view @ model)
  * camera animation           DepthRenderer/DepthRenderer/animation.py:30-118, composed as in
                               Benchmark/benchmark.py:875-886
  * fixed-function state       render.py:357-360 (GL_LINEAR, GL_CLAMP), :631-634 (cull back faces, depth test, clear to
                               opaque black), :787 (glReadPixels RGBA8, row 0 = bottom of the window)

PARITY UNPINNED: the reference executes this path through OpenGL (GLFW window, GLSL 4.60); a graphics stack is
unavailable on this box and the reference holds no golden frames, so nothing here can be checked against the
reference's own output.  The rasterisation rules the GL specification leaves to the implementation (sub-pixel snapping,
tie-breaking on shared edges, depth-buffer format, texel rounding) are fixed below and the CUDA rasteriser follows the
same rules bit for bit:
  window coordinates snapped to 1/256 pixel; a pixel is covered when its centre is inside the triangle under the
  top-left fill rule; depth = fp32 window z, GL_LESS, ties go to the triangle drawn first; perspective-correct uv;
  bilinear RGBA8 fetch with clamp-to-edge; colour = round-to-nearest-even of the fp32 filter result.
Every fp32 operation is a separately rounded IEEE operation in the order written (numpy float32 here, explicit
__fmul_rn / __fadd_rn / __fdiv_rn in csrc/render.cu).
"""
import math
from typing import Sequence, Tuple

import numpy as np

F = np.float32
SUB = 256                      # sub-pixel grid of the snapped window coordinates


# ------------------------------------------------------------------------------------------------ mesh
def mesh_from_depth(depth_map: np.ndarray, density: int, displacement: float = 1.0):
    """render.py:465-546: a (2^density + 1)^2 grid on the XY plane with the depth map's aspect ratio,
    z = (1 - depth / 255) * displacement sampled at the nearest pixel, uv in [0, 1], triangles (a, b, c), (c, b, d)."""
    if depth_map.ndim == 3:
        depth_map = depth_map[:, :, 0]
    height, width = depth_map.shape[:2]
    n = 2 ** density + 1
    x = np.linspace(-1, 1, n, dtype=np.float32)
    y = np.linspace(1, -1, n, dtype=np.float32)
    y = (height / width) * y - 0.5 * (1.0 - height / width) * y       # fp32: Python scalars adopt the array's dtype
    x_tex = np.linspace(0, 1, n, dtype=np.float32)
    y_tex = np.linspace(1, 0, n, dtype=np.float32)
    col_i, row_i = np.meshgrid(np.arange(n), np.arange(n))
    u = (col_i / n * width).astype(np.int64)                                         # `np.int` in the reference
    v = ((1 - row_i / n) * height - 1).astype(np.int64)
    z = 1.0 - depth_map[v, u] / 255.0
    vertices = np.stack([x[col_i], y[row_i], z], axis=-1).reshape(-1, 3).astype(np.float32)
    vertices[:, 2] *= np.float32(displacement)                                       # benchmark.py:949
    uv = np.stack([x_tex[col_i], y_tex[row_i]], axis=-1).reshape(-1, 2).astype(np.float32)
    a = (row_i[:-1, :-1] * n + col_i[:-1, :-1]).reshape(-1)
    b = ((row_i[:-1, :-1] + 1) * n + col_i[:-1, :-1]).reshape(-1)
    indices = np.stack([a, b, a + 1, a + 1, b, b + 1], axis=-1).reshape(-1).astype(np.uint32)
    return vertices, uv, indices


# ------------------------------------------------------------------------------------------------ matrices
def perspective_matrix(fov_y, aspect_ratio, near=0.01, far=1000.0):
    """utils.py:20-36 (the reference scales by fov_y itself, not by 1 / tan(fov_y / 2))."""
    return np.array([[fov_y / aspect_ratio, 0, 0, 0], [0, fov_y, 0, 0],
                     [0, 0, (far + near) / (near - far), (2 * near * far) / (near - far)], [0, 0, -1, 0]],
                    dtype=np.float32)


def rotation_matrix(angle, axis):
    """utils.py:48-81, axis in 'x', 'y', 'z'."""
    m = np.eye(4, dtype=np.float32)
    c, s = np.cos(angle), np.sin(angle)
    if axis == "x":
        m[1, 1], m[1, 2], m[2, 1], m[2, 2] = c, -s, s, c
    elif axis == "y":
        m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
    else:
        m[0, 0], m[0, 1], m[1, 0], m[1, 1] = c, -s, s, c
    return m


def translation_matrix(dx=0.0, dy=0.0, dz=0.0):
    m = np.eye(4, dtype=np.float32)
    m[0, 3], m[1, 3], m[2, 3] = dx, dy, dz
    return m


def benchmark_camera_mvps(width, height, n_frames, fps=60, fov_y=18):
    """The camera path of Benchmark/benchmark.py:865-886,1016-1018: Camera(fov_y=18) at dz = -10, animated by
    Compose([RotateAxisBounce(Y), RotateAxisBounce(X), Translate(X), Translate(Y)]); one mvp per frame, delta = 1/fps."""
    proj = perspective_matrix(fov_y, width / height)
    cam = translation_matrix(dz=-10)
    rot, lps = 2.5, 0.5 / 2.5
    out = []
    elapsed = 0.0
    for _ in range(n_frames):
        elapsed += 1.0 / fps
        ry = rotation_matrix(np.sin(2.0 * np.pi * (-lps * elapsed + 0.5)) * np.deg2rad(rot), "y")
        rx = rotation_matrix(np.sin(2.0 * np.pi * (-lps * elapsed + 0.5)) * np.deg2rad(rot / 5.0), "x")
        tx = translation_matrix(dx=np.sin(lps * elapsed * 2.0 * np.pi + 0.0) * 0.30)
        ty = translation_matrix(dy=np.sin(lps * elapsed * 2.0 * np.pi + 0.25 * 2.0 * np.pi) * 0.15)
        anim = np.eye(4, dtype=np.float32) @ ry @ rx @ tx @ ty                       # Compose.transform, animation.py:107-113
        view = cam @ anim                                                            # benchmark.py:1017
        out.append((proj @ view).astype(np.float32))                                 # render.py:75,811 (model = identity)
    return np.stack(out)


# ------------------------------------------------------------------------------------------------ vertex stage
def transform_vertices(vertices: np.ndarray, mvp: np.ndarray, width: int, height: int):
    """shader.vert:11 + the fixed-function viewport transform: per vertex (xi, yi) = window coordinates snapped to 1/256
    pixel (int32), zw = window depth, inv_w = 1 / clip w, w (all fp32).  Each product and sum rounded separately."""
    m = mvp.astype(np.float32)
    x, y, z = (vertices[:, i].astype(np.float32) for i in range(3))

    def row(r):
        return ((m[r, 0] * x + m[r, 1] * y) + m[r, 2] * z) + m[r, 3]

    cx, cy, cz, cw = row(0), row(1), row(2), row(3)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv_w = F(1.0) / cw
        ndx, ndy, ndz = cx / cw, cy / cw, cz / cw
    xw = (ndx + F(1.0)) * F(0.5 * width)
    yw = (ndy + F(1.0)) * F(0.5 * height)
    zw = ndz * F(0.5) + F(0.5)
    with np.errstate(invalid="ignore"):
        xi = np.rint(np.clip(xw * F(SUB), F(-2.0 ** 24), F(2.0 ** 24))).astype(np.int64)
        yi = np.rint(np.clip(yw * F(SUB), F(-2.0 ** 24), F(2.0 ** 24))).astype(np.int64)
    return xi, yi, zw.astype(np.float32), inv_w.astype(np.float32), cw.astype(np.float32)


def _edge_bias(ax, ay, bx, by):
    """Top-left fill rule for counter-clockwise triangles in y-up window coordinates: a pixel centre exactly on an
    edge belongs to the triangle only if the edge is a left edge (going down) or a top edge (horizontal, going left)."""
    dx, dy = bx - ax, by - ay
    return 0 if (dy < 0 or (dy == 0 and dx < 0)) else -1


def _triangle_fragments(xi, yi, tri, width, height):
    """Covered pixels of one triangle: (px, py, w0, w1, w2, area) with integer edge functions, or None when culled."""
    i0, i1, i2 = (int(t) for t in tri)
    x0, y0, x1, y1, x2, y2 = int(xi[i0]), int(yi[i0]), int(xi[i1]), int(yi[i1]), int(xi[i2]), int(yi[i2])
    area = (x1 - x0) * (y2 - y0) - (y1 - y0) * (x2 - x0)
    if area <= 0:                                             # GL_CULL_FACE (back faces) and degenerate triangles
        return None
    half = SUB // 2
    px_lo = max(0, -((-(min(x0, x1, x2) - half)) // SUB))     # ceil((min - 128) / 256)
    px_hi = min(width - 1, (max(x0, x1, x2) - half) // SUB)
    py_lo = max(0, -((-(min(y0, y1, y2) - half)) // SUB))
    py_hi = min(height - 1, (max(y0, y1, y2) - half) // SUB)
    if px_lo > px_hi or py_lo > py_hi:
        return None
    px, py = np.meshgrid(np.arange(px_lo, px_hi + 1, dtype=np.int64), np.arange(py_lo, py_hi + 1, dtype=np.int64))
    cx, cy = px * SUB + half, py * SUB + half
    w0 = (x2 - x1) * (cy - y1) - (y2 - y1) * (cx - x1)        # edge v1 -> v2, weight of v0
    w1 = (x0 - x2) * (cy - y2) - (y0 - y2) * (cx - x2)        # edge v2 -> v0, weight of v1
    w2 = (x1 - x0) * (cy - y0) - (y1 - y0) * (cx - x0)        # edge v0 -> v1, weight of v2
    inside = ((w0 + _edge_bias(x1, y1, x2, y2) >= 0) & (w1 + _edge_bias(x2, y2, x0, y0) >= 0) &
              (w2 + _edge_bias(x0, y0, x1, y1) >= 0))
    if not inside.any():
        return None
    return px[inside], py[inside], w0[inside], w1[inside], w2[inside], area


def _barycentrics(w0, w1, w2, area):
    a = F(area)                                               # int64 -> fp32, round to nearest even
    return w0.astype(np.float32) / a, w1.astype(np.float32) / a, w2.astype(np.float32) / a


def render(vertices, uv, indices, mvp, texture, width, height):
    """One frame: RGBA8 [height][width][4], row 0 = BOTTOM of the window (glReadPixels order, render.py:787).
    texture: uint8 [th][tw][4], row 0 = v 0 (glTexImage2D order, render.py:361)."""
    xi, yi, zw, inv_w, cw = transform_vertices(vertices, mvp, width, height)
    tris = indices.reshape(-1, 3).astype(np.int64)
    key = np.full((height, width), np.iinfo(np.uint64).max, dtype=np.uint64)
    # ---- pass 1: depth test (GL_LESS; equal depth -> the triangle drawn first), key = depth bits << 32 | triangle
    for t, tri in enumerate(tris):
        if not (cw[tri] > 0).all():                           # no near-plane clipping: such triangles are dropped
            continue
        fr = _triangle_fragments(xi, yi, tri, width, height)
        if fr is None:
            continue
        px, py, w0, w1, w2, area = fr
        b0, b1, b2 = _barycentrics(w0, w1, w2, area)
        z = (b0 * zw[tri[0]] + b1 * zw[tri[1]]) + b2 * zw[tri[2]]
        ok = (z >= F(0.0)) & (z <= F(1.0))                    # fragments outside the depth range are clipped
        k = (z.astype(np.float32).view(np.uint32).astype(np.uint64) << np.uint64(32)) | np.uint64(t)
        for x, y, kk, o in zip(px, py, k, ok):
            if o and kk < key[y, x]:
                key[y, x] = kk
    # ---- pass 2: shade the winning triangle of every pixel (shader.frag:7)
    frame = np.zeros((height, width, 4), dtype=np.uint8)
    frame[:, :, 3] = 255                                      # glClearColor(0, 0, 0, 1), render.py:634
    th, tw = texture.shape[:2]
    tex = texture.astype(np.float32)
    half = SUB // 2
    for y in range(height):
        for x in range(width):
            if key[y, x] == np.iinfo(np.uint64).max:
                continue
            tri = tris[int(key[y, x] & np.uint64(0xffffffff))]
            i0, i1, i2 = (int(v) for v in tri)
            x0, y0, x1, y1, x2, y2 = int(xi[i0]), int(yi[i0]), int(xi[i1]), int(yi[i1]), int(xi[i2]), int(yi[i2])
            cx, cy = x * SUB + half, y * SUB + half
            area = (x1 - x0) * (y2 - y0) - (y1 - y0) * (x2 - x0)
            w0 = (x2 - x1) * (cy - y1) - (y2 - y1) * (cx - x1)
            w1 = (x0 - x2) * (cy - y2) - (y0 - y2) * (cx - x2)
            w2 = (x1 - x0) * (cy - y0) - (y1 - y0) * (cx - x0)
            a = F(area)
            q0, q1, q2 = (F(w0) / a) * inv_w[i0], (F(w1) / a) * inv_w[i1], (F(w2) / a) * inv_w[i2]
            qs = (q0 + q1) + q2
            u = ((q0 * uv[i0, 0] + q1 * uv[i1, 0]) + q2 * uv[i2, 0]) / qs
            v = ((q0 * uv[i0, 1] + q1 * uv[i1, 1]) + q2 * uv[i2, 1]) / qs
            frame[y, x] = _texture_linear_clamp(tex, tw, th, u, v)
    return frame


def _texture_linear_clamp(tex, tw, th, u, v):
    """GL_LINEAR with the coordinates clamped to the edge texels (render.py:357-360)."""
    fx = u * F(tw) - F(0.5)
    fy = v * F(th) - F(0.5)
    x0f, y0f = np.floor(fx), np.floor(fy)
    ax, ay = F(fx - x0f), F(fy - y0f)
    x0 = int(min(max(x0f, -1.0), float(tw)))
    y0 = int(min(max(y0f, -1.0), float(th)))
    xa, xb = min(max(x0, 0), tw - 1), min(max(x0 + 1, 0), tw - 1)
    ya, yb = min(max(y0, 0), th - 1), min(max(y0 + 1, 0), th - 1)
    one = F(1.0)
    top = tex[ya, xa] * (one - ax) + tex[ya, xb] * ax
    bot = tex[yb, xa] * (one - ax) + tex[yb, xb] * ax
    c = top * (one - ay) + bot * ay
    return np.clip(np.rint(c), 0, 255).astype(np.uint8)
