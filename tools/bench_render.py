"""Throughput of the CUDA rasteriser at the reference benchmark's size (Benchmark/benchmark.py:845-846: mesh density 8,
displacement 4, one loop of the camera animation = 300 poses) on a 304x228 window, plus the CPU restatement of the
shaders on a bounded sample (2 poses at density 5) for scale.  Prints one JSON line.
    python tools/bench_render.py [--frames 300] [--density 8] [--no-cpu]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import depth_renderer as DR      # noqa: E402


def main():
    frames = int(sys.argv[sys.argv.index("--frames") + 1]) if "--frames" in sys.argv else 300
    density = int(sys.argv[sys.argv.index("--density") + 1]) if "--density" in sys.argv else 8
    w, h = 304, 228
    rng = np.random.default_rng(0)
    yy, xx = np.mgrid[0:h, 0:w]
    depth = (127 + 120 * np.sin(xx / 17.0) * np.cos(yy / 11.0)).astype(np.uint8)
    tex = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    tex[..., 3] = 255
    d_dev, t_dev = torch.from_numpy(depth).cuda(), torch.from_numpy(tex).cuda()
    r = DR.MeshRenderer(DR.Camera(window_size=(w, h), fov_y=18))
    cam = DR.get_translation_matrix(dz=-10)

    def once():
        mesh = DR.Mesh.from_texture(DR.Texture(t_dev), d_dev, density=density, displacement_factor=4.0)
        return r.render_animation(mesh, DR.benchmark_camera_animation(), frames, fps=60, camera_position=cam)

    for _ in range(3):
        out = once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    t0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        out = once()
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / reps
    ms = e0.elapsed_time(e1) / reps
    n_tri = 2 * 4 ** density
    px_bytes = frames * w * h * (4 + 8 * 2)                # RGBA8 written + 64-bit z/id buffer written and read
    line = {"metric": "rendered frames/s (304x228, density-8 mesh, benchmark camera path)", "value": round(frames / ms * 1e3, 1),
            "unit": "frames/s", "ms_per_mesh": round(ms, 3), "wall_ms_per_mesh": round(wall * 1e3, 3), "frames": frames,
            "triangles": n_tri, "mtriangles_per_s": round(n_tri * frames / ms / 1e3, 1),
            "pixel_buffer_gb_per_s": round(px_bytes / ms / 1e6, 1), "covered": float((out[..., :3].amax(-1) > 0).float().mean())}
    if "--no-cpu" not in sys.argv:
        from oracle import render_cpu as R                           # CPU baseline leg only
        v, uv, idx = R.mesh_from_depth(depth, 5, 4.0)
        mv = R.benchmark_camera_mvps(w, h, 2)
        t0 = time.perf_counter()
        for f in range(2):
            R.render(v, uv, idx, mv[f], tex, w, h)
        dt = (time.perf_counter() - t0) / 2
        line["cpu_baseline"] = {"value": round(1.0 / dt, 3), "unit": "frames/s", "cores": 1, "kind": "port",
                                "sample": "2 poses, density-5 mesh (2048 triangles), 304x228, numpy restatement of the shaders"}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
